#!/bin/bash
# Round 3, visit A (2 GPUs): full GPU suite incl. the multi-GPU tests, bench at N=2 with config.multi_gpu_check, ncu capture of the NIF kernel
# -> traffic json with source hashes, bench at N=1.
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu > gpurun_out/r3_gputests_2gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r3_gputests_2gpu.log
NG=$(nvidia-smi -L | wc -l)
if [ "$NG" -ge 2 ]; then
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 6 --warmup 3 --cli-spp 9000 > gpurun_out/r3_bench_n2.json 2> gpurun_out/r3_bench_n2.err; echo "bench n2 rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r3_bench_n2.json").read().strip().splitlines()[-1])
print("n2 value", d["value"], "e2e", d["e2e"]["value"]); print("multi_gpu_check", d["config"].get("multi_gpu_check")); print("cli_c2", {k: v for k, v in d["config"].get("cli_c2", {}).items() if k != "clock"})
PY
fi
[ -n "$ONLY_MULTI" ] && exit 0
SHORT="python bench.py --steps 1 --warmup 1 --samples-per-step 128 --no-cpu-baseline --no-extras"
timeout 600 $SHORT > gpurun_out/r3_bench_short.json 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:nif_tc -s 12 -c 1 -o gpurun_out/r3_nif_tc_bench -f $SHORT > gpurun_out/r3_ncu_full.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/r3_nif_tc_bench.ncu-rep --page raw --csv 2>/dev/null | python scratch/make_traffic_json.py "ncu --set full --clock-control none --import-source on -k regex:nif_tc -s 12 -c 1 $SHORT" > gpurun_out/r3_nif_tc_traffic.json; cat gpurun_out/r3_nif_tc_traffic.json
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r3_bench_n1.json 2> gpurun_out/r3_bench_n1.err; echo "bench n1 rc=$?"; cut -c1-300 gpurun_out/r3_bench_n1.json
