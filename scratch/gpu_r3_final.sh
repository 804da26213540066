#!/bin/bash
# Round-3 evidence on one B200: smoke, both bench arms, ncu launch list, --set full capture of the NIF kernel -> traffic json (with source hashes).
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2 | tee gpurun_out/r3_smoke.log
SHORT="python bench.py --steps 1 --warmup 1 --samples-per-step 128 --no-cpu-baseline --no-extras"
timeout 600 $SHORT > gpurun_out/r3_bench_short.json 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r3_bench_launches.csv $SHORT > gpurun_out/r3_ncu_launch.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:nif_tc -s 12 -c 1 -o gpurun_out/r3_nif_tc_bench -f $SHORT > gpurun_out/r3_ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i gpurun_out/r3_nif_tc_bench.ncu-rep --page raw --csv 2>/dev/null | python scratch/make_traffic_json.py "ncu --set full --clock-control none --import-source on -k regex:nif_tc -s 12 -c 1 $SHORT" > gpurun_out/r3_nif_tc_traffic.json; cp gpurun_out/r3_nif_tc_traffic.json profiles/r3_nif_tc_traffic.json; grep -c . gpurun_out/r3_nif_tc_traffic.json
ncu -i gpurun_out/r3_nif_tc_bench.ncu-rep --page raw --csv 2>/dev/null | python -c "
import sys,csv
rows=list(csv.reader(sys.stdin)); hdr=rows[0]; vals=rows[-1]
want=['Kernel Name','gpu__time_duration.sum','sm__pipe_tensor_cycles_active','sm__inst_executed_pipe_tensor','smsp__issue_active.avg.pct','dram__bytes_read.sum','dram__bytes_write.sum','launch__registers_per_thread','launch__grid_size','launch__block_size','sm__cycles_elapsed.max','sm__throughput.avg.pct','lts__t_sectors_op_red.sum','lts__t_sectors_op_atom.sum','l1tex__t_set_accesses_pipe_lsu_mem_global_op_red.sum']
for h,v in zip(hdr,vals):
    if any(h.startswith(w) for w in want): print(h, v)
" > gpurun_out/r3_nif_tc_ncu_metrics.txt
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r3_bench_reference_arm.json 2> gpurun_out/r3_bench_ref.err; echo "ref rc=$?"; cut -c1-250 gpurun_out/r3_bench_reference_arm.json
timeout 1200 python bench.py --steps 10 --warmup 3 > gpurun_out/r3_bench_n1.json 2> gpurun_out/r3_bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r3_bench_n1.json").read().strip().splitlines()[-1])
print("value %.4g e2e %.4g ms %.2f" % (d["value"], d["e2e"]["value"], d["ms_per_step"])); print(d["roofline"]); print(d["clocks"]); print(d.get("cpu_baseline"))
PY
