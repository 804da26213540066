#!/bin/bash
# One GPU visit: parity tests, bench, variant A/B, in-kernel timers, ncu launch list + full capture.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
for mode in ts10 ts5 tp10 tp5 ss1; do
  for n in 1048576 4194304; do
    echo -n "$mode $n " ; B200PT_NIF_MODE=$mode timeout 120 python scratch/nif_bench.py $n 20
  done
done > gpurun_out/ab.log 2>&1
cat gpurun_out/ab.log
timeout 120 python scratch/nif_prof.py 4194304 > gpurun_out/prof_ts10.log 2>&1; cat gpurun_out/prof_ts10.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"; cat gpurun_out/bench_n1.json
timeout 600 python bench.py --steps 1 --warmup 1 --samples-per-step 20 --no-cpu-baseline > gpurun_out/bench_short.json 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 1 --samples-per-step 20 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:nif_tc -s 3 -c 1 -o gpurun_out/nif_tc_ts10 -f python scratch/nif_bench.py 1048576 5 > gpurun_out/ncu_full.log 2>&1; echo "ncu full exit $?"
