#!/bin/bash
# Round-end evidence: full GPU parity suite, smoke, bench (both arms), ncu launch list and one --set full capture.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2 | tee gpurun_out/smoke.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"; cat gpurun_out/bench_n1.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref exit $?"; cat gpurun_out/bench_ref.json
SHORT="python bench.py --steps 1 --warmup 1 --samples-per-step 128 --no-cpu-baseline"
timeout 600 $SHORT > gpurun_out/bench_short.json 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $SHORT > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:nif_tc -s 12 -c 1 -o gpurun_out/nif_tc_mc10_bench -f $SHORT > gpurun_out/ncu_full.log 2>&1; echo "ncu full exit $?"
timeout 900 ncu --set full --clock-control none -k regex:trace_wavefront -s 12 -c 1 -o gpurun_out/trace_wavefront_bench -f $SHORT > gpurun_out/ncu_full_trace.log 2>&1; echo "ncu trace exit $?"
# second scene (Cornell box): CLI timing and one ncu capture of the pixel-stream kernel
CORNELL="host/b200_trace -o /tmp/cornell.png -w 1024 -h 1024 -s 256"
timeout 300 $CORNELL > gpurun_out/cornell_cli_1024.log 2>&1 && cat gpurun_out/cornell_cli_1024.log && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:cornell_pixels -c 1 -o gpurun_out/cornell_pixels -f $CORNELL > gpurun_out/ncu_cornell.log 2>&1; echo "ncu cornell exit $?"
python scratch/parity_probe.py 2>&1 | tail -3 | tee gpurun_out/parity_probe.log
