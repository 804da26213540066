#!/bin/bash
# Round-2 evidence on one B200: smoke, both bench arms, ncu launch list and --set full captures of the two hot kernels.
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2 | tee gpurun_out/r2_smoke.log
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err; echo "ref exit $?"
timeout 1200 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo "bench exit $?"; cut -c1-600 gpurun_out/r2_bench_n1.json
SHORT="python bench.py --steps 1 --warmup 1 --samples-per-step 128 --no-cpu-baseline --no-extras"
timeout 600 $SHORT > gpurun_out/r2_bench_short.json 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv $SHORT > gpurun_out/r2_ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:nif_tc -s 12 -c 1 -o gpurun_out/r2_nif_tc_bench -f $SHORT > gpurun_out/r2_ncu_full.log 2>&1; echo "ncu full exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:trace_wavefront -s 12 -c 1 -o gpurun_out/r2_trace_wavefront_bench -f $SHORT > gpurun_out/r2_ncu_full_trace.log 2>&1; echo "ncu trace exit $?"
for rep in gpurun_out/r2_nif_tc_bench gpurun_out/r2_trace_wavefront_bench; do
  ncu -i $rep.ncu-rep --page raw --csv 2>/dev/null | python -c "
import sys,csv
rows=list(csv.reader(sys.stdin)); hdr=rows[0]; vals=rows[-1]
want=['Kernel Name','gpu__time_duration.sum','sm__pipe_tensor_cycles_active','sm__inst_executed_pipe_tensor','smsp__thread_inst_executed_per_inst_executed.ratio','smsp__inst_executed.sum','smsp__issue_active.avg.pct','dram__bytes_read.sum','dram__bytes_write.sum','launch__registers_per_thread','launch__grid_size','launch__block_size','launch__cluster','lts__t_sectors_srcunit_tex_op_read.sum','sm__cycles_elapsed.max','smsp__cycles_active.avg','launch__shared_mem_per_block_dynamic','sm__throughput.avg.pct','gpc__cycles_elapsed.avg.per_second','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct']
for h,v in zip(hdr,vals):
    if any(h.startswith(w) for w in want): print(h, v)
" > $rep.summary.txt; done
python scratch/parity_probe.py 2>&1 | tail -3 | tee gpurun_out/r2_parity_probe.log
