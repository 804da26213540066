#!/bin/bash
for tv in v1 v2; do
  B200PT_TRACE=$tv B200PT_OVERLAP=0 ncu --set full --clock-control none --import-source on -k regex:trace_wavefront -s 4 -c 1 -o gpurun_out/r2_trace_$tv -f python bench.py --steps 1 --warmup 1 --samples-per-step 16 --no-cpu-baseline > gpurun_out/r2_trace_ncu_$tv.log 2>&1
  ncu -i gpurun_out/r2_trace_$tv.ncu-rep --page raw --csv 2>/dev/null | python -c "
import sys,csv
rows=list(csv.reader(sys.stdin))
hdr=rows[0]; vals=rows[2] if len(rows)>2 else rows[1]
want=['gpu__time_duration.sum','smsp__thread_inst_executed_per_inst_executed.ratio','smsp__inst_executed.sum','sm__inst_executed_pipe_fp64','smsp__issue_active.avg.pct_of_peak_sustained_active','sm__warps_active.avg.pct_of_peak_sustained_active','dram__bytes_read.sum','dram__bytes_write.sum','launch__registers_per_thread','smsp__warp_issue_stalled_barrier_per_warp_active.pct','smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio','smsp__average_warp_latency_issue_stalled_barrier','lts__t_bytes.sum','l1tex__t_bytes.sum','smsp__cycles_active.avg','sm__cycles_elapsed.max']
for h,v in zip(hdr,vals):
    if any(h.startswith(w) for w in want): print(h, v)
"
done
