"""ncu raw page (csv on stdin) of ONE nif_tc_kernel launch -> profiles-style traffic json with the hashes of the kernel sources
it was captured from (bench.py refuses a capture whose hashes differ from the tree's). Usage:
  ncu -i rep.ncu-rep --page raw --csv | python scratch/make_traffic_json.py "<capture command>" > gpurun_out/r3_nif_tc_traffic.json"""
import csv
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench

rows = list(csv.reader(sys.stdin))
hdr, vals = rows[0], rows[-1]
get = lambda k: next(v for h, v in zip(hdr, vals) if h == k)
num = lambda k: float(get(k).replace(",", ""))
units = dict(zip(hdr, rows[1]))
scale = lambda k: {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(units[k], 1)
rd, wr = num("dram__bytes_read.sum") * scale("dram__bytes_read.sum"), num("dram__bytes_write.sum") * scale("dram__bytes_write.sum")
dur = num("gpu__time_duration.sum") * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(units["gpu__time_duration.sum"], 1e-6)
out = {"kernel": get("Kernel Name"), "capture": sys.argv[1] if len(sys.argv) > 1 else "", "dram_bytes_read_per_launch": rd,
       "dram_bytes_write_per_launch": wr, "gpu_time_duration_ms": dur, "source_sha256": bench.nif_source_hashes()}
for k, name in (("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm_pipe_tensor_cycles_active_pct_of_elapsed"),):
    try:
        out[name] = num(k)
    except StopIteration:
        pass
print(json.dumps(out, indent=1))
