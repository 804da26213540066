#!/bin/bash
# NVTX evidence without nsys: ncu records the NVTX range stack each kernel was launched under.
A=/tmp/assets.extra; mkdir -p $A; cp tests/golden/nif_metadata.txt $A/; python -c "
import ipu_path_trace_b200 as pt; from ipu_path_trace_b200.nif import write_nif_bin; write_nif_bin('$A/nif.bin', pt.synthetic_model())"
CMD="host/b200_path_trace -o /tmp/nvtx.png --assets $A -w 1104 -h 1000 -s 24 --samples-per-step 8 --save-interval 2 --host-film"
$CMD > /dev/null 2>&1 && ncu --nvtx --metrics gpu__time_duration.sum --clock-control none -c 60 -o gpurun_out/r2_nvtx -f $CMD > gpurun_out/r2_nvtx_ncu.log 2>&1
ncu -i gpurun_out/r2_nvtx.ncu-rep --page raw --csv > gpurun_out/r2_nvtx_raw.csv 2>/dev/null
python - <<'PY'
import csv
rows = list(csv.reader(open("gpurun_out/r2_nvtx_raw.csv")))
hdr = rows[0]
print([h for h in hdr if "NVTX" in h.upper() or "Range" in h or "Domain" in h][:10])
iK = hdr.index("Kernel Name"); iT = [i for i, h in enumerate(hdr) if h.startswith("gpu__time_duration.sum")][0]
nv = [i for i, h in enumerate(hdr) if "Push/Pop" in h or "NVTX" in h.upper()]
out = open("gpurun_out/r2_nvtx_launches.csv", "w")
out.write("id,kernel,gpu_time_ns,nvtx_range_stack\n")
for r in rows[2:]:
    if len(r) != len(hdr): continue
    stack = " | ".join(r[i] for i in nv if r[i])
    out.write(f"{r[0]},{r[iK].split('(')[0]},{r[iT]},\"{stack}\"\n")
out.close()
print(open("gpurun_out/r2_nvtx_launches.csv").read()[:3000])
PY
