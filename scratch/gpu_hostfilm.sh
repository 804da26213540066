#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_host_cli.py -m gpu -x -q 2>&1 | tail -3
A=/tmp/assets.extra; mkdir -p $A; cp tests/golden/nif_metadata.txt $A/; python -c "
import ipu_path_trace_b200 as pt; from ipu_path_trace_b200.nif import write_nif_bin; write_nif_bin('$A/nif.bin', pt.synthetic_model())"
for mode in "" "--host-film" "--enable-load-balancing"; do
  host/b200_path_trace -o /tmp/hf.png --assets $A -w 1104 -h 1000 -s 6000 --samples-per-step 300 --save-interval 10 $mode 2> /tmp/hf.log
  echo "mode [$mode]: $(grep -E '\] Samples/sec:' /tmp/hf.log | cut -c30-) ; median step $(awk '/Completed render step/ {print $9}' /tmp/hf.log | sort -g | sed -n 10p)"
done
