#!/bin/bash
# A/B: alternate the two builds a few times on the same box
for i in 1 2 3; do
  for lib in libb200pt.so libb200pt_b.so; do
    B200PT_LIB=$PWD/ipu_path_trace_b200/$lib python scratch/nif_bench.py ${1:-4194304} 20 | sed "s/^/$lib /"
  done
done
