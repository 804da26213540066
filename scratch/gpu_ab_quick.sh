#!/bin/bash
A=$PWD/ipu_path_trace_b200/libb200pt_prev.so; B=$PWD/ipu_path_trace_b200/libb200pt.so
timeout 300 python -m pytest tests/test_gpu_nif.py -m gpu -x -q 2>&1 | tail -2
for i in 1 2; do for lib in $A $B; do echo -n "$(basename $lib) "; B200PT_LIB=$lib timeout 120 python scratch/nif_bench.py 4194304 20; done; done 2>&1
