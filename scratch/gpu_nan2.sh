#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_kat.py tests/test_gpu_trace.py tests/test_gpu_render.py -q -m gpu 2>&1 | tail -5
bash scratch/gpu_nan.sh 2>&1 | grep -E "non-finite|Samples/sec:"
python scratch/parity_probe.py 2>&1 | tail -2
