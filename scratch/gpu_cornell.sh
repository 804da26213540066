#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_cornell.py -q -m gpu 2>&1 | tail -15 | tee gpurun_out/cornell_test.log
CORNELL="host/b200_trace -o /tmp/cornell.png -w 1024 -h 1024 -s 256"
timeout 300 $CORNELL > gpurun_out/cornell_cli_1024.log 2>&1 && cat gpurun_out/cornell_cli_1024.log && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:cornell_pixels -c 1 -o gpurun_out/cornell_pixels_v7 -f $CORNELL > gpurun_out/ncu_cornell.log 2>&1; echo "ncu cornell exit $?"
host/b200_trace -o /tmp/c.png -w 1920 -h 1080 -s 1024 | tee gpurun_out/cornell_cli.log
