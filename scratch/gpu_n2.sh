#!/bin/bash
# 2-GPU refresh with the final code: full GPU suite (uses GPU 0), bench at N=2 under torchrun, reference arm under torchrun.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu 2>&1 | tail -4 | tee gpurun_out/pytest_gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_g2.json 2> gpurun_out/bench_g2.err; echo "N=2 exit $?"; tail -c 1500 gpurun_out/bench_g2.json
