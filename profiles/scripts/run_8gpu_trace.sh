#!/bin/bash
cd $GRAFT_REPO_ROOT
AMCLB200_TRACE=1 NCCL_DEBUG=WARN timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29641 bench.py --gpus 8 --steps 6 --warmup 3 --no-cpu-baseline --parity-particles 0 > gpurun_out/r2n_bench8_trace.json 2> gpurun_out/r2n_bench8_trace.err
