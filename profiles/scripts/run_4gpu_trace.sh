#!/bin/bash
cd $GRAFT_REPO_ROOT
AMCLB200_TRACE=1 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29631 bench.py --gpus 4 --steps 10 --warmup 3 --no-cpu-baseline --parity-particles 0 > gpurun_out/r2m_bench4_trace.json 2> gpurun_out/r2m_bench4_trace.err
