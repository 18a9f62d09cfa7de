#!/bin/bash
# bench line at N = 2 with the final laser kernel
cd $GRAFT_REPO_ROOT
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2ad_bench2.json 2> gpurun_out/r2ad_bench2.err
echo "bench rc $?" >> gpurun_out/r2ad_bench2.err
