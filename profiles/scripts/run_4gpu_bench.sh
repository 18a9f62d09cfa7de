#!/bin/bash
# bench line at N = 4 with the final laser kernel
cd $GRAFT_REPO_ROOT
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29614 bench.py --gpus 4 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2ae_bench4.json 2> gpurun_out/r2ae_bench4.err
echo "bench rc $?" >> gpurun_out/r2ae_bench4.err
