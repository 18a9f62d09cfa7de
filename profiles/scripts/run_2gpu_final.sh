#!/bin/bash
# final two-GPU pass of round 2: sharded tests (incl. the packed cloud of a sharded filter) and the bench line at N = 2
cd $GRAFT_REPO_ROOT
timeout 1200 python -m pytest tests/test_gpu_sharded.py -x -q > gpurun_out/r2z_pytest_sharded.log 2>&1
echo "pytest rc $?" >> gpurun_out/r2z_pytest_sharded.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2z_bench2.json 2> gpurun_out/r2z_bench2.err
echo "bench rc $?" >> gpurun_out/r2z_bench2.err
