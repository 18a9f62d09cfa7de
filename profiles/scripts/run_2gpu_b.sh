#!/bin/bash
cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests/test_gpu_sharded.py -x -q > gpurun_out/r2j_pytest_sharded.log 2>&1
echo "pytest rc $?" >> gpurun_out/r2j_pytest_sharded.log
AMCLB200_TRACE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2j_bench2.json 2> gpurun_out/r2j_bench2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 20 --warmup 3 --no-cpu-baseline --parity-particles 0 > gpurun_out/r2j_bench2_notrace.json 2> gpurun_out/r2j_bench2_notrace.err
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_small_filters.py -x -q > gpurun_out/r2j_pytest_1gpu.log 2>&1
echo "pytest rc $?" >> gpurun_out/r2j_pytest_1gpu.log
AMCLB200_TRACE=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2j_bench1.json 2> gpurun_out/r2j_bench1.err
