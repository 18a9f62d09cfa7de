#!/bin/bash
# 8 GPUs: K5 at N = 8 (traced and untraced) and N = 4
cd $GRAFT_REPO_ROOT
nvidia-smi -L | wc -l > gpurun_out/r2i_gpus.txt
AMCLB200_TRACE=1 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2i_bench8_trace.json 2> gpurun_out/r2i_bench8_trace.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29622 bench.py --gpus 8 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2i_bench8.json 2> gpurun_out/r2i_bench8.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29623 bench.py --gpus 4 --steps 20 --warmup 3 --no-cpu-baseline --parity-particles 0 > gpurun_out/r2i_bench4.json 2> gpurun_out/r2i_bench4.err
SHARD_N=100003 SHARD_P2P=1 SHARD_MODEL=lf_recover SHARD_NCCL=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29624 tests/sharded_worker.py > gpurun_out/r2i_worker8.log 2>&1
