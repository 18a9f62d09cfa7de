#!/bin/bash
cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2g_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/r2g_pytest.log
MIRROR=1 python profiles/scripts/k1_latency.py > gpurun_out/r2g_k1.txt 2>&1
MIRROR=0 python profiles/scripts/k1_latency.py >> gpurun_out/r2g_k1.txt 2>&1
AMCLB200_TRACE=1 MIRROR=0 python profiles/scripts/k1_latency.py 2>&1 | grep trace | tail -3 >> gpurun_out/r2g_k1.txt
AMCLB200_TRACE=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err
