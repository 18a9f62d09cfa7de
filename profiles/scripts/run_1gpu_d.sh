#!/bin/bash
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -x -q -k "resample or cumulative or full_update or recovery" > gpurun_out/r2p_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/r2p_pytest.log
AMCLB200_TRACE=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --parity-particles 0 > gpurun_out/r2p_bench.json 2> gpurun_out/r2p_bench.err
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"pick_and_copy|bucket_heads|maxscan|sample_table|cs_carries|cc_brick" -c 24 --csv --log-file gpurun_out/r2p_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --parity-particles 0 > /dev/null 2>&1
