#!/bin/bash
# distance-transform kernels after the O(1)-per-cell rewrite: parity, timings, ncu
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_parity.py -q -k "edt" > gpurun_out/r2q_pytest.log 2>&1
echo "pytest edt rc $?" >> gpurun_out/r2q_pytest.log
timeout 900 python -m pytest tests/test_gpu_inflation.py tests/test_gpu_node_io.py tests/test_gpu_edge_cases.py tests/test_gpu_native.py tests/test_zz_no_alloc.py -q >> gpurun_out/r2q_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/r2q_pytest.log
timeout 600 python profiles/run_configs.py maps > gpurun_out/r2q_maps.txt 2>&1
AMCLB200_EDT=search timeout 600 python profiles/scripts/edt_once.py > gpurun_out/r2q_search.txt 2>&1
timeout 600 python profiles/scripts/edt_once.py > gpurun_out/r2q_fast.txt 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"edt_|infl_" -c 12 -o gpurun_out/r2q_edt python profiles/scripts/edt_once.py > gpurun_out/r2q_ncu.log 2>&1
ncu -i gpurun_out/r2q_edt.ncu-rep --page raw --csv > gpurun_out/r2q_edt_raw.csv 2>/dev/null
