#!/bin/bash
cd $GRAFT_REPO_ROOT
python profiles/scripts/lf16.py > gpurun_out/r2l_lf_main.txt 2>&1
AMCLB200_LIB=$PWD/scratch/libamcl_block8.so python profiles/scripts/lf16.py > gpurun_out/r2l_lf_block8.txt 2>&1
AMCLB200_LIB=$PWD/scratch/libamcl_block8.so timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edge_cases.py -x -q > gpurun_out/r2l_pytest_block8.log 2>&1
echo "rc $?" >> gpurun_out/r2l_pytest_block8.log
AMCLB200_LIB=$PWD/scratch/libamcl_block8.so timeout 300 ncu --set full --clock-control none -k regex:likelihood_field -s 1 -c 1 -o gpurun_out/r2_lf_block8 -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --parity-particles 0 > gpurun_out/r2l_ncu.log 2>&1
ncu -i gpurun_out/r2_lf_block8.ncu-rep --page raw --csv > gpurun_out/r2_lf_block8_raw.csv 2>/dev/null
