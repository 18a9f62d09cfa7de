#!/bin/bash
cd $GRAFT_REPO_ROOT
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"cc_brick|cs_carries" -c 6 -o gpurun_out/r2_cc -f python bench.py --steps 1 --warmup 0 --no-cpu-baseline --parity-particles 0 > gpurun_out/r2k_ncu.log 2>&1
ncu -i gpurun_out/r2_cc.ncu-rep --page raw --csv > gpurun_out/r2_cc_raw.csv 2>/dev/null
ncu -i gpurun_out/r2_cc.ncu-rep --page source --csv > gpurun_out/r2_cc_src.csv 2>/dev/null
bash profiles/scripts/run_sanitizer.sh > gpurun_out/r2k_sanitizer_tail.log 2>&1
python profiles/run_configs.py > gpurun_out/r2k_configs.log 2>&1
MIRROR=1 python profiles/scripts/k1_latency.py > gpurun_out/r2k_k1.txt 2>&1
MIRROR=0 python profiles/scripts/k1_latency.py >> gpurun_out/r2k_k1.txt 2>&1
