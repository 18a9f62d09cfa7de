#!/bin/bash
cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2h_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/r2h_pytest.log
MIRROR=1 python profiles/scripts/k1_latency.py > gpurun_out/r2h_k1.txt 2>&1
MIRROR=0 python profiles/scripts/k1_latency.py >> gpurun_out/r2h_k1.txt 2>&1
AMCLB200_TRACE=1 MIRROR=0 python profiles/scripts/k1_latency.py 2>&1 | grep trace | tail -3 >> gpurun_out/r2h_k1.txt
# ncu --set full of the stage kernels of one bench step (K5) and of one K1 update
timeout 900 ncu --set full --clock-control none --import-source on \
  -k regex:"pick_and_copy|cc_brick|cluster_accumulate|motion_kernel|edt_col_pass|edt_row_bits|cs_carries|cs_apply|cs_tile|bin_insert|find_kld_cut|cc_flatten|cc_component_min|normalise_weights" \
  -c 30 -o gpurun_out/r2_stages -f python bench.py --steps 1 --warmup 0 --no-cpu-baseline --parity-particles 0 > gpurun_out/r2h_ncu.log 2>&1
ncu -i gpurun_out/r2_stages.ncu-rep --page raw --csv > gpurun_out/r2_stages_raw.csv 2>/dev/null
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"cs_small" -s 3 -c 1 -o gpurun_out/r2_cs_small -f env MIRROR=0 python profiles/scripts/k1_latency.py > gpurun_out/r2h_ncu2.log 2>&1
ncu -i gpurun_out/r2_cs_small.ncu-rep --page source --csv > gpurun_out/r2_cs_small_src.csv 2>/dev/null
ncu -i gpurun_out/r2_stages.ncu-rep --page source --csv -k regex:cs_carries > gpurun_out/r2_cs_carries_src.csv 2>/dev/null
