#!/bin/bash
# final single-GPU pass of round 2: whole GPU suite, the bench line, ncu of the laser kernel, launch list, K1 latency
cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r2ac_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/r2ac_pytest.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2ac_bench1.json 2> gpurun_out/r2ac_bench1.err
echo "bench rc $?" >> gpurun_out/r2ac_bench1.err
timeout 300 python profiles/scripts/k1_latency.py > gpurun_out/r2ac_k1.txt 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:likelihood_field -s 1 -c 1 -o gpurun_out/r2ac_lf -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --parity-particles 0 > gpurun_out/r2ac_ncu.log 2>&1
ncu -i gpurun_out/r2ac_lf.ncu-rep --page raw --csv > gpurun_out/r2ac_lf_raw.csv 2>/dev/null
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2ac_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --parity-particles 0 > gpurun_out/r2ac_launches.log 2>&1
