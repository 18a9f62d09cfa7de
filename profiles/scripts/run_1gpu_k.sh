#!/bin/bash
# laser kernel, texture fetches + unrolled beam loop: the shipped build and neighbours
cd $GRAFT_REPO_ROOT
V=$GRAFT_REPO_ROOT/navigation_b200/lib/variants
run() { echo "variant $1" >> gpurun_out/r2w_variants.txt; AMCLB200_LIB=$2 timeout 300 python profiles/scripts/lf16.py >> gpurun_out/r2w_variants.txt 2>&1; }
run default $GRAFT_REPO_ROOT/navigation_b200/lib/libamcl_b200.so
for v in c256u3 t384u8 t384u5 tp3 t320u8; do run $v $V/libamcl_b200_$v.so; done
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edge_cases.py -q -x -k "sensor or laser or likelihood or lf or skip or beam" > gpurun_out/r2w_pytest.log 2>&1
echo "rc $?" >> gpurun_out/r2w_pytest.log
