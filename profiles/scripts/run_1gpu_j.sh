#!/bin/bash
# laser kernel: fewer warps with more fetches in flight per lane
cd $GRAFT_REPO_ROOT
V=$GRAFT_REPO_ROOT/navigation_b200/lib/variants
run() { echo "variant $1 TEXQ $2" >> gpurun_out/r2v_variants.txt; AMCLB200_LIB=$V/libamcl_b200_$1.so AMCLB200_LF_TEXQ=$2 timeout 300 python profiles/scripts/lf16.py >> gpurun_out/r2v_variants.txt 2>&1; }
run t448u3 8
run t448u4 8
run t384u4 8
run t384u6 8
