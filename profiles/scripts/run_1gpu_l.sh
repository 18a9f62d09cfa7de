#!/bin/bash
# laser kernel: fetches for TEXQ of the 8 particles of a lane, loads for the rest, 3 (or 4) beams in flight
cd $GRAFT_REPO_ROOT
V=$GRAFT_REPO_ROOT/navigation_b200/lib/variants
run() { echo "variant $1 TEXQ $2" >> gpurun_out/r2y_variants.txt; AMCLB200_LIB=$V/libamcl_b200_$1.so AMCLB200_LF_TEXQ=$2 timeout 300 python profiles/scripts/lf16.py >> gpurun_out/r2y_variants.txt 2>&1; }
run split 4
run split 5
run split 6
run split 7
run split_u4 6
run split_u4 5
