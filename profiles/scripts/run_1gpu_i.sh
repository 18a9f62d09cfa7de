#!/bin/bash
# laser kernel: deeper unrolling of the beam loop with texture fetches; 4 particles per warp with 24 warps
cd $GRAFT_REPO_ROOT
V=$GRAFT_REPO_ROOT/navigation_b200/lib/variants
run() { echo "variant $1 TEXQ $2" >> gpurun_out/r2u_variants.txt; AMCLB200_LIB=$V/libamcl_b200_$1.so AMCLB200_LF_TEXQ=$2 timeout 300 python profiles/scripts/lf16.py >> gpurun_out/r2u_variants.txt 2>&1; }
run unroll3 8
run unroll4 8
run p4u4 4
run p4u2 4
run p4u4 0
