#!/bin/bash
# laser kernel on the 2-D array texture: beams in flight per lane
cd $GRAFT_REPO_ROOT
V=$GRAFT_REPO_ROOT/navigation_b200/lib/variants
run() { echo "variant $1" >> gpurun_out/r2ab_variants.txt; AMCLB200_LIB=$2 timeout 300 python profiles/scripts/lf16.py >> gpurun_out/r2ab_variants.txt 2>&1; }
run default $GRAFT_REPO_ROOT/navigation_b200/lib/libamcl_b200.so
for v in d2u2 d2u4 d2u5 d2t384u6; do run $v $V/libamcl_b200_$v.so; done
