#!/bin/bash
# laser kernel: code-generation variants (profiles/scripts/build_lf_variants.sh), loads (TEXQ 0) and fetches (TEXQ 8)
cd $GRAFT_REPO_ROOT
for v in default cg noalloc twophase unroll1 unroll2; do
  lib=$GRAFT_REPO_ROOT/navigation_b200/lib/variants/libamcl_b200_$v.so
  [ $v = default ] && lib=$GRAFT_REPO_ROOT/navigation_b200/lib/libamcl_b200.so
  for q in 0 8; do
    echo "variant $v TEXQ $q" >> gpurun_out/r2t_variants.txt
    AMCLB200_LIB=$lib AMCLB200_LF_TEXQ=$q timeout 300 python profiles/scripts/lf16.py >> gpurun_out/r2t_variants.txt 2>&1
  done
done
