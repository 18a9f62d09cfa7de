#!/bin/bash
# laser kernel through the texture pipe: more variants, the slice size of an 8-GPU run, repeatability
cd $GRAFT_REPO_ROOT
for q in 6 7 8 0 8 0; do
  echo "TEXQ $q" >> gpurun_out/r2s_texq.txt
  AMCLB200_LF_TEXQ=$q timeout 300 python profiles/scripts/lf16.py >> gpurun_out/r2s_texq.txt 2>&1
done
for q in 0 8; do
  echo "TEXQ $q 2Mi" >> gpurun_out/r2s_texq.txt
  LF16_N=2097152 AMCLB200_LF_TEXQ=$q timeout 300 python profiles/scripts/lf16.py >> gpurun_out/r2s_texq.txt 2>&1
done
