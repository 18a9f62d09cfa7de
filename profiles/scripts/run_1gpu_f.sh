#!/bin/bash
# laser kernel: TEXQ of the 8 gathers per lane through the texture pipe (AMCLB200_LF_TEXQ)
cd $GRAFT_REPO_ROOT
for q in 0 2 3 4 8; do
  echo "TEXQ $q" >> gpurun_out/r2r_texq.txt
  AMCLB200_LF_TEXQ=$q timeout 300 python profiles/scripts/lf16.py >> gpurun_out/r2r_texq.txt 2>&1
done
AMCLB200_LF_TEXQ=4 timeout 600 python -m pytest tests/test_gpu_parity.py -q -x -k "sensor or laser or likelihood or lf" > gpurun_out/r2r_pytest.log 2>&1
echo "rc $?" >> gpurun_out/r2r_pytest.log
