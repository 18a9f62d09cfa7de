#!/bin/bash
# laser kernel: the view as a 2-D array texture (texel fetch with integer coordinates, hardware clamp)
cd $GRAFT_REPO_ROOT
echo "default (linear texture)" >> gpurun_out/r2aa_tex2d.txt
timeout 300 python profiles/scripts/lf16.py >> gpurun_out/r2aa_tex2d.txt 2>&1
echo "2-D array texture" >> gpurun_out/r2aa_tex2d.txt
AMCLB200_LF_TEX2D=1 AMCLB200_LF_TEXQ=100 timeout 300 python profiles/scripts/lf16.py >> gpurun_out/r2aa_tex2d.txt 2>&1
AMCLB200_LF_TEX2D=1 AMCLB200_LF_TEXQ=100 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_edge_cases.py -q -k "sensor or laser or likelihood or lf or skip or beam or edge or off" > gpurun_out/r2aa_pytest.log 2>&1
echo "rc $?" >> gpurun_out/r2aa_pytest.log
