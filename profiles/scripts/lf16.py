"""Laser-kernel time at 16 Mi particles x 1080 beams (LF and LFP); AMCLB200_LIB selects a variant build."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from profiles.run_configs import sensor_case
N = int(os.environ.get("LF16_N", 1 << 24))
for model in os.environ.get("LF16_MODELS", "lf,lfp").split(","):
    r = sensor_case("x", 4000, N, 1080, model, resample=False)
    print(model, N, r["kernel_ms"], r["sensor_ms"])
