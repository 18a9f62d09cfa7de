"""Laser-kernel time at 16 Mi particles x 1080 beams (LF and LFP); AMCLB200_LIB selects a variant build."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from profiles.run_configs import sensor_case
for model in ("lf", "lfp"):
    r = sensor_case("x", 4000, 1 << 24, 1080, model, resample=False)
    print(model, r["kernel_ms"], r["sensor_ms"])
