"""One K4-style update (likelihood_field_prob + resample) for launch-list profiling."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from profiles.run_configs import sensor_case
import json
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 22
print(json.dumps(sensor_case("K4", 4000, n, 1080, sys.argv[2] if len(sys.argv) > 2 else "lfp", resample=True)))
import ctypes as C
from navigation_b200 import library
t = (C.c_int * 2)(); library.load().amclb200_cumsum_last_tiles(t); print("dirty tiles", t[0], "of", t[1])
