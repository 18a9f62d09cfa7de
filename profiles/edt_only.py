"""map_update_cspace alone on a 4000x4000 map (for launch-list profiling of the EDT kernels)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from navigation_b200 import library, synth
from navigation_b200.harness import Side
lib = library.load()
occ = synth.rooms_map(int(sys.argv[1]) if len(sys.argv) > 1 else 4000, seed=1)
side = Side(lib, occ)
for _ in range(3):
    side.update_cspace(2.0)
print("edt ms", lib.amclb200_map_last_edt_ms(side.map))
