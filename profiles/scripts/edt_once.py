"""One map_update_cspace (4000^2, radius 40 cells) and one inflation pass (radius 11
cells) for an ncu capture of the distance-transform kernels."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from navigation_b200 import library, synth          # noqa: E402
from navigation_b200.harness import Side            # noqa: E402

lib = library.load()
size = 4000
occ = synth.rooms_map(size, seed=1)
side = Side(lib, occ)
lib.amclb200_map_set_mirror(side.map, 0)
side.update_cspace(2.0)
side.update_cspace(2.0)
print("edt kernels ms", lib.amclb200_map_last_edt_ms(side.map))
g = torch.from_numpy(np.where(occ == 1, 254, 0).astype(np.uint8)).cuda()
lib.amclb200_inflate_costs(g.data_ptr(), 1, size, size, 0, 0, size, size, 0.05, 0.2, 0.55, 10.0, 0)
print("inflation kernels ms", lib.amclb200_inflate_last_ms())
side.close()
