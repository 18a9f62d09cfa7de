"""Two full updates (motion + likelihood_field + resample with clusters) at sizes that reach
both resampling paths, for compute-sanitizer (memcheck / racecheck)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from navigation_b200 import abi, library, synth
from navigation_b200.harness import Side, make_scan
lib = library.load()
lib.amclb200_set_rng(library.RNG_PHILOX, 11)
occ = synth.rooms_map(400, seed=3)
pose = synth.room_centre_pose(400, 0.05)
for n, kind in ((3000, "tracking"), (20000, "global"), (20000, "tracking")):
    side = Side(lib, occ)
    side.alloc(100, n)
    side.make_laser(60, model="lf")
    side.make_odom(model=abi.ODOM_MODEL_DIFF)
    scan = make_scan(side, pose, 60, range_max=12.0)
    cloud = synth.uniform_cloud(occ, 0.05, n, 3) if kind == "global" else synth.gaussian_cloud(pose, (0.3, 0.3, 0.15), n, 3)
    side.set_samples(cloud)
    for step in range(2):
        side.update_action((0.1, 0.0, 0.02), (0.1, 0.0, 0.02))
        side.update_sensor(scan, 12.0)
        side.resample()
    print(n, kind, "n_after", side.cur.sample_count, "k", side.cur.kdtree.contents.leaf_count, "clusters", side.cur.cluster_count)
    side.close()
print("SANITIZE-CASE-DONE")
