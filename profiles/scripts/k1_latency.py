"""K1-sized update (5000 particles x 180 beams, 1000^2 map): wall time per API call and launches."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
from navigation_b200 import abi, library, synth
from navigation_b200.harness import Side
from profiles.run_configs import gpu_scan
lib = library.load()
lib.amclb200_set_rng(library.RNG_PHILOX, 7)
for cloud_kind in ("global", "tracking"):
    occ = synth.rooms_map(1000, seed=1)
    side = Side(lib, occ)
    n = 5000
    side.alloc(100, n)
    mirror = int(os.environ.get("MIRROR", "1"))
    lib.amclb200_set_mirror(side.pf, mirror)
    side.make_laser(180, model="lf")
    side.make_odom(model=abi.ODOM_MODEL_DIFF)
    pose = synth.room_centre_pose(1000, 0.05)
    scan = gpu_scan(side, pose, 180, 30.0)
    cloud = synth.uniform_cloud(occ, 0.05, n, 3) if cloud_kind == "global" else synth.gaussian_cloud(pose, (0.3, 0.3, 0.15), n, 3)
    pristine = torch.from_numpy(cloud).cuda()
    acc = np.zeros(4)
    reps = 50
    for it in range(reps + 5):
        lib.amclb200_pf_load_device(side.pf, pristine.data_ptr(), n)
        torch.cuda.synchronize()
        l0 = lib.amclb200_launch_count()
        t0 = time.perf_counter(); side.update_action((0.1, 0, 0.02), (0.1, 0, 0.02))
        t1 = time.perf_counter(); side.update_sensor(scan, 30.0)
        t2 = time.perf_counter(); side.resample()
        t3 = time.perf_counter()
        if it >= 5:
            acc += (t1 - t0, t2 - t1, t3 - t2, lib.amclb200_launch_count() - l0)
    acc /= reps
    print("%s mirror=%d: action %.1f us, sensor %.1f us, resample %.1f us, total %.1f us, %.0f launches, n_after %d clusters %d"
          % (cloud_kind, mirror, acc[0] * 1e6, acc[1] * 1e6, acc[2] * 1e6, acc[:3].sum() * 1e6, acc[3], side.cur.sample_count, side.cur.cluster_count))
    side.close()
