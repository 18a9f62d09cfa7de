"""CPU: the flat-file replay (navigation_b200.replay, runFromBag's role) driving
the COMPILED REFERENCE through the same per-scan loop the CUDA library is driven
with -- the host logic (thresholds, scan conversion, resample interval,
arg-max cluster) is exercised without a GPU."""
import math

import numpy as np

from navigation_b200 import driver, replay, synth


def test_log_round_trip_and_reference_replay(ref, tmp_path):
    occ = synth.rooms_map(300, seed=3)
    path_pts = [(2.5, 2.5, 0.0)] + [(2.5 + 0.25 * k, 2.5, 0.0) for k in range(1, 16)]
    bag, bear = driver.simulate_run(ref, occ, 0.05, path_pts, n_beams=40, seed=2)
    f = str(tmp_path / "log.npz")
    replay.save_log(f, bag, bear, 12.0, occ=occ, scale=0.05)
    log = replay.load_log(f)
    assert np.array_equal(log["odom"], np.array([b[0] for b in bag]))
    assert np.array_equal(log["ranges"], np.array([b[1] for b in bag]))
    assert np.array_equal(log["occ"], occ) and log["scan"][3] == 12.0
    ref.tap_set_ignore_srand(1)
    ref.tap_seed(11)
    track, updates = replay.replay(ref, log, min_particles=200, max_particles=1500, laser_max_beams=20,
                                   init_pose=(2.55, 2.45, 0.02), init_cov=(0.05, 0.05, 0.02),
                                   d_thresh=0.2, a_thresh=0.2, resample_interval=1)
    ref.tap_set_ignore_srand(0)
    assert updates >= 12 and track[-1]["n"] >= 200
    err, aerr = replay.score(track)
    odom_err = math.hypot(log["odom"][-1][0] - log["truth"][-1][0], log["odom"][-1][1] - log["truth"][-1][1])
    assert err < 0.25 and aerr < 0.15 and err < odom_err + 0.25
