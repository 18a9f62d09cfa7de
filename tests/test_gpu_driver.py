"""End-to-end: the ROS-free per-scan loop (navigation_b200.driver, restating
AmclNode::laserReceived) replays a synthetic "bag" -- drifting odometry plus
noisy scans ray-cast from the ground truth -- and must track the robot the way
the reference's basic_localization tests demand (final pose within a
tolerance), on the CUDA library and, with the same drand48 seed, on the
compiled reference; the two pose estimates must agree."""
import ctypes as C
import math

import numpy as np
import pytest

from navigation_b200 import driver, library, synth

pytestmark = pytest.mark.gpu


def _path():
    pts = []
    x, y, th = 2.5, 2.5, 0.0
    for _ in range(22):                      # through the door into the next room
        x += 0.25
        pts.append((x, y, th))
    for _ in range(6):                       # turn left on the spot
        th += math.pi / 12
        pts.append((x, y, th))
    for _ in range(5):
        y += 0.25
        pts.append((x, y, th))
    return [(2.5, 2.5, 0.0)] + pts


def _run(lib, bag, bear, occ, seed, is_ref, dist=None):
    # neither side may re-seed the shared drand48 stream behind the test's back
    if is_ref:
        lib.tap_set_ignore_srand(1)
        lib.tap_seed(seed)
    else:
        libc = C.CDLL(None)
        libc.srand48.argtypes = [C.c_long]
        lib.amclb200_set_reseed(0)
        libc.srand48(seed)
    d = driver.AmclDriver(lib, occ, min_particles=500, max_particles=3000, laser_max_beams=30,
                          alpha_slow=0.0, alpha_fast=0.0, init_pose=(2.6, 2.4, 0.05),
                          init_cov=(0.1, 0.1, 0.03), d_thresh=0.2, a_thresh=0.2, resample_interval=1)
    sy, sx = occ.shape
    if is_ref:
        dist = np.empty((sy, sx))
        lib.amcl_c_map_get_occ_dist(d.map, dist.ctypes.data)
    elif dist is not None:
        # like for like: evaluate on the reference's (wavefront) distances, SURVEY C1
        lib.amcl_c_map_set_occ_dist(d.map, dist.ctypes.data, 2.0)
        lib.amclb200_map_push_distances(d.map)
    est = []
    inc = bear[1] - bear[0]
    for odom, ranges, truth in bag:
        out = d.laser_received(odom, ranges, bear[0], inc, 0.0, 12.0)
        if out["pose"] is not None:
            est.append((np.array(out["pose"]), truth, out["n"]))
    updates = d.updates
    d.close()
    if is_ref:
        lib.tap_set_ignore_srand(0)
    else:
        lib.amclb200_set_reseed(1)
    return est, updates, dist


def test_tracking_run_matches_truth_and_reference(gpu_lib, ref):
    gpu_lib.amclb200_set_rng(library.RNG_DRAND48, 0)
    occ = synth.rooms_map(400, seed=3)
    bag, bear = driver.simulate_run(gpu_lib, occ, 0.05, _path(), n_beams=60, seed=1)
    est_ref, up_ref, dist = _run(ref, bag, bear, occ, 77, True)
    est_gpu, up_gpu, _ = _run(gpu_lib, bag, bear, occ, 77, False, dist)
    assert up_gpu == up_ref and up_gpu >= 25
    pose, truth, n = est_gpu[-1]
    err = math.hypot(pose[0] - truth[0], pose[1] - truth[1])
    aerr = abs(driver.angle_diff(pose[2], truth[2]))
    assert err < 0.25 and aerr < 0.15, (err, aerr)       # basic_localization-style gate
    # the odometry alone has drifted further than that
    odom_err = math.hypot(bag[-1][0][0] - truth[0], bag[-1][0][1] - truth[1])
    assert odom_err > err
    # same seed, same scans: the reference lands in the same place
    pr, tr, nr = est_ref[-1]
    assert math.hypot(pr[0] - truth[0], pr[1] - truth[1]) < 0.25
    assert len(est_gpu) == len(est_ref)
    for (pg, _, ng), (pr, _, nr) in zip(est_gpu, est_ref):
        assert ng == nr                                   # KLD sample counts step by step
        assert np.allclose(pg, pr, rtol=0, atol=1e-6)


def test_flat_file_replay_equals_in_memory_run(gpu_lib, tmp_path):
    """runFromBag's role (amcl_node.cpp:676-766): a run written to a flat file and
    replayed from it gives the same estimates as feeding the scans directly."""
    from navigation_b200 import replay

    gpu_lib.amclb200_set_rng(library.RNG_DRAND48, 0)
    occ = synth.rooms_map(400, seed=3)
    bag, bear = driver.simulate_run(gpu_lib, occ, 0.05, _path(), n_beams=60, seed=1)
    path = str(tmp_path / "run.npz")
    replay.save_log(path, bag, bear, 12.0, occ=occ, scale=0.05)
    log = replay.load_log(path)
    assert log["ranges"].shape == (len(bag), 60)
    params = dict(min_particles=500, max_particles=3000, laser_max_beams=30, alpha_slow=0.0, alpha_fast=0.0,
                  init_pose=(2.6, 2.4, 0.05), init_cov=(0.1, 0.1, 0.03), d_thresh=0.2, a_thresh=0.2,
                  resample_interval=1)
    libc = C.CDLL(None)
    libc.srand48.argtypes = [C.c_long]
    gpu_lib.amclb200_set_reseed(0)
    libc.srand48(5)
    track, updates = replay.replay(gpu_lib, log, **params)
    libc.srand48(5)
    est, up2, _ = _run_plain(gpu_lib, bag, bear, occ, params)
    gpu_lib.amclb200_set_reseed(1)
    assert updates == up2 and len(track) == len(est)
    assert np.allclose(np.array([t["pose"] for t in track]), np.array(est), rtol=0, atol=1e-12)
    err, aerr = replay.score(track)
    assert err < 0.25 and aerr < 0.15


def _run_plain(lib, bag, bear, occ, params):
    d = driver.AmclDriver(lib, occ, **params)
    inc = bear[1] - bear[0]
    est = []
    for odom, ranges, truth in bag:
        out = d.laser_received(odom, ranges, bear[0], inc, 0.0, 12.0)
        if out["pose"] is not None:
            est.append(out["pose"])
    up = d.updates
    d.close()
    return est, up, None
