"""Generates tests/golden/amcl_golden.npz from the compiled, UNMODIFIED reference
(oracle/_ref/libamcl_ref.so, built by oracle/Makefile from /root/reference).

Run in the build container (where /root/reference is mounted):
    python tests/golden/make_golden.py
The fixture is committed; tests/test_oracle_golden.py checks the C restatement
against it and tests/test_gpu_parity.py checks the CUDA path against it, so both
gates survive on machines without the reference tree.
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from navigation_b200 import synth  # noqa: E402
from oracle import oracle  # noqa: E402
from tests.scenario import Side, make_scan  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "amcl_golden.npz")

MAP_SIZE = 160
SCALE = 0.05
RANGE_MAX = 6.0
ALPHAS = (0.2, 0.25, 0.3, 0.15, 0.1)


def main():
    ref = oracle.ref()
    assert ref is not None, "reference not built: run `make -C oracle ref`"
    g = {}
    occ = synth.rooms_map(MAP_SIZE, seed=7, room=50, door=(20, 30), clutter_mod=700,
                          unknown_ring=True)
    g["occ"] = occ
    side = Side(ref, occ, scale=SCALE)
    side.update_cspace(1.0)
    g["cspace_ref_dist_1p0"] = side.occ_dist()
    pose = synth.room_centre_pose(MAP_SIZE, SCALE, room=50)
    g["true_pose"] = pose
    B = 48
    scan = make_scan(side, pose, B, range_max=RANGE_MAX)
    scan[7, 0] = np.nan
    g["scan"] = scan
    n = 256
    cloud = synth.gaussian_cloud(pose, (0.3, 0.3, 0.2), n, 5)
    w = np.random.default_rng(2).uniform(0.2, 1.8, n)
    cloud[:, 3] = w / w.sum()
    g["cloud"] = cloud

    # --- laser models (a6, a7, a8): weights after pf_update_sensor -------------
    for model, beams in (("lf", 24), ("lfp", 48), ("beam", 48)):
        side.alloc(10, n)
        side.make_laser(beams, model=model, laser_pose=(0.12, -0.03, 0.05), max_occ_dist=1.0)
        if model != "beam":
            g["dist_used_" + model] = side.occ_dist()
        sc = scan.copy()
        if model == "beam":
            sc[7, 0] = 2.0
        side.set_samples(cloud)
        side.update_sensor(sc, RANGE_MAX)
        g["weights_" + model] = side.samples()[:, 3]
        g["neff_" + model] = np.array([side.cur.n_effective, side.pf.contents.w_slow,
                                       side.pf.contents.w_fast])
        ref.amcl_c_laser_delete(side.laser)
        side.laser = None
        ref.pf_free(side.pf)
        side.pf = None

    # --- beam skip (a7 second pass) --------------------------------------------
    tight = synth.gaussian_cloud(pose, (0.04, 0.04, 0.02), n, 6)
    g["cloud_tight"] = tight
    sc = scan.copy()
    sc[~(sc[:, 0] < RANGE_MAX), 0] = 1.5
    sc[10:14, 0] *= 0.5  # a "person" in front of the wall: these beams disagree
    g["scan_skip"] = sc
    side.alloc(10, n)
    side.make_laser(B, model="lfp", beamskip=(1, 0.5, 0.3, 0.9), max_occ_dist=1.0)
    side.set_samples(tight)
    side.pf.contents.sets[side.pf.contents.current_set].converged = 1
    side.update_sensor(sc, RANGE_MAX)
    g["weights_lfp_skip"] = side.samples()[:, 3]
    ref.amcl_c_laser_delete(side.laser)
    side.laser = None
    ref.pf_free(side.pf)
    side.pf = None

    # --- motion models (a3) -------------------------------------------------------
    for model in range(4):
        side.alloc(10, n)
        side.make_odom(model=model, alphas=ALPHAS)
        side.set_samples(cloud)
        side.seed(40 + model)
        ref.tap_log_begin(100000)
        side.update_action((1.0, 0.5, 0.3), (0.3, 0.02, 0.05))
        m = ref.tap_log_size()
        log = np.empty(m)
        ref.tap_log_copy(log.ctypes.data, m)
        ref.tap_log_end()
        g["motion_%d" % model] = side.samples()[:, :3]
        g["motion_draws_%d" % model] = np.array([m])
        ref.amcl_c_odom_delete(side.odom)
        side.odom = None
        ref.pf_free(side.pf)
        side.pf = None

    # --- resample + clusters (a10-a12) ---------------------------------------------
    for tag, cl in (("track", cloud), ("global", synth.uniform_cloud(occ, SCALE, n, 3))):
        cl = cl.copy()
        w = np.random.default_rng(9).gamma(0.6, size=n)
        cl[:, 3] = w / w.sum()
        g["resample_in_" + tag] = cl
        side.alloc(20, n, pop_err=0.02, pop_z=0.99)
        side.set_samples(cl)
        side.seed(900)
        side.resample()
        out = side.samples()
        g["resample_out_" + tag] = out
        g["resample_k_" + tag] = np.array([side.cur.kdtree.contents.leaf_count,
                                           side.cur.cluster_count,
                                           side.pf.contents.converged])
        tab = side.cluster_table()
        g["clusters_" + tag] = np.array([[wt, *mean, cov[0, 0], cov[0, 1], cov[1, 0], cov[1, 1],
                                          cov[2, 2]] for wt, mean, cov in tab])
        g["cluster_counts_" + tag] = np.array([side.cur.clusters[c].count for c in range(len(tab))])
        cv = np.array(side.cur.cov.tolist())
        g["setstats_" + tag] = np.array([*side.cur.mean.tolist(), cv[0, 0], cv[0, 1], cv[1, 0],
                                         cv[1, 1], cv[2, 2]])
        ref.pf_free(side.pf)
        side.pf = None

    # --- pf_init (a2) ---------------------------------------------------------------
    side.alloc(10, n)
    ref.tap_set_ignore_srand(1)
    side.seed(31)
    side.init_gaussian((0.5, -0.4, 0.7), (0.2, 0.1, 0.05))
    ref.tap_set_ignore_srand(0)
    g["init_gaussian"] = side.samples()
    ref.pf_free(side.pf)
    side.pf = None
    side.close()

    np.savez_compressed(OUT, **g)
    print("wrote", OUT, os.path.getsize(OUT), "bytes,", len(g), "arrays")


if __name__ == "__main__":
    main()
