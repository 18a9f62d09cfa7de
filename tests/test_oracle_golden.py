"""The C restatement (oracle/amcl_oracle.c) against the committed fixture
tests/golden/amcl_golden.npz, which was produced by the compiled unmodified
reference (tests/golden/make_golden.py).  Needs neither a GPU nor the reference
tree.  All comparisons are bit-exact: the port restates the same fp64
expressions in the same order with the same libm.
"""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "amcl_golden.npz")
SCALE = 0.05
RANGE_MAX = 6.0
ALPHAS = np.array([0.2, 0.25, 0.3, 0.15, 0.1])


@pytest.fixture(scope="module")
def g():
    return dict(np.load(GOLD))


def test_exact_edt_bounds_reference_wavefront(g, port):
    occ = g["occ"]
    R = port.orc_cell_radius(1.0, SCALE)
    assert R == 20
    exact = oracle.occ_dist_from_d2(oracle.edt_exact(occ, R), SCALE, 1.0)
    refd = g["cspace_ref_dist_1p0"]
    assert np.all(exact <= refd)
    assert np.mean(exact != refd) < 0.05
    assert np.max(refd - exact) <= SCALE + 1e-12


@pytest.mark.parametrize("model,mid,beams", [("lf", 1, 24), ("lfp", 2, 48), ("beam", 0, 48)])
def test_laser_models(g, port, model, mid, beams):
    cloud = g["cloud"].copy()
    scan = g["scan"].copy()
    if model == "beam":
        scan[7, 0] = 2.0
    dist = g["dist_used_" + model] if model != "beam" else None
    lp = oracle.laser_params(laser_pose=(0.12, -0.03, 0.05))
    total = oracle.laser_model(mid, cloud, lp, beams, g["occ"], dist, SCALE, (0.0, 0.0), 1.0,
                               scan, RANGE_MAX)
    ws, wf = C.c_double(0.0), C.c_double(0.0)
    neff = port.orc_sensor_finish(cloud, cloud.shape[0], total, 0.0, 0.0, C.byref(ws), C.byref(wf))
    assert np.array_equal(cloud[:, 3], g["weights_" + model])
    assert np.array_equal(np.array([neff, ws.value, wf.value]), g["neff_" + model])


def test_lfp_beam_skip(g, port):
    cloud = g["cloud_tight"].copy()
    n = cloud.shape[0]
    tmp = np.zeros((n, 48))
    lp = oracle.laser_params()
    total = oracle.laser_model(2, cloud, lp, 48, g["occ"], g["dist_used_lfp"], SCALE, (0.0, 0.0),
                               1.0, g["scan_skip"], RANGE_MAX, do_beamskip=1, converged=1,
                               temp_obs=tmp)
    ws, wf = C.c_double(0.0), C.c_double(0.0)
    port.orc_sensor_finish(cloud, n, total, 0.0, 0.0, C.byref(ws), C.byref(wf))
    assert np.array_equal(cloud[:, 3], g["weights_lfp_skip"])


@pytest.mark.parametrize("model", [0, 1, 2, 3])
def test_motion(g, port, model):
    cloud = g["cloud"].copy()
    port.orc_srand48(40 + model)
    port.orc_reset_count()
    port.orc_motion(cloud, cloud.shape[0], model, ALPHAS, np.array([1.0, 0.5, 0.3]),
                    np.array([0.3, 0.02, 0.05]))
    assert port.orc_draw_count() == int(g["motion_draws_%d" % model][0])
    assert np.array_equal(cloud[:, :3], g["motion_%d" % model])


@pytest.mark.parametrize("tag", ["track", "global"])
def test_resample_and_clusters(g, port, tag):
    cl = g["resample_in_" + tag].copy()
    n = cl.shape[0]
    out = np.empty((n, 4))
    src = np.empty(n, np.int32)
    k = C.c_int()
    port.orc_srand48(900)
    cnt = port.orc_resample(cl, n, out, src, 20, n, 0.02, 0.99, 0.0, None, None, C.byref(k))
    want = g["resample_out_" + tag]
    assert cnt == want.shape[0]
    assert np.array_equal(out[:cnt], want)
    kk, ncl, conv = [int(v) for v in g["resample_k_" + tag]]
    assert k.value == kk
    labels = np.empty(cnt, np.int32)
    got_ncl = port.orc_cluster_labels(out[:cnt].copy(), cnt, labels, None)
    assert got_ncl == ncl
    stats = np.empty((max(ncl, 1), 18))
    st = np.empty(8)
    port.orc_cluster_stats(out[:cnt].copy(), cnt, labels, ncl, stats, st)
    assert np.array_equal(stats[:ncl, 1:10], g["clusters_" + tag])
    assert np.array_equal(stats[:ncl, 0].astype(np.int64), g["cluster_counts_" + tag])
    assert np.array_equal(st, g["setstats_" + tag])
    assert port.orc_update_converged(out[:cnt].copy(), cnt, 0.5) == conv


def test_pf_init(g, port):
    want = g["init_gaussian"]
    n = want.shape[0]
    mine = np.empty((n, 4))
    port.orc_srand48(31)
    port.orc_pf_init(mine, n, np.array([0.5, -0.4, 0.7]), np.diag([0.2, 0.1, 0.05]).copy(), 0)
    assert np.array_equal(mine, want)
