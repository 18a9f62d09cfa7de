"""GPU parity at the sizes BASELINE.json names (K2, K4) against the CPU oracle
port, plus the small gaps round 1 left (pf_get_cep_stats and pf_init on the
device).  Same gates as tests/test_gpu_parity.py: integers bit-exact, floating
point within 1e-5 relative.  All calls go through the C ABI.
"""
import ctypes as C

import numpy as np
import pytest

from navigation_b200 import abi, library, synth
from oracle import oracle
from tests.scenario import Side
from tests.test_gpu_parity import RTOL, _check_resample_against_oracle, rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture()
def lib(gpu_lib):
    gpu_lib.amclb200_set_rng(library.RNG_DRAND48, 0)
    return gpu_lib


def gpu_scan(lib, side, pose, beams, range_max):
    bear = synth.scan_bearings(beams)
    rays = np.stack([np.full_like(bear, pose[0]), np.full_like(bear, pose[1]), pose[2] + bear], 1).copy()
    out = np.empty(beams)
    lib.amclb200_map_calc_ranges(side.map, rays.ctypes.data, beams, range_max, out.ctypes.data)
    return np.stack([np.minimum(out, range_max), bear], 1).copy()


# ------------------------------------------------------------ K2: beam model ----

def test_k2_full_size_beam_model_vs_oracle(lib, port):
    """BASELINE config 2 at its own size: 100 000 particles x 360 beams on a
    2000 x 2000 map, beam model (amcl_laser.cpp:144-213 + map_range.c:38-120),
    every weight against the oracle port (36 M rays on one host core)."""
    size, n, beams, rmax = 2000, 100000, 360, 30.0
    occ = synth.rooms_map(size, seed=1)
    side = Side(lib, occ)
    pose = synth.room_centre_pose(size, 0.05)
    scan = gpu_scan(lib, side, pose, beams, rmax)
    cloud = synth.uniform_cloud(occ, 0.05, n, 3)
    cloud[: n // 2] = synth.gaussian_cloud(pose, (0.5, 0.5, 0.26), n // 2, 4)
    cloud[:, 3] = 1.0 / n
    side.alloc(100, n)
    side.make_laser(beams, model="beam")
    side.set_samples(cloud)
    side.update_sensor(scan, rmax)
    got = side.samples()[:, 3]
    mine = cloud.copy()
    total = oracle.laser_model(0, mine, oracle.laser_params(), beams, occ, None, 0.05, (0.0, 0.0), 2.0,
                               scan, rmax)
    ws, wf = C.c_double(0.0), C.c_double(0.0)
    neff = port.orc_sensor_finish(mine, n, total, 0.0, 0.0, C.byref(ws), C.byref(wf))
    assert rel_err(got, mine[:, 3]) < RTOL
    assert abs(side.cur.n_effective - neff) / neff < RTOL
    # the ray caster itself, at this map size: ranges bit-exact (a ray whose end
    # cell flips by an ulp of the device sin/cos may differ; none is expected)
    rng = np.random.default_rng(0)
    m = 20000
    half = size * 0.05 / 2
    rays = np.stack([rng.uniform(-half, half, m), rng.uniform(-half, half, m), rng.uniform(-4, 4, m)], 1).copy()
    out = np.empty(m)
    lib.amclb200_map_calc_ranges(side.map, rays.ctypes.data, m, rmax, out.ctypes.data)
    want = np.array([port.orc_calc_range(occ, size, size, 0.0, 0.0, 0.05, x, y, a, rmax) for x, y, a in rays])
    assert np.mean(out == want) > 0.9995
    side.close()


# ------------------------------------------- K4: resample + clusters at scale ----

@pytest.mark.parametrize("model,n,seed", [("lf", 1 << 20, 11), ("lf", 1 << 22, 12), ("lfp", 1 << 20, 13)])
def test_k4_resample_at_scale_vs_oracle(lib, port, model, n, seed):
    """BASELINE config 4: 1080-beam sensor update on the 4000^2 map, then
    pf_update_resample (pf.c:363-510) and pf_cluster_stats (pf.c:559-693).  The
    GPU's own post-sensor set goes to the oracle's O(N log N) resampler with the
    same drand48 stream: n, leaf_count and every source index bit-exact, the
    cluster partition by bijection, statistics within 1e-5.  The likelihood_field
    weights of a global cloud keep n in the millions, k in the 10^5..10^6 range
    and tens of thousands of clusters (dense histogram at scale, the > 4096
    cluster ranking path); likelihood_field_prob collapses the set onto a few bins."""
    size, beams, rmax = 4000, 1080, 30.0
    occ = synth.rooms_map(size, seed=1)
    side = Side(lib, occ)
    lib.amclb200_map_set_mirror(side.map, 0)
    side.alloc(100, n)
    side.make_laser(beams, model=model)
    pose = synth.room_centre_pose(size, 0.05)
    scan = gpu_scan(lib, side, pose, beams, rmax)
    cloud = synth.uniform_cloud(occ, 0.05, n, seed)
    cloud[:, 3] = 1.0 / n
    side.set_samples(cloud)
    side.update_sensor(scan, rmax)
    after = side.samples()
    assert np.array_equal(after[:, :3], cloud[:, :3])
    assert abs(after[:, 3].sum() - 1.0) < 1e-9
    side.close()
    cnt, k, ncl = _check_resample_against_oracle(lib, port, occ, after, 100, 0.01, 0.99, 70 + seed)
    if model == "lf":
        assert cnt == n and k > 100000
        assert ncl > 4096 or n > (1 << 20)      # 4 Mi samples fill the bins: fewer, larger clusters
    else:
        assert cnt < n


# ------------------------------------------------- a12: pf_get_cep_stats ----

@pytest.mark.parametrize("n", [300, 200000])
def test_cep_stats_vs_oracle(lib, port, n):
    """pf_get_cep_stats (pf.c:701-740) on the device-resident set."""
    occ = synth.rooms_map(80, seed=1, room=40, door=(10, 20))
    side = Side(lib, occ)
    cloud = synth.gaussian_cloud((1.0, 2.0, 0.0), (0.4, 0.2, 0.1), n, 8)
    w = np.random.default_rng(3).gamma(0.8, size=n)
    cloud[:, 3] = w / w.sum()
    side.alloc(10, n)
    side.set_samples(cloud)
    mean = abi.pf_vector_t()
    var = C.c_double()
    lib.pf_get_cep_stats(side.pf, C.byref(mean), C.byref(var))
    out = np.empty(3)
    port.orc_cep_stats(cloud, n, out)
    assert abs(mean.v[0] - out[0]) <= RTOL * abs(out[0])
    assert abs(mean.v[1] - out[1]) <= RTOL * abs(out[1])
    assert mean.v[2] == 0.0
    assert abs(var.value - out[2]) <= RTOL * abs(out[2])
    side.close()


# ---------------------------------------------------------- a2: pf_init ----

def test_pf_init_matches_reference_bit_exact(lib, ref):
    """pf_init (pf.c:136-174): with the library's own re-seeding switched off
    and the reference's srand48 calls ignored, both sides draw the Gaussian cloud
    from the same drand48 stream -- the node-default covariance has a repeated
    eigenvalue, so this also pins the eigenbasis (eig3.c).  Poses bit-exact."""
    occ = synth.rooms_map(200, seed=6, room=50, door=(20, 30))
    out = []
    for L in (ref, lib):
        side = Side(L, occ)
        side.alloc(50, 5000)
        if L is ref:
            L.tap_set_ignore_srand(1)
        else:
            L.amclb200_set_reseed(0)
        side.seed(31)
        side.init_gaussian((0.7, -0.4, 0.3), (0.25, 0.25, 0.0685))
        out.append((side.samples(), side.cur.cluster_count, side.cur.kdtree.contents.leaf_count,
                    np.array(side.cur.mean.tolist()), np.array(side.cur.cov.tolist())))
        if L is ref:
            L.tap_set_ignore_srand(0)
        else:
            L.amclb200_set_reseed(1)
        side.close()
    (sa, ca, ka, ma, va), (sb, cb, kb, mb, vb) = out
    assert np.array_equal(sa, sb)
    assert ca == cb and ka == kb
    assert np.allclose(ma, mb, rtol=1e-9, atol=1e-12)
    assert np.allclose(va, vb, rtol=1e-7, atol=1e-12)


def test_push_distances_reports_values_without_a_code(lib):
    """amclb200_map_push_distances: a distance that is not sqrt(di^2 + dj^2) * scale
    is the caller's error -- -1 and the device map as it was, not an abort."""
    occ = synth.rooms_map(120, seed=1, room=40, door=(10, 20))
    side = Side(lib, occ)
    side.update_cspace(1.0)
    good = side.occ_dist()
    bad = good.copy()
    bad[17, 23] = 0.123                          # not a cell distance
    lib.amcl_c_map_set_occ_dist(side.map, np.ascontiguousarray(bad).ctypes.data, 1.0)
    assert lib.amclb200_map_push_distances(side.map) == -1
    lib.amcl_c_map_set_occ_dist(side.map, np.ascontiguousarray(np.sqrt(7.0) * 0.05 + 0 * good).ctypes.data, 1.0)
    assert lib.amclb200_map_push_distances(side.map) == -1      # 7 is not a sum of two squares
    lib.amclb200_map_pull_distances(side.map)                    # the device map still holds the transform
    assert np.array_equal(side.occ_dist(), good)
    side.close()
