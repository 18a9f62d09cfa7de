"""GPU vs the compiled reference on the awkward inputs the reference's code
paths special-case: NaN and max-range readings, scans with nothing to
integrate, ragged beam counts (step > 1), particles off the map, a
one-particle filter, min_samples == max_samples.  Both sides get the same
calls through the same binder; the GPU side evaluates on the reference's own
(wavefront) distances so that only the laser model is compared."""
import numpy as np
import pytest

from navigation_b200 import abi, library, synth
from tests.scenario import Side, make_scan

pytestmark = pytest.mark.gpu
RTOL = 1e-9


def both(gpu_lib, ref, occ):
    gpu_lib.amclb200_set_rng(library.RNG_DRAND48, 0)
    return Side(ref, occ), Side(gpu_lib, occ)


def lasers(r, g, max_beams, model, **kw):
    """Same laser on both sides.  SetModelLikelihoodField* runs map_update_cspace
    (amcl_laser.cpp:96-98, :113-115), so the reference's wavefront distances are
    handed to the GPU side AFTER the models are set."""
    r.make_laser(max_beams, model=model, **kw)
    g.make_laser(max_beams, model=model, **kw)
    if model != "beam":
        g.set_occ_dist(r.occ_dist(), 2.0)


def weights_after(side, cloud, scan, rmax, n_min=10):
    side.alloc(n_min, cloud.shape[0])
    side.set_samples(cloud)
    ok = side.update_sensor(scan, rmax)
    return ok, side.samples()[:, 3].copy(), side.cur.n_effective


def assert_close(a, b, floor=1e-290):
    scale = np.maximum(np.abs(b), floor)
    assert np.max(np.abs(a - b) / scale) < RTOL


@pytest.mark.parametrize("model", ["lf", "lfp", "beam"])
def test_nan_and_max_range_readings(gpu_lib, ref, model):
    occ = synth.rooms_map(300, seed=2)
    r, g = both(gpu_lib, ref, occ)
    pose = synth.room_centre_pose(300, 0.05)
    scan = make_scan(r, pose, 61, range_max=8.0)
    scan[3, 0] = np.nan if model != "beam" else 7.0      # the beam model's mixture propagates NaN on both sides
    scan[10:14, 0] = 8.0                                 # at range_max: skipped by lf / lfp, scored by beam
    scan[20, 0] = 1e9                                    # beyond range_max
    cloud = synth.gaussian_cloud(pose, (0.3, 0.3, 0.2), 700, 1)
    cloud[::9] = synth.uniform_cloud(occ, 0.05, len(cloud[::9]), 2)
    cloud[:, 3] = 1.0 / len(cloud)
    lasers(r, g, 61, model, laser_pose=(0.1, 0.02, -0.03))
    out = [weights_after(s, cloud, scan, 8.0) for s in (r, g)]
    assert out[0][0] == out[1][0]
    assert_close(out[1][1], out[0][1])
    assert abs(out[1][2] - out[0][2]) <= RTOL * out[0][2]
    r.close(); g.close()


@pytest.mark.parametrize("model", ["lf", "lfp"])
def test_scan_with_nothing_to_integrate(gpu_lib, ref, model):
    """Every reading at or beyond range_max: lf/lfp skip them all, weights stay uniform."""
    occ = synth.rooms_map(200, seed=2)
    r, g = both(gpu_lib, ref, occ)
    scan = np.stack([np.full(30, 12.0), synth.scan_bearings(30)], 1).copy()
    cloud = synth.uniform_cloud(occ, 0.05, 333, 4)
    cloud[:, 3] = 1.0 / 333
    lasers(r, g, 30, model)
    out = [weights_after(s, cloud, scan, 12.0) for s in (r, g)]
    assert_close(out[1][1], out[0][1])
    assert np.allclose(out[1][1], 1.0 / 333, rtol=1e-12)
    r.close(); g.close()


@pytest.mark.parametrize("model,beams,max_beams", [("lf", 100, 30), ("lfp", 100, 30), ("beam", 77, 20),
                                                   ("lf", 31, 30), ("lfp", 31, 30)])
def test_ragged_beam_counts(gpu_lib, ref, model, beams, max_beams):
    """range_count not a multiple of the step (amcl_laser.cpp:246-250, :321)."""
    occ = synth.rooms_map(300, seed=5)
    r, g = both(gpu_lib, ref, occ)
    pose = synth.room_centre_pose(300, 0.05)
    scan = make_scan(r, pose, beams, range_max=9.0)
    cloud = synth.gaussian_cloud(pose, (0.2, 0.2, 0.1), 500, 3)
    cloud[:, 3] = 1.0 / 500
    lasers(r, g, max_beams, model)
    out = [weights_after(s, cloud, scan, 9.0) for s in (r, g)]
    assert_close(out[1][1], out[0][1])
    r.close(); g.close()


@pytest.mark.parametrize("model", ["lf", "lfp", "beam"])
def test_particles_off_the_map_and_single_particle(gpu_lib, ref, model):
    occ = synth.rooms_map(200, seed=6)
    r, g = both(gpu_lib, ref, occ)
    pose = synth.room_centre_pose(200, 0.05)
    scan = make_scan(r, pose, 40, range_max=6.0)
    cloud = np.array([[pose[0], pose[1], pose[2], 0.25], [50.0, -70.0, 1.0, 0.25],
                      [-5.2, 4.9, -2.0, 0.25], [1e6, 1e6, 0.3, 0.25]])
    lasers(r, g, 40, model)
    for c in (cloud, cloud[:1].copy()):
        c = c.copy()
        c[:, 3] = 1.0 / len(c)
        out = [weights_after(s, c, scan, 6.0, n_min=1) for s in (r, g)]
        assert_close(out[1][1], out[0][1])
    r.close(); g.close()


def test_min_equals_max_samples_resample(gpu_lib, ref):
    """KLD bound clamped from both sides: n stays max_samples whatever k is."""
    occ = synth.rooms_map(200, seed=7)
    r, g = both(gpu_lib, ref, occ)
    pose = synth.room_centre_pose(200, 0.05)
    cloud = synth.gaussian_cloud(pose, (0.05, 0.05, 0.02), 400, 5)
    w = np.random.default_rng(1).gamma(0.5, size=400)
    cloud[:, 3] = w / w.sum()
    res = []
    for s in (r, g):
        s.alloc(400, 400)
        s.set_samples(cloud)
        s.seed(9)
        s.resample()
        res.append((s.cur.sample_count, s.cur.kdtree.contents.leaf_count, s.samples().copy()))
    assert res[0][0] == res[1][0] == 400 and res[0][1] == res[1][1]
    assert np.array_equal(res[0][2][:, :3], res[1][2][:, :3])
    assert np.allclose(res[0][2][:, 3], res[1][2][:, 3], rtol=1e-15)
    r.close(); g.close()
