"""Node-sized filters (amcl_node runs 500 .. 10 000 particles): the paths the
library takes only at that size -- the one-launch exact cumulative table, and
pf_update_resample's single trip to the host (hash histogram, the cut read from
the device, the launch chain replayed as a CUDA graph).  Same gates as the other
parity tests: integers bit-exact against the CPU oracle, statistics within 1e-5."""
import ctypes as C

import numpy as np
import pytest

from navigation_b200 import library, synth
from tests.scenario import Side
from tests.test_gpu_parity import RTOL, _check_resample_against_oracle, _weights

pytestmark = pytest.mark.gpu


@pytest.fixture()
def lib(gpu_lib):
    gpu_lib.amclb200_set_rng(library.RNG_DRAND48, 0)
    return gpu_lib


@pytest.mark.parametrize("kind,n,seed", [("gamma", 64, 1), ("gamma", 100, 2), ("uniform", 777, 3),
                                         ("dyadic", 4097, 4), ("tiny", 3000, 5), ("underflow", 9000, 6),
                                         ("posterior", 10000, 7), ("gamma", 16383, 8), ("underflow", 16384, 9)])
def test_small_cumulative_table_bit_exact(lib, port, kind, n, seed):
    """pf.c:399-402 through the one-block table (64 <= n <= 16384)."""
    occ = synth.rooms_map(80, seed=1, room=40, door=(10, 20))
    side = Side(lib, occ)
    side.alloc(10, n)
    cloud = np.zeros((n, 4))
    cloud[:, 3] = _weights(kind, n, seed)
    side.set_samples(cloud)
    before = lib.amclb200_cumsum_fallbacks()
    got = np.empty(n + 1)
    assert lib.amclb200_pf_cumulative_table(side.pf, got.ctypes.data, n + 1) == n + 1
    want = np.empty(n + 1)
    port.orc_cumsum(cloud, n, want)
    assert np.array_equal(got, want)
    assert lib.amclb200_cumsum_fallbacks() == before
    side.close()


def test_small_table_survives_bad_weights(lib, port):
    """Negative weights defeat the binade reasoning: the block's own recurrence
    check must notice and walk the table sequentially."""
    occ = synth.rooms_map(80, seed=1, room=40, door=(10, 20))
    n = 3000
    side = Side(lib, occ)
    side.alloc(10, n)
    cloud = np.zeros((n, 4))
    cloud[:, 3] = np.random.default_rng(9).normal(size=n)
    side.set_samples(cloud)
    before = lib.amclb200_cumsum_fallbacks()
    got = np.empty(n + 1)
    lib.amclb200_pf_cumulative_table(side.pf, got.ctypes.data, n + 1)
    want = np.empty(n + 1)
    port.orc_cumsum(cloud, n, want)
    assert np.array_equal(got, want)
    assert lib.amclb200_cumsum_fallbacks() == before + 1
    side.close()


def test_many_clusters_fall_back_to_the_long_ranking(lib, port):
    """A global cloud on a large map: thousands of one-bin clusters, more than the
    short root list of the single-trip path ranks (4096)."""
    occ = synth.rooms_map(2000, seed=3)
    n = 16000
    cloud = synth.uniform_cloud(occ, 0.05, n, 5)
    cloud[:, 3] = 1.0 / n
    cnt, k, ncl = _check_resample_against_oracle(lib, port, occ, cloud, 100, 0.01, 0.99, 77)
    assert cnt == n and ncl > 4096


def test_same_filter_resampled_again_after_its_parameters_change(lib, port):
    """The captured launch chain bakes pop_err / pop_z / min_samples in; amcl_node
    pokes them through the struct (amcl_node.cpp:615-616), so the next update must
    see the new values: every call is compared with the oracle run on its inputs."""
    occ = synth.rooms_map(400, seed=3)
    pose = synth.room_centre_pose(400, 0.05)
    n = 6000
    side = Side(lib, occ)
    side.alloc(100, n, pop_err=0.01, pop_z=0.99)
    for rnd, (pop_err, pop_z, min_s) in enumerate([(0.01, 0.99, 100), (0.01, 0.99, 100), (0.05, 0.95, 100),
                                                   (0.05, 0.95, 2500)]):
        side.pf.contents.pop_err = pop_err
        side.pf.contents.pop_z = pop_z
        side.pf.contents.min_samples = min_s
        cloud = synth.gaussian_cloud(pose, (0.3, 0.3, 0.15), n, 20 + rnd)
        w = np.random.default_rng(rnd).gamma(0.7, size=n)
        cloud[:, 3] = w / w.sum()
        side.set_samples(cloud)
        side.seed(300 + rnd)
        side.resample()
        out = side.samples()
        ob = np.empty((n, 4))
        osrc = np.empty(n, np.int32)
        k = C.c_int()
        port.orc_srand48(300 + rnd)
        cnt = port.orc_resample(cloud, n, ob, osrc, min_s, n, pop_err, pop_z, 0.0, None, None, C.byref(k))
        assert out.shape[0] == cnt
        assert side.cur.kdtree.contents.leaf_count == k.value
        assert np.array_equal(out, ob[:cnt])
        labels = np.empty(cnt, np.int32)
        assert side.cur.cluster_count == port.orc_cluster_labels(ob[:cnt].copy(), cnt, labels, None)
        assert side.pf.contents.converged == port.orc_update_converged(ob[:cnt].copy(), cnt, 0.5)
        tot = sum(t[0] for t in side.cluster_table())
        assert abs(tot - 1.0) < 1e-9
    side.close()
