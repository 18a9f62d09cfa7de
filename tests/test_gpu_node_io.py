"""GPU: map ingest (N4) and the published particle cloud (N2) through the C ABI,
against the oracle (oracle/node_io_oracle.c) and the golden vectors of the
reference's own map_server tests."""
import ctypes as C
import os

import numpy as np
import pytest

from navigation_b200 import library, synth
from oracle import oracle
from tests.scenario import Side

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "map_server_golden.npz")


@pytest.fixture(scope="module")
def g():
    return dict(np.load(GOLD))


def ingest(lib, pixels, negate, occ_th, free_th, mode, alpha_mask=None, res=0.1, origin=(0.0, 0.0, 0.0)):
    px = np.ascontiguousarray(pixels, dtype=np.uint8)
    h, w = px.shape[:2]
    ch = 1 if px.ndim == 2 else px.shape[2]
    if alpha_mask is None:
        alpha_mask = ch in (2, 4)
    grid = np.empty((h, w), dtype=np.int8)
    org = (C.c_double * 3)(*origin)
    m = lib.amclb200_map_from_image(px.ctypes.data, w, h, w * ch, ch, int(alpha_mask), res, org,
                                    int(negate), occ_th, free_th, oracle.MAP_MODES[mode], grid.ctypes.data)
    return m, grid


def occ_states(m):
    mp = m.contents
    n = mp.size_x * mp.size_y
    return np.array([mp.cells[k].occ_state for k in range(n)], dtype=np.int8).reshape(mp.size_y, mp.size_x)


def test_ingest_reference_test_images(gpu_lib, g):
    lib = gpu_lib
    for image, params, expect in (("testmap_png", "params_testmap", "expect_testmap"),
                                  ("testmap_bmp", "params_testmap", "expect_testmap"),
                                  ("testmap2_png", "params_testmap2", "expect_testmap2")):
        negate, occ_th, free_th = g[params]
        m, grid = ingest(lib, g["pixels_" + image], bool(negate), occ_th, free_th, "trinary",
                         origin=(-2.0, -3.0, -1.0))
        assert np.array_equal(grid, g[expect])
        assert np.array_equal(occ_states(m), oracle.convert_map(g[expect]))
        mp = m.contents
        # convertMap: origin moved to the map centre (amcl_node.cpp:996-997)
        assert mp.origin_x == -2.0 + (mp.size_x // 2) * 0.1 and mp.origin_y == -3.0 + (mp.size_y // 2) * 0.1
        lib.map_free(m)


def test_ingest_spectrum_modes(gpu_lib, g):
    px = g["pixels_spectrum_png"]
    for mode in ("trinary", "scale", "raw"):
        m, grid = ingest(gpu_lib, px, False, 0.65, 0.1, mode)
        assert np.array_equal(grid, oracle.image_to_grid(px, False, 0.65, 0.1, mode))
        gpu_lib.map_free(m)
    m, grid = ingest(gpu_lib, px, False, 0.65, 0.1, "trinary")
    u = grid.astype(np.uint8)
    assert (np.sum(u == 100), np.sum(u == 0), np.sum(u == 255)) == tuple(g["spectrum_trinary_counts"])
    gpu_lib.map_free(m)


@pytest.mark.parametrize("channels,alpha_mask", [(1, False), (3, False), (4, True), (4, False), (2, True)])
def test_ingest_random_images_bit_exact(gpu_lib, channels, alpha_mask):
    rng = np.random.default_rng(channels * 7 + alpha_mask)
    shape = (1103, 1517) if channels == 1 else (1103, 1517, channels)
    px = rng.integers(0, 256, size=shape, dtype=np.uint8)
    for mode in ("trinary", "scale", "raw"):
        for negate in (False, True):
            want = oracle.image_to_grid(px, negate, 0.65, 0.196, mode, has_alpha_mask=alpha_mask)
            m, grid = ingest(gpu_lib, px, negate, 0.65, 0.196, mode, alpha_mask=alpha_mask)
            assert np.array_equal(grid, want), (mode, negate)
            gpu_lib.map_free(m)


def test_ingested_map_drives_the_filter_like_an_uploaded_one(gpu_lib):
    """The occupancy the kernel left on the device serves the distance transform
    and the ray caster exactly like one uploaded from host cells."""
    lib = gpu_lib
    occ = synth.rooms_map(400, seed=5, unknown_ring=True)
    # an image that maps back to this occupancy: occupied -> black, free -> white, unknown -> mid grey
    img = np.where(occ == 1, 0, np.where(occ == -1, 254, 205)).astype(np.uint8)[::-1]
    m, grid = ingest(lib, img, False, 0.65, 0.196, "trinary", res=0.05,
                     origin=(-(400 // 2) * 0.05, -(400 // 2) * 0.05, 0.0))
    assert np.array_equal(oracle.convert_map(grid), occ)
    side = Side(lib, occ)
    side.update_cspace(1.0)
    lib.map_update_cspace(m, 1.0)
    got = np.empty((400, 400))
    lib.amcl_c_map_get_occ_dist(m, got.ctypes.data)
    assert np.array_equal(got, side.occ_dist())
    rng = np.random.default_rng(1)
    rays = np.stack([rng.uniform(-9, 9, 2000), rng.uniform(-9, 9, 2000), rng.uniform(-4, 4, 2000)], 1).copy()
    a, b = np.empty(2000), np.empty(2000)
    lib.amclb200_map_calc_ranges(m, rays.ctypes.data, 2000, 8.0, a.ctypes.data)
    lib.amclb200_map_calc_ranges(side.map, rays.ctypes.data, 2000, 8.0, b.ctypes.data)
    assert np.array_equal(a, b)
    side.close()
    lib.map_free(m)


def test_device_authoritative_map(gpu_lib):
    """amclb200_map_set_mirror(map, 0): map_update_cspace neither re-reads the host
    cells nor writes occ_dist; amclb200_map_pull_distances fills them on demand."""
    lib = gpu_lib
    occ = synth.rooms_map(300, seed=8)
    img = np.where(occ == 1, 0, np.where(occ == -1, 254, 205)).astype(np.uint8)[::-1]
    m, _ = ingest(lib, img, False, 0.65, 0.196, "trinary", res=0.05)
    lib.amclb200_map_set_mirror(m, 0)
    lib.map_update_cspace(m, 1.5)
    got = np.empty((300, 300))
    lib.amcl_c_map_get_occ_dist(m, got.ctypes.data)
    assert np.all(got == 0.0) and m.contents.max_occ_dist == 1.5
    lib.amclb200_map_pull_distances(m)
    lib.amcl_c_map_get_occ_dist(m, got.ctypes.data)
    side = Side(lib, occ)
    side.update_cspace(1.5)
    assert np.array_equal(got, side.occ_dist())
    side.close()
    lib.map_free(m)


@pytest.mark.parametrize("n,stride", [(1000, 1), (100003, 7), (5, 10)])
def test_pack_cloud_vs_oracle(gpu_lib, n, stride):
    lib = gpu_lib
    lib.amclb200_set_rng(library.RNG_DRAND48, 0)
    occ = synth.rooms_map(200, seed=1)
    side = Side(lib, occ)
    side.alloc(10, n)
    rng = np.random.default_rng(n)
    cloud = np.column_stack([rng.uniform(-4, 4, n), rng.uniform(-4, 4, n), rng.uniform(-9, 9, n), np.full(n, 1.0 / n)])
    side.set_samples(cloud)
    want = oracle.pack_cloud(cloud, stride)
    out = np.empty((want.shape[0] + 3, 7))
    cnt = lib.amclb200_pf_pack_cloud(side.pf, stride, out.shape[0], out.ctypes.data)
    assert cnt == want.shape[0]
    got = out[:cnt]
    assert np.array_equal(got[:, :3], want[:, :3])
    assert np.array_equal(got[:, 3:5], want[:, 3:5])          # exact zeros
    assert np.max(np.abs(got[:, 5:] - want[:, 5:])) < 4e-16   # device vs host sin/cos
    cnt2 = lib.amclb200_pf_pack_cloud(side.pf, stride, 2, out.ctypes.data)
    assert cnt2 == min(2, want.shape[0])
    side.close()
