"""CPU: the restated map ingest / cloud packing (oracle/node_io_oracle.c)
against the golden vectors of the reference's own unit tests
(map_server/test/test_constants.cpp and utest.cpp, committed as
tests/golden/map_server_golden.npz by tests/golden/make_map_golden.py)."""
import os

import numpy as np
import pytest

from oracle import oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "map_server_golden.npz")


@pytest.fixture(scope="module")
def g():
    return dict(np.load(GOLD))


@pytest.mark.parametrize("image", ["testmap_png", "testmap_bmp"])
def test_loader_matches_reference_constants(g, image):
    negate, occ_th, free_th = g["params_testmap"]
    grid = oracle.image_to_grid(g["pixels_" + image], bool(negate), occ_th, free_th, "trinary")
    assert np.array_equal(grid, g["expect_testmap"])          # utest.cpp:43-58, :62-78


def test_loader_negate_testmap2(g):
    negate, occ_th, free_th = g["params_testmap2"]
    grid = oracle.image_to_grid(g["pixels_testmap2_png"], bool(negate), occ_th, free_th, "trinary")
    assert np.array_equal(grid, g["expect_testmap2"])         # rtest.cpp with testmap2.yaml


def test_loader_modes_on_spectrum(g):
    """utest.cpp:121-152."""
    px = g["pixels_spectrum_png"]
    tri = oracle.image_to_grid(px, False, 0.65, 0.1, "trinary").astype(np.uint8)
    n100, n0, n255 = g["spectrum_trinary_counts"]
    assert (np.sum(tri == 100), np.sum(tri == 0), np.sum(tri == 255)) == (n100, n0, n255)
    sc = oracle.image_to_grid(px, False, 0.65, 0.1, "scale").astype(np.uint8)
    assert np.sum(sc == 100) == n100 and np.sum(sc == 0) == n0
    assert np.sum((sc >= 1) & (sc < 100)) == n255
    raw = oracle.image_to_grid(px, False, 0.65, 0.1, "raw").astype(np.uint8)
    assert np.array_equal(np.bincount(raw.ravel(), minlength=256), np.ones(256, dtype=np.int64))


def test_convert_map(g):
    grid = np.array([[0, 100, -1, 50, 99, 1]], dtype=np.int8)
    assert oracle.convert_map(grid).tolist() == [[-1, 1, 0, 0, 0, 0]]  # amcl_node.cpp:1002-1009


def test_pack_cloud_layout():
    rng = np.random.default_rng(3)
    s = np.column_stack([rng.normal(size=50), rng.normal(size=50), rng.uniform(-7, 7, 50), np.full(50, 0.02)])
    out = oracle.pack_cloud(s, stride=3)
    assert out.shape == (17, 7)
    assert np.array_equal(out[:, 0], s[::3, 0]) and np.array_equal(out[:, 1], s[::3, 1])
    assert np.all(out[:, 2:5] == 0)
    assert np.allclose(out[:, 5], np.sin(s[::3, 2] / 2), rtol=0, atol=1e-16)
    assert np.allclose(out[:, 6], np.cos(s[::3, 2] / 2), rtol=0, atol=1e-16)
    assert oracle.pack_cloud(s, stride=1, max_poses=5).shape == (5, 7)
