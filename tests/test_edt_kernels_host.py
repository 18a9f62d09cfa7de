"""The O(1)-per-cell distance-transform kernels (navigation_b200/csrc/edt_squares.cuh)
compiled for the host -- one OS thread per CUDA thread, tests/native/cuda_thread_shim.hpp
-- and run with the library's own launch geometry (edt_plan), against the oracle's
exact transform (oracle/amcl_oracle.c orc_edt_exact, the restatement of
map_update_cspace, map_cspace.cpp:120-176).  Integer work: bit-exact.  This is a
check of the kernel source where there is no GPU; the GPU tests
(tests/test_gpu_parity.py, test_k3_full_size_edt_bit_exact) check the compiled kernels.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from navigation_b200 import synth
from oracle import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def host_kernels(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("edt_host") / "edt_kernels_host.so")
    build = subprocess.run(["g++", "-O1", "-std=c++17", "-shared", "-fPIC", "-pthread", "-Wno-unknown-pragmas",
                            "-I" + os.path.join(ROOT, "tests", "native"),
                            "-I" + os.path.join(ROOT, "navigation_b200", "csrc"),
                            os.path.join(ROOT, "tests", "native", "edt_kernels_host.cpp"), "-o", so],
                           capture_output=True, text=True)
    assert build.returncode == 0, build.stderr[-3000:]
    lib = C.CDLL(so)
    lib.edt_kernels_host.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
    return lib


def tiled_codes(lib, occ, R, pack16):
    sy, sx = occ.shape
    tx, ty = (sx + 4) // 4, (sy + 4) // 4
    code = np.full(tx * ty * 16, 0x7777, np.uint16)
    assert lib.edt_kernels_host(occ.ctypes.data, sx, sy, R, pack16, code.ctypes.data) == 0
    return code.reshape(ty, tx, 4, 4).transpose(0, 2, 1, 3).reshape(ty * 4, tx * 4)


def expected_codes(occ, R):
    """d2 where <= R^2, R^2 + 1 beyond, R^2 + 2 in the padding outside the map."""
    sy, sx = occ.shape
    d2 = oracle.edt_exact(occ, R)
    tx, ty = (sx + 4) // 4, (sy + 4) // 4
    want = np.full((ty * 4, tx * 4), R * R + 2, np.int64)
    want[:sy, :sx] = np.where(d2 < 0, R * R + 1, d2)
    return want


def scattered(size_y, size_x, seed):
    rng = np.random.default_rng(seed)
    occ = np.full((size_y, size_x), -1, np.int8)
    occ[rng.random((size_y, size_x)) < 0.002] = 1
    occ[rng.random((size_y, size_x)) < 0.01] = 0          # unknown cells are not obstacles
    occ[size_y // 2, : size_x // 3] = 1
    return occ


@pytest.mark.parametrize("size_y,size_x,R,pack16", [
    (70, 208, 40, 1),      # the node's default radius (2.0 m / 0.05 m), rows a multiple of 16 cells
    (131, 203, 40, 0),     # ragged width and height: partial tiles, partial patches
    (64, 128, 7, 1),       # exactly one patch
    (65, 127, 0, 0),       # radius 0
    (120, 64, 160, 1),     # the largest radius the uint16 sums allow
    (40, 48, 3, 1),
])
def test_kernels_on_host_vs_oracle(host_kernels, size_y, size_x, R, pack16):
    occ = scattered(size_y, size_x, seed=size_x)
    assert np.array_equal(tiled_codes(host_kernels, occ, R, pack16), expected_codes(occ, R))


def test_kernels_on_host_rooms_empty_and_full(host_kernels):
    occ = np.ascontiguousarray(synth.rooms_map(160, seed=1, room=40, door=(10, 20)))
    assert np.array_equal(tiled_codes(host_kernels, occ, 12, 1), expected_codes(occ, 12))
    empty = np.full((48, 64), -1, np.int8)
    assert np.array_equal(tiled_codes(host_kernels, empty, 9, 1), expected_codes(empty, 9))
    full = np.ones((48, 64), np.int8)
    assert np.array_equal(tiled_codes(host_kernels, full, 9, 1), expected_codes(full, 9))


def test_radius_above_the_fast_limit_is_refused(host_kernels):
    occ = np.full((8, 16), -1, np.int8)
    code = np.zeros(((8 + 4) // 4) * ((16 + 4) // 4) * 16, np.uint16)
    limit = host_kernels.edt_kernels_host_max_radius()
    assert limit == 160
    assert host_kernels.edt_kernels_host(occ.ctypes.data, 16, 8, limit + 1, 1, code.ctypes.data) == -1


# ------------------------------------------------------- costmap inflation ----

def host_inflate(lib, grid, window, resolution, inscribed, radius, weight, inflate_unknown):
    """amclb200_inflate_costs's arithmetic around the kernels (inflation_dev.cu): cell
    radius (costmap_2d.cpp:181-185) and cost by squared cell distance
    (inflation_layer.h:104-119 through the oracle's orc_inflation_cost)."""
    g = np.array(grid, dtype=np.uint8, order="C", copy=True)
    sy, sx = g.shape
    w = (0, 0, sx, sy) if window is None else window
    R = int(max(0.0, np.ceil(radius / resolution)))
    lut = np.zeros(R * R + 1, np.uint8)
    for dx in range(R + 1):
        for dy in range(dx, R + 1):
            if dx * dx + dy * dy <= R * R:
                lut[dx * dx + dy * dy] = oracle.port().orc_inflation_cost(float(np.hypot(dx, dy)), resolution,
                                                                         inscribed, weight)
    lib.infl_kernels_host.argtypes = [C.c_void_p] + [C.c_int] * 7 + [C.c_void_p, C.c_int]
    rc = lib.infl_kernels_host(g.ctypes.data, sx, sy, w[0], w[1], w[2], w[3], R, lut.ctypes.data, int(inflate_unknown))
    assert rc >= 0, rc
    return g


@pytest.mark.parametrize("seed,inflate_unknown", [(0, False), (1, True)])
def test_inflation_kernels_on_host_vs_oracle(host_kernels, seed, inflate_unknown):
    """The cases of tests/test_gpu_inflation.py::test_random_grids_and_windows on a
    smaller grid: whole map, interior windows, windows at the corners, a one-cell
    and an empty window."""
    rng = np.random.default_rng(seed)
    g = rng.choice(np.array([0, 0, 0, 0, 0, 0, 60, 128, 252, 253, 255], dtype=np.uint8), size=(150, 203))
    g[rng.random(g.shape) < 0.01] = 254
    for window in (None, (30, 20, 80, 60), (0, 0, 5, 5), (195, 140, 203, 150), (100, 70, 101, 71), (50, 50, 50, 50)):
        got = host_inflate(host_kernels, g, window, 0.05, 0.12, 0.55, 10.0, inflate_unknown)
        want = oracle.inflate(g, window, 0.05, 0.12, 0.55, 10.0, inflate_unknown, exact=True)
        assert np.array_equal(got, want), window


def test_inflation_kernels_on_host_large_radius(host_kernels):
    g = np.zeros((90, 140), np.uint8)
    g[40, 60] = 254
    g[5, 130] = 254
    g[80:, :10] = 255
    for unknown in (False, True):
        got = host_inflate(host_kernels, g, (20, 10, 120, 70), 0.05, 0.3, 3.0, 3.0, unknown)     # 60 cells
        want = oracle.inflate(g, (20, 10, 120, 70), 0.05, 0.3, 3.0, 3.0, unknown, exact=True)
        assert np.array_equal(got, want)
