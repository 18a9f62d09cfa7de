"""GPU: amclb200_inflate_costs (costmap_2d::InflationLayer::updateCosts) against
the oracle.  Bit-exact against the nearest-lethal-cell restatement; against the
reference's bin sweep the GPU cost is never lower and differs in a few cells
per 10^5 (the sweep's carried source is not always the nearest obstacle -- the
same kind of gap as SURVEY.md section 0, C1).  The known answers of the
reference's own tests (costmap_2d/test/inflation_tests.cpp) are asserted on the
GPU output directly."""
import numpy as np
import pytest
import torch

from oracle import oracle

pytestmark = pytest.mark.gpu
NO_INFORMATION, LETHAL, INSCRIBED = 255, 254, 253


def gpu_inflate(lib, grid, window=None, resolution=1.0, inscribed=0.0, radius=0.55, weight=10.0,
                inflate_unknown=False, on_device=False):
    g = np.array(grid, dtype=np.uint8, order="C", copy=True)
    sy, sx = g.shape
    w = (0, 0, sx, sy) if window is None else window
    if on_device:
        t = torch.from_numpy(g).cuda()
        lib.amclb200_inflate_costs(t.data_ptr(), 1, sx, sy, w[0], w[1], w[2], w[3], resolution, inscribed,
                                   radius, weight, int(inflate_unknown))
        return t.cpu().numpy()
    lib.amclb200_inflate_costs(g.ctypes.data, 0, sx, sy, w[0], w[1], w[2], w[3], resolution, inscribed,
                               radius, weight, int(inflate_unknown))
    return g


def grid_with(shape, lethal):
    g = np.zeros(shape, dtype=np.uint8)
    for (x, y) in lethal:
        g[y, x] = LETHAL
    return g


W = 1.0    # costmap_2d/test/inflation_tests.launch:4 sets inflation/cost_scaling_factor = 1


def test_reference_test_cases(gpu_lib):
    """inflation_tests.cpp:137-282 and :377-410 on the GPU output."""
    out = gpu_inflate(gpu_lib, grid_with((10, 10), [(0, 0)]), None, 1.0, 2.1, 4.1, W)
    assert out[0, 0] == LETHAL and out[0, 1] == INSCRIBED and out[0, 2] == INSCRIBED
    assert out[0, 3] < INSCRIBED and out[1, 2] < INSCRIBED and out[1, 1] == INSCRIBED
    assert np.sum(out == NO_INFORMATION) == 0
    out = gpu_inflate(gpu_lib, grid_with((100, 100), [(50, 50)]), None, 1.0, 5.0, 10.5, W)
    for i in range(0, 6):
        for (x, y) in ((50 + i, 50), (50 - i, 50), (50, 50 + i), (50, 50 - i)):
            assert out[y, x] >= INSCRIBED
    for i in range(6, 12):
        assert out[50, 50 + i] == oracle.port().orc_inflation_cost(float(i), 1.0, 5.0, W)
    out = gpu_inflate(gpu_lib, grid_with((10, 10), [(4, 4), (5, 5)]), None, 1.0, 2.1, 4.1, W)
    for (mx, my) in ((4, 4), (5, 5)):                       # validatePointInflation
        for y in range(10):
            for x in range(10):
                d = np.hypot(x - mx, y - my)
                if d <= 4.1:
                    assert out[y, x] >= oracle.port().orc_inflation_cost(float(d), 1.0, 2.1, W)
    assert np.array_equal(out, oracle.inflate(grid_with((10, 10), [(4, 4), (5, 5)]), None, 1.0, 2.1, 4.1, W))
    # testInflation3 (:377-410): counts after one update, no change after a second
    out = gpu_inflate(gpu_lib, grid_with((10, 10), [(5, 5)]), None, 1.0, 1.0, 3.0, W)
    assert np.sum(out != 0) == 29 and np.sum(out == LETHAL) == 1 and np.sum(out == INSCRIBED) == 4
    assert np.array_equal(gpu_inflate(gpu_lib, out, None, 1.0, 1.0, 3.0, W), out)


def test_loader_and_inflation_chained_on_the_static_test_map(gpu_lib):
    """inflation_tests.cpp:284-303 and :348-372: TenByTen.pgm through the device map
    ingest, 100 -> lethal as StaticLayer does, then the device inflation: 20 lethal and
    28 inscribed cells, and the L-shape assertions of testInflation2."""
    import ctypes as C
    import os
    g = dict(np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "map_server_golden.npz")))
    px = np.ascontiguousarray(g["pixels_tenbyten_pgm"], dtype=np.uint8)
    negate, occ_th, free_th = g["params_tenbyten"]
    grid = np.empty(px.shape, dtype=np.int8)
    m = gpu_lib.amclb200_map_from_image(px.ctypes.data, px.shape[1], px.shape[0], px.shape[1], 1, 0, 1.0,
                                        (C.c_double * 3)(0, 0, 0), int(negate), occ_th, free_th, 0, grid.ctypes.data)
    gpu_lib.map_free(m)
    master = np.where(grid == 100, LETHAL, 0).astype(np.uint8)
    out = gpu_inflate(gpu_lib, master, None, 1.0, 1.0, 1.0, W)
    assert np.sum(out == LETHAL) == 20 and np.sum(out == INSCRIBED) == 28
    for (x, y) in ((1, 1), (2, 1), (2, 2)):
        master[y, x] = LETHAL
    out = gpu_inflate(gpu_lib, master, None, 1.0, 1.0, 1.0, W)
    assert out[3, 2] == INSCRIBED and out[3, 3] == INSCRIBED


@pytest.mark.parametrize("seed,inflate_unknown,on_device", [(0, False, False), (1, True, False), (2, False, True)])
def test_random_grids_and_windows(gpu_lib, seed, inflate_unknown, on_device):
    rng = np.random.default_rng(seed)
    g = rng.choice(np.array([0, 0, 0, 0, 0, 0, 60, 128, 252, 253, 255], dtype=np.uint8), size=(317, 431))
    g[rng.random(g.shape) < 0.01] = LETHAL
    for window in (None, (30, 20, 80, 60), (0, 0, 5, 5), (420, 300, 431, 317), (200, 100, 201, 101), (50, 50, 50, 50)):
        got = gpu_inflate(gpu_lib, g, window, 0.05, 0.12, 0.55, 10.0, inflate_unknown, on_device)
        exact = oracle.inflate(g, window, 0.05, 0.12, 0.55, 10.0, inflate_unknown, exact=True)
        sweep = oracle.inflate(g, window, 0.05, 0.12, 0.55, 10.0, inflate_unknown)
        assert np.array_equal(got, exact), window
        if not inflate_unknown:
            assert np.all(got >= sweep)
        assert np.mean(got != sweep) < 1e-3


def test_edge_cases(gpu_lib):
    g = np.zeros((40, 50), dtype=np.uint8)
    assert np.array_equal(gpu_inflate(gpu_lib, g, None, 0.05, 0.1, 0.0), g)          # radius 0: early return
    assert np.array_equal(gpu_inflate(gpu_lib, g, None, 0.05, 0.1, 0.55), g)         # no lethal cell
    full = np.full((40, 50), LETHAL, dtype=np.uint8)
    assert np.array_equal(gpu_inflate(gpu_lib, full, None, 0.05, 0.1, 0.55), full)
    unk = np.full((40, 50), NO_INFORMATION, dtype=np.uint8)
    unk[20, 25] = LETHAL
    for flag in (False, True):
        assert np.array_equal(gpu_inflate(gpu_lib, unk, None, 0.05, 0.12, 0.55, 10.0, flag),
                              oracle.inflate(unk, None, 0.05, 0.12, 0.55, 10.0, flag, exact=True))
    one = np.array([[LETHAL]], dtype=np.uint8)
    assert gpu_inflate(gpu_lib, one, None, 0.05, 0.1, 0.55)[0, 0] == LETHAL


def test_full_size_properties(gpu_lib):
    """4000 x 4000 at 0.05 m (the K3 map): idempotent, monotone, and every cell
    within the inscribed radius of a lethal cell is at least INSCRIBED."""
    from navigation_b200 import synth
    occ = synth.rooms_map(4000, seed=1)
    g = np.where(occ == 1, LETHAL, 0).astype(np.uint8)
    once = gpu_inflate(gpu_lib, g, None, 0.05, 0.2, 0.55, 10.0)
    twice = gpu_inflate(gpu_lib, once, None, 0.05, 0.2, 0.55, 10.0)
    assert np.array_equal(once, twice)
    assert np.all(once >= g)
    ys, xs = np.nonzero(g[1000:1100, 1000:1100] == LETHAL)
    for y, x in zip(ys[:50], xs[:50]):
        patch = once[1000 + y - 4:1000 + y + 5, 1000 + x - 4:1000 + x + 5]
        yy, xx = np.mgrid[-4:5, -4:5]
        assert np.all(patch[np.hypot(yy, xx) * 0.05 <= 0.2] >= INSCRIBED)
    sub = oracle.inflate(g[:300, :300], None, 0.05, 0.2, 0.55, 10.0, exact=True)
    assert np.array_equal(once[:280, :280], sub[:280, :280])


def test_radius_above_the_register_block_limit(gpu_lib):
    """200 cells > kFastRadius (160): the per-cell search kernels (infl_row_dist, infl_col_apply)."""
    g = np.zeros((260, 300), dtype=np.uint8)
    g[130, 150] = LETHAL
    g[10, 290] = LETHAL
    got = gpu_inflate(gpu_lib, g, (100, 100, 200, 160), 0.01, 0.05, 2.0, 3.0)
    want = oracle.inflate(g, (100, 100, 200, 160), 0.01, 0.05, 2.0, 3.0, exact=True)
    assert np.array_equal(got, want)
