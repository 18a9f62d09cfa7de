"""CPU: the restated InflationLayer::updateCosts (oracle/node_io_oracle.c) against
the known answers of the reference's own tests
(/root/reference/costmap_2d/test/inflation_tests.cpp), restated on bare grids:
the tests there place lethal cells through an ObstacleLayer on an all-free
master grid, which is what the arrays below are.  The rostest sets
inflation/cost_scaling_factor = 1 (costmap_2d/test/inflation_tests.launch:4), hence W below."""
import numpy as np
import pytest

from oracle import oracle

NO_INFORMATION, LETHAL, INSCRIBED, FREE = 255, 254, 253, 0
W = 1.0


def cost(d, res, inscribed, weight=W):
    return oracle.port().orc_inflation_cost(float(d), res, inscribed, weight)


def grid_with(shape, lethal):
    g = np.zeros(shape, dtype=np.uint8)
    for (x, y) in lethal:
        g[y, x] = LETHAL
    return g


def test_adjacent_to_obstacle_can_still_move():
    """inflation_tests.cpp:137-161 (inscribed 2.1, inflation radius 4.1, resolution 1)."""
    out = oracle.inflate(grid_with((10, 10), [(0, 0)]), None, 1.0, 2.1, 4.1, W)
    assert out[0, 0] == LETHAL
    assert out[0, 1] == INSCRIBED and out[0, 2] == INSCRIBED
    assert out[0, 3] < INSCRIBED and out[1, 2] < INSCRIBED
    assert out[1, 1] == INSCRIBED


def test_inflation_should_not_create_unknowns():
    """inflation_tests.cpp:163-182."""
    out = oracle.inflate(grid_with((10, 10), [(0, 0)]), None, 1.0, 2.1, 4.1, W)
    assert np.sum(out == NO_INFORMATION) == 0


def test_cost_function_correctness():
    """inflation_tests.cpp:188-247 (inscribed 5.0, inflation radius 10.5, obstacle at (50, 50))."""
    out = oracle.inflate(grid_with((100, 100), [(50, 50)]), None, 1.0, 5.0, 10.5, W)
    for i in range(0, 6):
        for (x, y) in ((50 + i, 50), (50 - i, 50), (50, 50 + i), (50, 50 - i)):
            assert out[y, x] >= INSCRIBED
    for i in range(6, 12):
        assert out[50, 50 + i] == cost(i, 1.0, 5.0)


def validate_point_inflation(out, mx, my, res, inscribed, radius_cells):
    """validatePointInflation (inflation_tests.cpp:73-135): every cell within the
    radius of the obstacle costs at least what that obstacle alone implies."""
    sy, sx = out.shape
    for y in range(sy):
        for x in range(sx):
            d = np.hypot(x - mx, y - my)
            if d <= radius_cells:
                assert out[y, x] >= cost(d, res, inscribed)


def test_inflation_order_correctness():
    """inflation_tests.cpp:256-282: two diagonal obstacles."""
    out = oracle.inflate(grid_with((10, 10), [(4, 4), (5, 5)]), None, 1.0, 2.1, 4.1, W)
    validate_point_inflation(out, 4, 4, 1.0, 2.1, 4.1)
    validate_point_inflation(out, 5, 5, 1.0, 2.1, 4.1)


def test_inflation3_counts_and_idempotence():
    """inflation_tests.cpp:377-410: empty 10 x 10 map, inscribed 1, inflation radius 3, obstacle at (5, 5)."""
    g = grid_with((10, 10), [(5, 5)])
    for exact in (False, True):
        out = oracle.inflate(g, None, 1.0, 1.0, 3.0, W, exact=exact)
        assert np.sum(out != FREE) == 29
        assert np.sum(out == LETHAL) == 1 and np.sum(out == INSCRIBED) == 4
        again = oracle.inflate(out, None, 1.0, 1.0, 3.0, W, exact=exact)     # "should see no change"
        assert np.array_equal(again, out)


def static_master():
    """The master grid inflation_tests.cpp's testInflation starts from: map_server's
    rule on TenByTen.pgm (tests/golden/map_server_golden.npz), then StaticLayer's
    100 -> LETHAL_OBSTACLE, everything else free (track_unknown_space false)."""
    import os
    g = dict(np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "map_server_golden.npz")))
    negate, occ_th, free_th = g["params_tenbyten"]
    grid = oracle.image_to_grid(g["pixels_tenbyten_pgm"], bool(negate), occ_th, free_th, "trinary")
    return np.where(grid == 100, LETHAL, FREE).astype(np.uint8)


def test_inflation_on_the_static_test_map():
    """inflation_tests.cpp:284-303 (testInflation, first update; radii 1, 1, 1): 20 lethal and 28
    inscribed cells -- the map_server loader and the inflation layer chained."""
    out = oracle.inflate(static_master(), None, 1.0, 1.0, 1.0, W)
    assert np.sum(out == LETHAL) == 20 and np.sum(out == INSCRIBED) == 28


def test_inflation2_l_shape_on_the_static_map():
    """inflation_tests.cpp:348-372 (testInflation2)."""
    m = static_master()
    for (x, y) in ((1, 1), (2, 1), (2, 2)):
        m[y, x] = LETHAL
    out = oracle.inflate(m, None, 1.0, 1.0, 1.0, W)
    assert out[3, 2] == INSCRIBED and out[3, 3] == INSCRIBED


@pytest.mark.parametrize("seed,inflate_unknown", [(0, False), (1, True), (2, False)])
def test_sweep_equals_nearest_obstacle_form(seed, inflate_unknown):
    """The bin sweep and the exhaustive nearest-lethal-cell form agree on random
    grids with unknown cells and pre-existing costs (windowed update included)."""
    rng = np.random.default_rng(seed)
    g = rng.choice(np.array([0, 0, 0, 0, 0, 0, 60, 128, 252, 253, 255], dtype=np.uint8), size=(90, 120))
    g[rng.random(g.shape) < 0.01] = LETHAL
    for window in (None, (30, 20, 80, 60), (0, 0, 5, 5), (110, 80, 120, 90)):
        a = oracle.inflate(g, window, 0.05, 0.12, 0.55, 10.0, inflate_unknown)
        b = oracle.inflate(g, window, 0.05, 0.12, 0.55, 10.0, inflate_unknown, exact=True)
        assert np.all(b >= a) or inflate_unknown     # exact distance <= propagated distance
        assert np.mean(a != b) < 1e-3
