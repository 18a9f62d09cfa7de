"""TEST INFRASTRUCTURE: loaders for the two CPU oracles.

* ``port()``  -- oracle/_build/libamcl_oracle.so, the C restatement in
  amcl_oracle.c (always buildable; travels to the GPU box as a built .so).
* ``ref()``   -- oracle/_ref/libamcl_ref.so, the unmodified reference compiled
  by oracle/Makefile (exists wherever it was built while /root/reference was
  mounted; None otherwise).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs import this module.  The product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "_build", "libamcl_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libamcl_ref.so")

_port = None
_ref = None

f64p = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
i8p = np.ctypeslib.ndpointer(dtype=np.int8, flags="C_CONTIGUOUS")


def build(verbose=False):
    """Compile the port and, when the reference tree is mounted, the reference."""
    out = subprocess.run(["make", "-C", HERE, "all"], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + out.stdout + out.stderr)
    if verbose:
        print(out.stdout)


def port():
    global _port
    if _port is not None:
        return _port
    if not os.path.exists(PORT_SO):
        build()
    lib = C.CDLL(PORT_SO)
    sig = {
        "orc_srand48": (None, [C.c_long]),
        "orc_drand48": (C.c_double, []),
        "orc_get_state": (C.c_uint64, []),
        "orc_set_state": (None, [C.c_uint64]),
        "orc_draw_count": (C.c_uint64, []),
        "orc_reset_count": (None, []),
        "orc_ran_gaussian": (C.c_double, [C.c_double]),
        "orc_ran_gaussian_parts": (None, [C.POINTER(C.c_double), C.POINTER(C.c_double)]),
        "orc_coord_add": (None, [f64p, f64p, f64p]),
        "orc_eigen3": (None, [f64p, f64p, f64p]),
        "orc_pf_init": (None, [f64p, C.c_int, f64p, f64p, C.c_long]),
        "orc_cell_radius": (C.c_int, [C.c_double, C.c_double]),
        "orc_edt_exact": (None, [i8p, C.c_int, C.c_int, C.c_int, i32p]),
        "orc_edt_bruteforce": (None, [i8p, C.c_int, C.c_int, C.c_int, i32p]),
        "orc_d2_to_occ_dist": (None, [i32p, C.c_size_t, C.c_double, C.c_double, f64p]),
        "orc_calc_range": (C.c_double, [i8p, C.c_int, C.c_int] + [C.c_double] * 7),
        "orc_laser_model": (
            C.c_double,
            [C.c_int, f64p, C.c_int, f64p, C.c_int, C.c_int, C.c_int, C.c_void_p, i8p,
             C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double, C.c_double,
             f64p, C.c_int, C.c_double],
        ),
        "orc_sensor_finish": (
            C.c_double,
            [f64p, C.c_int, C.c_double, C.c_double, C.c_double,
             C.POINTER(C.c_double), C.POINTER(C.c_double)],
        ),
        "orc_motion": (None, [f64p, C.c_int, C.c_int, f64p, f64p, f64p]),
        "orc_resample_limit": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_double, C.c_double]),
        "orc_bin_key": (None, [f64p, i32p]),
        "orc_resample": (
            C.c_int,
            [f64p, C.c_int, f64p, i32p, C.c_int, C.c_int, C.c_double, C.c_double,
             C.c_double, C.c_void_p, C.c_void_p, C.POINTER(C.c_int)],
        ),
        "orc_cumsum": (None, [f64p, C.c_int, f64p]),
        "orc_cluster_labels": (C.c_int, [f64p, C.c_int, i32p, C.POINTER(C.c_int)]),
        "orc_cluster_stats": (None, [f64p, C.c_int, i32p, C.c_int, f64p, f64p]),
        "orc_update_converged": (C.c_int, [f64p, C.c_int, C.c_double]),
        "orc_cep_stats": (None, [f64p, C.c_int, f64p]),
        "orc_bin_first_occurrence": (C.c_int, [f64p, C.c_int, i32p]),
        "orc_image_to_grid": (None, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                     C.c_double, C.c_double, C.c_int, C.c_void_p]),
        "orc_convert_map": (None, [C.c_void_p, C.c_size_t, i8p]),
        "orc_pack_cloud": (C.c_int, [f64p, C.c_int, C.c_int, C.c_int, f64p]),
        "orc_inflation_cost": (C.c_ubyte, [C.c_double] * 4),
        "orc_cell_distance": (C.c_uint, [C.c_double, C.c_double]),
        "orc_inflate": (None, [C.c_void_p] + [C.c_int] * 6 + [C.c_double] * 4 + [C.c_int]),
        "orc_inflate_exact": (None, [C.c_void_p] + [C.c_int] * 6 + [C.c_double] * 4 + [C.c_int]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _port = lib
    return lib


def ref():
    """The compiled reference (bound with the shared ABI binder) or None."""
    global _ref
    if _ref is not None:
        return _ref
    if not os.path.exists(REF_SO):
        if os.path.exists("/root/reference/amcl/src/amcl/pf/pf.c"):
            build()
        if not os.path.exists(REF_SO):
            return None
    import sys

    sys.path.insert(0, os.path.dirname(HERE))
    from navigation_b200 import abi

    lib = abi.bind(C.CDLL(REF_SO))
    lib.tap_seed.argtypes = [C.c_long]
    lib.tap_seed.restype = None
    lib.tap_set_ignore_srand.argtypes = [C.c_int]
    lib.tap_get_state.restype = C.c_uint64
    lib.tap_set_state.argtypes = [C.c_uint64]
    lib.tap_draw_count.restype = C.c_uint64
    lib.tap_log_begin.argtypes = [C.c_uint64]
    lib.tap_log_size.restype = C.c_uint64
    lib.tap_log_copy.argtypes = [C.c_void_p, C.c_uint64]
    lib.drand48.restype = C.c_double
    _ref = lib
    return lib


# ---- numpy-level helpers over the port ---------------------------------------

def laser_params(laser_pose=(0.0, 0.0, 0.0), z_hit=0.95, z_short=0.1, z_max=0.05,
                 z_rand=0.05, sigma_hit=0.2, lambda_short=0.1, beam_skip_distance=0.5,
                 beam_skip_threshold=0.3, beam_skip_error_threshold=0.9):
    return np.array([laser_pose[0], laser_pose[1], laser_pose[2], z_hit, z_short, z_max,
                     z_rand, sigma_hit, lambda_short, beam_skip_distance,
                     beam_skip_threshold, beam_skip_error_threshold], dtype=np.float64)


def laser_model(model, samples, lp, max_beams, occ, occ_dist, scale, origin, max_occ_dist,
                ranges_rb, range_max, do_beamskip=0, converged=0, temp_obs=None):
    """Runs one of the three laser models in place on samples [n,4]; returns total."""
    lib = port()
    sy, sx = occ.shape
    od = occ_dist.ctypes.data if occ_dist is not None else None
    to = temp_obs.ctypes.data if temp_obs is not None else None
    return lib.orc_laser_model(model, samples, samples.shape[0], lp, max_beams, do_beamskip,
                               converged, to, occ, od, sx, sy, origin[0], origin[1], scale,
                               max_occ_dist, np.ascontiguousarray(ranges_rb), ranges_rb.shape[0],
                               range_max)


def edt_exact(occ, R):
    sy, sx = occ.shape
    d2 = np.empty((sy, sx), dtype=np.int32)
    port().orc_edt_exact(occ, sx, sy, R, d2)
    return d2


def occ_dist_from_d2(d2, scale, max_occ_dist):
    out = np.empty(d2.shape, dtype=np.float64)
    port().orc_d2_to_occ_dist(np.ascontiguousarray(d2), d2.size, scale, max_occ_dist, out)
    return out


# ---- either side of the filter (node_io_oracle.c) ---------------------------------
MAP_MODES = {"trinary": 0, "scale": 1, "raw": 2}


def image_to_grid(pixels, negate=False, occ_th=0.65, free_th=0.196, mode="trinary", has_alpha_mask=None):
    """map_server's loader on decoded pixels (h, w) or (h, w, channels) uint8 -> int8 grid (h, w), bottom row first."""
    px = np.ascontiguousarray(pixels, dtype=np.uint8)
    h, w = px.shape[:2]
    ch = 1 if px.ndim == 2 else px.shape[2]
    if has_alpha_mask is None:
        has_alpha_mask = ch in (2, 4)
    grid = np.empty((h, w), dtype=np.int8)
    port().orc_image_to_grid(px.ctypes.data, w, h, w * ch, ch, int(has_alpha_mask), int(negate),
                             occ_th, free_th, MAP_MODES[mode], grid.ctypes.data)
    return grid


def convert_map(grid):
    g = np.ascontiguousarray(grid, dtype=np.int8)
    out = np.empty(g.shape, dtype=np.int8)
    port().orc_convert_map(g.ctypes.data, g.size, out)
    return out


def pack_cloud(samples, stride=1, max_poses=None):
    s = np.ascontiguousarray(samples, dtype=np.float64)
    n = s.shape[0]
    cap = (n + stride - 1) // stride if max_poses is None else max_poses
    out = np.empty((max(cap, 1), 7), dtype=np.float64)
    cnt = port().orc_pack_cloud(s, n, stride, cap, out)
    return out[:cnt]


def inflate(master, window=None, resolution=1.0, inscribed_radius=0.0, inflation_radius=0.55,
            weight=10.0, inflate_unknown=False, exact=False):
    """InflationLayer::updateCosts on a (size_y, size_x) uint8 cost grid; returns a new grid.
    window = (min_i, min_j, max_i, max_j), default the whole map."""
    g = np.array(master, dtype=np.uint8, order="C", copy=True)
    sy, sx = g.shape
    w = (0, 0, sx, sy) if window is None else window
    fn = port().orc_inflate_exact if exact else port().orc_inflate
    fn(g.ctypes.data, sx, sy, w[0], w[1], w[2], w[3], resolution, inscribed_radius, inflation_radius,
       weight, int(inflate_unknown))
    return g
