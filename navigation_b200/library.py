"""Loader for libamcl_b200.so (the CUDA implementation of the AMCL hot path).

There is no CPU fallback: if the shared library is missing, ``load()`` raises,
and the library itself aborts when no CUDA device is present the first time a
map or filter is created.
"""
import ctypes as C
import os
import subprocess

from . import abi

HERE = os.path.dirname(os.path.abspath(__file__))
# AMCLB200_LIB names another build of the same library (kernel-variant experiments)
LIB_PATH = os.environ.get("AMCLB200_LIB") or os.path.join(HERE, "lib", "libamcl_b200.so")
CSRC = os.path.join(HERE, "csrc")

_lib = None

# Part 3 of include/amcl_b200.h
EXTENSION_SYMBOLS = {
    "amclb200_version": (C.c_char_p, []),
    "amclb200_set_stream": (None, [C.c_void_p]),
    "amclb200_set_rng": (None, [C.c_int, C.c_uint64]),
    "amclb200_set_reseed": (None, [C.c_int]),
    "amclb200_get_reseed": (C.c_int, []),
    "amclb200_set_mirror": (None, [abi.pf_p, C.c_int]),
    "amclb200_pf_pull_samples": (None, [abi.pf_p]),
    "amclb200_pf_push_samples": (None, [abi.pf_p]),
    "amclb200_pf_device_samples": (C.c_void_p, [abi.pf_p, C.c_int]),
    "amclb200_map_push_distances": (C.c_int, [abi.map_p]),
    "amclb200_pf_resample_sources": (C.c_int, [abi.pf_p, C.c_void_p, C.c_int]),
    "amclb200_pf_cluster_labels": (C.c_int, [abi.pf_p, C.c_void_p, C.c_int]),
    "amclb200_pf_cumulative_table": (C.c_int, [abi.pf_p, C.c_void_p, C.c_int]),
    "amclb200_pf_last_timings": (None, [abi.pf_p, C.POINTER(C.c_float)]),
    "amclb200_map_last_edt_ms": (C.c_double, [abi.map_p]),
    "amclb200_map_calc_ranges": (None, [abi.map_p, C.c_void_p, C.c_int, C.c_double, C.c_void_p]),
    "amclb200_pf_load_device": (None, [abi.pf_p, C.c_void_p, C.c_int]),
    "amclb200_pf_set_shard": (None, [abi.pf_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "amclb200_pf_enable_peer_access": (None, [abi.pf_p]),
    "amclb200_nccl_unique_id": (C.c_int, [C.c_void_p]),
    "amclb200_pf_set_shard_nccl": (None, [abi.pf_p, C.c_int, C.c_int, C.c_void_p]),
    "amclb200_pf_load_device_shard": (None, [abi.pf_p, C.c_void_p, C.c_int]),
    "amclb200_pf_shard_range": (None, [abi.pf_p, C.c_int, C.POINTER(C.c_int)]),
    "amclb200_launch_count": (C.c_uint64, []),
    "amclb200_pf_best_cluster": (C.c_int, [abi.pf_p, abi.c_double_p, C.POINTER(abi.pf_vector_t), C.POINTER(abi.pf_matrix_t)]),
    "amclb200_cumsum_fallbacks": (C.c_uint64, []),
    "amclb200_set_parallel_cumsum": (None, [C.c_int]),
    "amclb200_cumsum_last_tiles": (None, [C.POINTER(C.c_int)]),
    "amclb200_pf_sensor_kernel_stats": (None, [abi.pf_p, C.POINTER(C.c_double)]),
    "amclb200_measure_gather_rate": (C.c_double, [C.c_uint64, C.c_int]),
    "amclb200_edt_codes_static": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "amclb200_edt_static_max_cells": (C.c_uint64, []),
    "amclb200_map_from_image": (abi.map_p, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double,
                                            C.POINTER(C.c_double), C.c_int, C.c_double, C.c_double, C.c_int,
                                            C.c_void_p]),
    "amclb200_pf_pack_cloud": (C.c_int, [abi.pf_p, C.c_int, C.c_int, C.c_void_p]),
    "amclb200_inflate_costs": (C.c_int, [C.c_void_p] + [C.c_int] * 7 + [C.c_double] * 4 + [C.c_int]),
    "amclb200_inflate_last_ms": (C.c_float, []),
    "amclb200_map_set_mirror": (None, [abi.map_p, C.c_int]),
    "amclb200_map_pull_distances": (None, [abi.map_p]),
}

RNG_PHILOX, RNG_DRAND48 = 0, 1


def build(verbose=False):
    """nvcc-compile the library in-tree for sm_100a (works without a GPU)."""
    out = subprocess.run(["make", "-C", CSRC, "-j8"], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("libamcl_b200 build failed:\n" + out.stdout[-4000:] + out.stderr[-4000:])
    if verbose:
        print(out.stdout[-2000:])
    return LIB_PATH


def load():
    """dlopen the library and attach prototypes.  Raises if it was never built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback for the AMCL hot path)" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    abi.bind(lib, tables=(abi.REFERENCE_SYMBOLS, abi.FACADE_SYMBOLS, EXTENSION_SYMBOLS))
    _lib = lib
    return lib


def declared_symbols():
    """Every C symbol include/amcl_b200.h + amcl_abi.h declare."""
    extra = [
        "pf_copy_set", "copy_set", "pf_vector_zero", "pf_vector_finite", "pf_vector_fprintf",
        "pf_vector_add", "pf_vector_sub", "pf_matrix_zero", "pf_matrix_finite",
        "pf_matrix_fprintf", "eigen_decomposition", "pf_pdf_gaussian_alloc",
        "pf_pdf_gaussian_free", "pf_pdf_gaussian_sample", "pf_kdtree_alloc", "pf_kdtree_free",
        "pf_kdtree_clear", "pf_kdtree_insert", "pf_kdtree_cluster", "pf_kdtree_get_prob",
        "pf_kdtree_get_cluster", "map_load_occ", "amcl_c_uniform_pose",
    ]
    names = list(abi.REFERENCE_SYMBOLS) + list(abi.FACADE_SYMBOLS) + list(EXTENSION_SYMBOLS) + extra
    return sorted(set(names))
