"""ctypes view of the reference's C ABI for the AMCL hot path.

The structures mirror include/amcl_b200/amcl_abi.h (which in turn restates
/root/reference/amcl/include/amcl/{pf/pf.h:59-143, pf/pf_kdtree.h:37-77,
map/map.h:46-79}).  ``bind(lib)`` attaches prototypes for every reference
entry point (pf.h:147-203, map.h:88-111) and for the plain-C facade over the
C++ sensor classes (navigation_b200/csrc/class_capi.cpp).

The same binder is used for this repo's CUDA library and -- in tests only --
for the compiled reference (oracle/_ref/libamcl_ref.so), because both export
the same symbols.  Nothing here computes anything.
"""
import ctypes as C

c_double_p = C.POINTER(C.c_double)


class pf_vector_t(C.Structure):
    _fields_ = [("v", C.c_double * 3)]

    def tolist(self):
        return [self.v[0], self.v[1], self.v[2]]


class pf_matrix_t(C.Structure):
    _fields_ = [("m", (C.c_double * 3) * 3)]

    def tolist(self):
        return [[self.m[i][j] for j in range(3)] for i in range(3)]


class pf_sample_t(C.Structure):
    _fields_ = [("pose", pf_vector_t), ("weight", C.c_double)]


class pf_cluster_t(C.Structure):
    _fields_ = [
        ("count", C.c_int),
        ("weight", C.c_double),
        ("mean", pf_vector_t),
        ("cov", pf_matrix_t),
        ("m", C.c_double * 4),
        ("c", (C.c_double * 2) * 2),
    ]


class pf_kdtree_t(C.Structure):
    _fields_ = [
        ("size", C.c_double * 3),
        ("root", C.c_void_p),
        ("node_count", C.c_int),
        ("node_max_count", C.c_int),
        ("nodes", C.c_void_p),
        ("leaf_count", C.c_int),
    ]


class pf_sample_set_t(C.Structure):
    _fields_ = [
        ("sample_count", C.c_int),
        ("samples", C.POINTER(pf_sample_t)),
        ("kdtree", C.POINTER(pf_kdtree_t)),
        ("cluster_count", C.c_int),
        ("cluster_max_count", C.c_int),
        ("clusters", C.POINTER(pf_cluster_t)),
        ("mean", pf_vector_t),
        ("cov", pf_matrix_t),
        ("converged", C.c_int),
        ("n_effective", C.c_double),
    ]


pf_init_model_fn_t = C.CFUNCTYPE(pf_vector_t, C.c_void_p)


class pf_t(C.Structure):
    _fields_ = [
        ("min_samples", C.c_int),
        ("max_samples", C.c_int),
        ("pop_err", C.c_double),
        ("pop_z", C.c_double),
        ("limit_cache", C.c_void_p),
        ("current_set", C.c_int),
        ("sets", pf_sample_set_t * 2),
        ("w_slow", C.c_double),
        ("w_fast", C.c_double),
        ("alpha_slow", C.c_double),
        ("alpha_fast", C.c_double),
        ("random_pose_fn", C.c_void_p),
        ("random_pose_data", C.c_void_p),
        ("dist_threshold", C.c_double),
        ("converged", C.c_int),
        ("selective_resampling", C.c_int),
    ]


class map_cell_t(C.Structure):
    _fields_ = [("occ_state", C.c_int), ("occ_dist", C.c_double)]


class map_t(C.Structure):
    _fields_ = [
        ("origin_x", C.c_double),
        ("origin_y", C.c_double),
        ("scale", C.c_double),
        ("size_x", C.c_int),
        ("size_y", C.c_int),
        ("cells", C.POINTER(map_cell_t)),
        ("max_occ_dist", C.c_double),
    ]


assert C.sizeof(pf_sample_t) == 32
assert C.sizeof(pf_cluster_t) == 176
assert C.sizeof(pf_sample_set_t) == 152
assert C.sizeof(pf_t) == 408
assert C.sizeof(map_cell_t) == 16
assert C.sizeof(map_t) == 48

pf_p = C.POINTER(pf_t)
set_p = C.POINTER(pf_sample_set_t)
map_p = C.POINTER(map_t)

# name -> (restype, argtypes).  Reference entry points first.
REFERENCE_SYMBOLS = {
    # pf.h:147-203
    "pf_alloc": (pf_p, [C.c_int, C.c_int, C.c_double, C.c_double, C.c_void_p, C.c_void_p]),
    "pf_free": (None, [pf_p]),
    "pf_init": (None, [pf_p, pf_vector_t, pf_matrix_t]),
    "pf_init_model": (None, [pf_p, C.c_void_p, C.c_void_p]),
    "pf_update_action": (None, [pf_p, C.c_void_p, C.c_void_p]),
    "pf_update_sensor": (None, [pf_p, C.c_void_p, C.c_void_p]),
    "pf_update_resample": (None, [pf_p]),
    "pf_set_selective_resampling": (None, [pf_p, C.c_int]),
    "pf_get_cep_stats": (None, [pf_p, C.POINTER(pf_vector_t), c_double_p]),
    "pf_get_cluster_stats": (C.c_int, [pf_p, C.c_int, c_double_p, C.POINTER(pf_vector_t), C.POINTER(pf_matrix_t)]),
    "pf_cluster_stats": (None, [pf_p, set_p]),
    "pf_update_converged": (C.c_int, [pf_p]),
    "pf_init_converged": (None, [pf_p]),
    # pf_pdf.h:64-79, pf_vector.h:52-88
    "pf_ran_gaussian": (C.c_double, [C.c_double]),
    "pf_vector_coord_add": (pf_vector_t, [pf_vector_t, pf_vector_t]),
    "pf_vector_coord_sub": (pf_vector_t, [pf_vector_t, pf_vector_t]),
    "pf_matrix_unitary": (None, [C.POINTER(pf_matrix_t), C.POINTER(pf_matrix_t), pf_matrix_t]),
    # map.h:88-111
    "map_alloc": (map_p, []),
    "map_free": (None, [map_p]),
    "map_get_cell": (C.POINTER(map_cell_t), [map_p, C.c_double, C.c_double, C.c_double]),
    "map_update_cspace": (None, [map_p, C.c_double]),
    "map_calc_range": (C.c_double, [map_p, C.c_double, C.c_double, C.c_double, C.c_double]),
}

# Plain-C facade over the C++ classes (class_capi.cpp; same source in both libs).
FACADE_SYMBOLS = {
    "amcl_c_map_from_occupancy": (map_p, [C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double]),
    "amcl_c_map_get_occ_dist": (None, [map_p, C.c_void_p]),
    "amcl_c_map_set_occ_dist": (None, [map_p, C.c_void_p, C.c_double]),
    "amcl_c_freespace_build": (C.c_int, [map_p]),
    "amcl_c_uniform_pose_fn": (C.c_void_p, []),
    "amcl_c_odom_new": (C.c_void_p, []),
    "amcl_c_odom_delete": (None, [C.c_void_p]),
    "amcl_c_odom_set_model": (None, [C.c_void_p, C.c_int] + [C.c_double] * 5),
    "amcl_c_odom_update_action": (C.c_int, [C.c_void_p, pf_p, c_double_p, c_double_p]),
    "amcl_c_laser_new": (C.c_void_p, [C.c_int, map_p]),
    "amcl_c_laser_delete": (None, [C.c_void_p]),
    "amcl_c_laser_clone": (C.c_void_p, [C.c_void_p]),
    "amcl_c_laser_set_pose": (None, [C.c_void_p] + [C.c_double] * 3),
    "amcl_c_laser_set_model_beam": (None, [C.c_void_p] + [C.c_double] * 7),
    "amcl_c_laser_set_model_lf": (None, [C.c_void_p] + [C.c_double] * 4),
    "amcl_c_laser_set_model_lfp": (None, [C.c_void_p] + [C.c_double] * 4 + [C.c_int] + [C.c_double] * 3),
    "amcl_c_laser_update_sensor": (C.c_int, [C.c_void_p, pf_p, C.c_void_p, C.c_int, C.c_double]),
    "amcl_c_pf_current_set": (set_p, [pf_p]),
    "amcl_c_sizeof_pf": (C.c_int, []),
    "amcl_c_sizeof_laser": (C.c_int, []),
    "amcl_c_sizeof_odom": (C.c_int, []),
}

ODOM_MODEL_DIFF, ODOM_MODEL_OMNI, ODOM_MODEL_DIFF_CORRECTED, ODOM_MODEL_OMNI_CORRECTED = range(4)


def bind(lib, tables=(REFERENCE_SYMBOLS, FACADE_SYMBOLS)):
    """Attach restype/argtypes; raises AttributeError on a missing symbol."""
    for table in tables:
        for name, (res, args) in table.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
    return lib
