"""ctypes harness over the *reference's own API* (pf_alloc, AMCLLaser, AMCLOdom
... through the plain-C facade of include/amcl_b200.h Part 2).  It drives
whichever library it is given -- this repo's CUDA library, or the compiled
reference (oracle/_ref) in tests and in bench.py's CPU legs -- with the same
calls, which is what lets parity tests and the two bench arms read alike.
"""
import ctypes as C

import numpy as np

from navigation_b200 import abi, synth

LASER_DEFAULTS = dict(z_hit=0.95, z_short=0.1, z_max=0.05, z_rand=0.05, sigma_hit=0.2,
                      lambda_short=0.1, max_occ_dist=2.0)


class Side:
    """One library + one map + one filter, addressed through the reference API."""

    def __init__(self, lib, occ, scale=0.05, origin=(0.0, 0.0)):
        self.lib = lib
        self.is_gpu = hasattr(lib, "amclb200_version")
        self.occ = np.ascontiguousarray(occ, dtype=np.int8)
        sy, sx = self.occ.shape
        self.map = lib.amcl_c_map_from_occupancy(self.occ.ctypes.data, sx, sy, scale,
                                                 origin[0], origin[1])
        self.scale = scale
        self.origin = origin
        self.pf = None
        self.laser = None
        self.odom = None
        self._keep = []

    # -- RNG -----------------------------------------------------------------
    def seed(self, s):
        """Seed the drand48 stream this side draws from."""
        if self.is_gpu:
            libc = C.CDLL(None)
            libc.srand48.argtypes = [C.c_long]
            libc.srand48(s)
        else:
            self.lib.tap_seed(s)

    # -- map -------------------------------------------------------------------
    def update_cspace(self, max_occ_dist=2.0):
        self.lib.map_update_cspace(self.map, max_occ_dist)

    def occ_dist(self):
        sy, sx = self.occ.shape
        out = np.empty((sy, sx), dtype=np.float64)
        self.lib.amcl_c_map_get_occ_dist(self.map, out.ctypes.data)
        return out

    def set_occ_dist(self, dist, max_occ_dist):
        d = np.ascontiguousarray(dist, dtype=np.float64)
        self.lib.amcl_c_map_set_occ_dist(self.map, d.ctypes.data, max_occ_dist)
        if self.is_gpu:
            rc = self.lib.amclb200_map_push_distances(self.map)
            assert rc == 0, "occ_dist values the device codes cannot express"

    # -- filter ------------------------------------------------------------------
    def alloc(self, min_samples, max_samples, alpha_slow=0.0, alpha_fast=0.0,
              pop_err=0.01, pop_z=0.99):
        self.lib.amcl_c_freespace_build(self.map)
        self.pf = self.lib.pf_alloc(min_samples, max_samples, alpha_slow, alpha_fast,
                                    self.lib.amcl_c_uniform_pose_fn(),
                                    C.cast(self.map, C.c_void_p))
        # the node pokes these after pf_alloc (amcl_node.cpp:915-916)
        self.pf.contents.pop_err = pop_err
        self.pf.contents.pop_z = pop_z
        return self.pf

    @property
    def cur(self):
        return self.pf.contents.sets[self.pf.contents.current_set]

    def samples(self):
        """Copy of the current set as [n, 4] (x, y, theta, w)."""
        if self.is_gpu:
            self.lib.amclb200_pf_pull_samples(self.pf)
        s = self.cur
        n = s.sample_count
        arr = np.ctypeslib.as_array(C.cast(s.samples, C.POINTER(C.c_double)), shape=(n, 4))
        return arr.copy()

    def set_samples(self, arr):
        arr = np.ascontiguousarray(arr, dtype=np.float64)
        s = self.pf.contents.sets[self.pf.contents.current_set]
        n = arr.shape[0]
        assert n <= self.pf.contents.max_samples
        self.pf.contents.sets[self.pf.contents.current_set].sample_count = n
        C.memmove(s.samples, arr.ctypes.data, arr.nbytes)
        if self.is_gpu:
            self.lib.amclb200_pf_push_samples(self.pf)

    def init_gaussian(self, mean, cov_diag):
        m = abi.pf_vector_t()
        m.v[:] = list(mean)
        cv = abi.pf_matrix_t()
        for k in range(3):
            cv.m[k][k] = cov_diag[k]
        self.lib.pf_init(self.pf, m, cv)

    def init_uniform(self):
        self.lib.pf_init_model(self.pf, self.lib.amcl_c_uniform_pose_fn(),
                               C.cast(self.map, C.c_void_p))

    # -- sensors -------------------------------------------------------------------
    def make_laser(self, max_beams, model="lf", laser_pose=(0.0, 0.0, 0.0), beamskip=None, **kw):
        p = dict(LASER_DEFAULTS)
        p.update(kw)
        self.laser = self.lib.amcl_c_laser_new(max_beams, self.map)
        self.lib.amcl_c_laser_set_pose(self.laser, *laser_pose)
        if model == "lf":
            self.lib.amcl_c_laser_set_model_lf(self.laser, p["z_hit"], p["z_rand"],
                                               p["sigma_hit"], p["max_occ_dist"])
        elif model == "lfp":
            bs = beamskip or (0, 0.5, 0.3, 0.9)
            self.lib.amcl_c_laser_set_model_lfp(self.laser, p["z_hit"], p["z_rand"],
                                                p["sigma_hit"], p["max_occ_dist"],
                                                int(bs[0]), bs[1], bs[2], bs[3])
        else:
            self.lib.amcl_c_laser_set_model_beam(self.laser, p["z_hit"], p["z_short"],
                                                 p["z_max"], p["z_rand"], p["sigma_hit"],
                                                 p["lambda_short"], 0.0)
        return self.laser

    def update_sensor(self, ranges_rb, range_max):
        rb = np.ascontiguousarray(ranges_rb, dtype=np.float64)
        return self.lib.amcl_c_laser_update_sensor(self.laser, self.pf, rb.ctypes.data,
                                                   rb.shape[0], range_max)

    def make_odom(self, model=abi.ODOM_MODEL_DIFF, alphas=(0.2, 0.2, 0.2, 0.2, 0.2)):
        self.odom = self.lib.amcl_c_odom_new()
        self.lib.amcl_c_odom_set_model(self.odom, model, *alphas)
        return self.odom

    def update_action(self, pose, delta):
        p = (C.c_double * 3)(*pose)
        d = (C.c_double * 3)(*delta)
        return self.lib.amcl_c_odom_update_action(self.odom, self.pf, p, d)

    def resample(self):
        self.lib.pf_update_resample(self.pf)

    def cluster_table(self):
        """[(weight, mean[3], cov[3][3])] via pf_get_cluster_stats (pf.c:744-761)."""
        out = []
        k = 0
        while True:
            w = C.c_double()
            m = abi.pf_vector_t()
            cv = abi.pf_matrix_t()
            if not self.lib.pf_get_cluster_stats(self.pf, k, C.byref(w), C.byref(m), C.byref(cv)):
                break
            out.append((w.value, np.array(m.tolist()), np.array(cv.tolist())))
            k += 1
        return out

    def close(self):
        if self.laser:
            self.lib.amcl_c_laser_delete(self.laser)
        if self.odom:
            self.lib.amcl_c_odom_delete(self.odom)
        if self.pf:
            self.lib.pf_free(self.pf)
        self.lib.map_free(self.map)
        self.pf = self.laser = self.odom = None


def make_scan(lib_side, pose, n_beams, range_max=30.0, perturb=None):
    """Scan(B) of SURVEY.md 8(d): ranges from map_calc_range at the true pose;
    readings at/over range_max are reported as range_max."""
    bear = synth.scan_bearings(n_beams)
    rng = np.array([lib_side.lib.map_calc_range(lib_side.map, pose[0], pose[1],
                                                pose[2] + b, range_max) for b in bear])
    rng = np.minimum(rng, range_max)
    if perturb is not None:
        rng = perturb(rng)
    return np.stack([rng, bear], axis=1).copy()


def match_clusters(a, b):
    """Pair clusters of two tables (different label numbering) by nearest
    (mean x, mean y, mean heading, weight); asserts the pairing is a bijection."""
    assert len(a) == len(b)
    pairs = []
    for ia, (wa, ma, _) in enumerate(a):
        d = [float(np.sum((ma[:2] - mb[:2]) ** 2) + (wa - wb) ** 2 +
                   abs(np.angle(np.exp(1j * (ma[2] - mb[2])))) ** 2) for (wb, mb, _) in b]
        pairs.append((ia, int(np.argmin(d))))
    assert len(set(p[1] for p in pairs)) == len(b), "cluster tables do not pair up one-to-one"
    return pairs
