"""ROS-free equivalent of AmclNode's per-scan loop.

`AmclDriver.laser_received` restates what `AmclNode::laserReceived` does with
the filter (/root/reference/amcl/src/amcl_node.cpp:1128-1529) minus ROS, tf and
message types: odometry delta and update thresholds (:1195-1213), action update
(:1233-1247), scan -> AMCLLaserData conversion with the min/max range rules
(:1296-1327), sensor update (:1329), resampling every resample_interval-th
update (:1336-1340) and the arg-max-weight cluster as the pose estimate
(:1366-1435).  Map set-up follows handleMapMessage (:877-964).

It drives any library that exports the reference API -- this repo's CUDA
library, or (in tests) the compiled reference -- through navigation_b200.abi,
so whole trajectories can be replayed on both and compared the way the
reference's own basic_localization tests do (final pose within a tolerance).
"""
import ctypes as C
import math

import numpy as np

from . import abi

LASER_MODELS = {"beam": 0, "likelihood_field": 1, "likelihood_field_prob": 2}


def _normalize(z):
    return math.atan2(math.sin(z), math.cos(z))


def angle_diff(a, b):
    """amcl_node.cpp:107-122."""
    a = _normalize(a)
    b = _normalize(b)
    d1 = a - b
    d2 = 2 * math.pi - abs(d1)
    if d1 > 0:
        d2 *= -1.0
    return d1 if abs(d1) < abs(d2) else d2


class AmclDriver:
    # defaults are the node's (amcl_node.cpp:365-456)
    def __init__(self, lib, occ, scale=0.05, origin=(0.0, 0.0), *, min_particles=100,
                 max_particles=5000, kld_err=0.01, kld_z=0.99, alpha_slow=0.001, alpha_fast=0.1,
                 odom_model=abi.ODOM_MODEL_DIFF, odom_alphas=(0.2, 0.2, 0.2, 0.2, 0.2),
                 laser_model="likelihood_field", laser_max_beams=30, z_hit=0.95, z_short=0.1,
                 z_max=0.05, z_rand=0.05, sigma_hit=0.2, lambda_short=0.1,
                 laser_likelihood_max_dist=2.0, laser_min_range=-1.0, laser_max_range=-1.0,
                 laser_pose=(0.0, 0.0, 0.0), d_thresh=0.2, a_thresh=math.pi / 6.0,
                 resample_interval=2, selective_resampling=False, do_beamskip=False,
                 beam_skip_distance=0.5, beam_skip_threshold=0.3, beam_skip_error_threshold=0.9,
                 init_pose=(0.0, 0.0, 0.0), init_cov=(0.5 * 0.5, 0.5 * 0.5, (math.pi / 12) ** 2)):
        self.lib = lib
        occ = np.ascontiguousarray(occ, dtype=np.int8)
        sy, sx = occ.shape
        # handleMapMessage: convertMap, free-space index, pf_alloc, struct pokes, pf_init
        self.map = lib.amcl_c_map_from_occupancy(occ.ctypes.data, sx, sy, scale, origin[0], origin[1])
        lib.amcl_c_freespace_build(self.map)
        self.pf = lib.pf_alloc(min_particles, max_particles, alpha_slow, alpha_fast,
                               lib.amcl_c_uniform_pose_fn(), C.cast(self.map, C.c_void_p))
        lib.pf_set_selective_resampling(self.pf, 1 if selective_resampling else 0)
        self.pf.contents.pop_err = kld_err
        self.pf.contents.pop_z = kld_z
        self.set_initial_pose(init_pose, init_cov)
        self.odom = lib.amcl_c_odom_new()
        lib.amcl_c_odom_set_model(self.odom, odom_model, *odom_alphas)
        self.laser = lib.amcl_c_laser_new(laser_max_beams, self.map)
        lib.amcl_c_laser_set_pose(self.laser, *laser_pose)
        kind = LASER_MODELS[laser_model]
        if kind == 0:
            lib.amcl_c_laser_set_model_beam(self.laser, z_hit, z_short, z_max, z_rand, sigma_hit,
                                            lambda_short, 0.0)
        elif kind == 2:
            lib.amcl_c_laser_set_model_lfp(self.laser, z_hit, z_rand, sigma_hit,
                                           laser_likelihood_max_dist, int(do_beamskip),
                                           beam_skip_distance, beam_skip_threshold,
                                           beam_skip_error_threshold)
        else:
            lib.amcl_c_laser_set_model_lf(self.laser, z_hit, z_rand, sigma_hit,
                                          laser_likelihood_max_dist)
        self.d_thresh = d_thresh
        self.a_thresh = a_thresh
        self.resample_interval = resample_interval
        self.laser_min_range = laser_min_range
        self.laser_max_range = laser_max_range
        self.pf_odom_pose = None          # pf_init_ == False
        self.resample_count = 0
        self.force_update = False
        self.pose = None
        self.cov = None
        self.updates = 0

    # -- services ---------------------------------------------------------------
    def set_initial_pose(self, pose, cov_diag):
        """applyInitialPose (amcl_node.cpp:1621-1632)."""
        m = abi.pf_vector_t()
        m.v[:] = list(pose)
        cv = abi.pf_matrix_t()
        for k in range(3):
            cv.m[k][k] = cov_diag[k]
        self.lib.pf_init(self.pf, m, cv)
        self.pf_odom_pose = None

    def global_localization(self):
        """globalLocalizationCallback (amcl_node.cpp:1088-1102)."""
        self.lib.pf_init_model(self.pf, self.lib.amcl_c_uniform_pose_fn(), C.cast(self.map, C.c_void_p))
        self.pf_odom_pose = None

    def request_nomotion_update(self):
        self.force_update = True

    # -- the per-scan loop ----------------------------------------------------------
    def laser_received(self, odom_pose, ranges, angle_min, angle_increment, range_min, range_max):
        """One scan.  Returns dict(updated, resampled, pose, cov, n)."""
        lib, pf = self.lib, self.pf
        pose = [float(odom_pose[0]), float(odom_pose[1]), float(odom_pose[2])]
        update = False
        delta = [0.0, 0.0, 0.0]
        force_publication = False
        if self.pf_odom_pose is not None:
            delta = [pose[0] - self.pf_odom_pose[0], pose[1] - self.pf_odom_pose[1],
                     angle_diff(pose[2], self.pf_odom_pose[2])]
            update = (abs(delta[0]) > self.d_thresh or abs(delta[1]) > self.d_thresh or
                      abs(delta[2]) > self.a_thresh or self.force_update)
            self.force_update = False
        if self.pf_odom_pose is None:
            self.pf_odom_pose = pose
            update = True                       # "should update sensor data"
            force_publication = True
            self.resample_count = 0
        elif update:
            p = (C.c_double * 3)(*pose)
            d = (C.c_double * 3)(*delta)
            lib.amcl_c_odom_update_action(self.odom, pf, p, d)
        resampled = False
        if update:
            rmax = min(range_max, self.laser_max_range) if self.laser_max_range > 0.0 else range_max
            rmin = max(range_min, self.laser_min_range) if self.laser_min_range > 0.0 else range_min
            r = np.asarray(ranges, dtype=np.float64)
            out = np.empty((r.size, 2), dtype=np.float64)
            dbl_max = np.finfo(np.float64).max
            # short readings -> max range; beyond max range -> "infinitely far" (skipped)
            out[:, 0] = np.where(r <= rmin, rmax, np.where(r > rmax, dbl_max, r))
            out[:, 1] = angle_min + np.arange(r.size) * angle_increment
            lib.amcl_c_laser_update_sensor(self.laser, pf, out.ctypes.data, r.size, rmax)
            self.pf_odom_pose = pose
            self.updates += 1
            self.resample_count += 1
            if self.resample_count % self.resample_interval == 0:
                lib.pf_update_resample(pf)
                resampled = True
        if resampled or force_publication:
            self._read_hypothesis()
        cur = pf.contents.sets[pf.contents.current_set]
        return dict(updated=update, resampled=resampled, pose=self.pose, cov=self.cov,
                    n=cur.sample_count)

    def _read_hypothesis(self):
        """Arg-max-weight cluster (amcl_node.cpp:1369-1394); covariance of the whole set (:1421-1435)."""
        lib, pf = self.lib, self.pf
        best_w, best = 0.0, None
        w = C.c_double()
        m = abi.pf_vector_t()
        cv = abi.pf_matrix_t()
        k = 0
        while lib.pf_get_cluster_stats(pf, k, C.byref(w), C.byref(m), C.byref(cv)):
            if w.value > best_w:
                best_w, best = w.value, (m.v[0], m.v[1], m.v[2])
            k += 1
        if best is not None:
            cur = pf.contents.sets[pf.contents.current_set]
            self.pose = best
            self.cov = (cur.cov.m[0][0], cur.cov.m[1][1], cur.cov.m[2][2])

    def close(self):
        self.lib.amcl_c_laser_delete(self.laser)
        self.lib.amcl_c_odom_delete(self.odom)
        self.lib.pf_free(self.pf)
        self.lib.map_free(self.map)


def simulate_run(lib_for_rays, occ, scale, path, n_beams=60, fov=1.5 * math.pi, range_max=12.0,
                 odom_drift=(1.03, 0.02), noise_sigma=0.01, seed=0):
    """A synthetic "bag": ground-truth poses along `path`, odometry that drifts
    (scaled translation + heading bias per metre), and scans ray-cast from the
    ground truth with Gaussian range noise.  Rays are cast by the given
    library's map_calc_range.  Returns a list of (odom_pose, ranges, truth)."""
    sy, sx = occ.shape
    m = lib_for_rays.amcl_c_map_from_occupancy(np.ascontiguousarray(occ, np.int8).ctypes.data, sx, sy,
                                               scale, 0.0, 0.0)
    rng = np.random.default_rng(seed)
    bear = -0.5 * fov + np.arange(n_beams) * (fov / (n_beams - 1))
    out = []
    odom = np.array(path[0], dtype=np.float64)
    prev = np.array(path[0], dtype=np.float64)
    for truth in path:
        truth = np.array(truth, dtype=np.float64)
        step = truth - prev
        dist = math.hypot(step[0], step[1])
        heading_err = odom_drift[1] * dist
        c, s = math.cos(odom[2] - prev[2]), math.sin(odom[2] - prev[2])
        odom = odom + np.array([odom_drift[0] * (c * step[0] - s * step[1]),
                                odom_drift[0] * (s * step[0] + c * step[1]),
                                angle_diff(truth[2], prev[2]) + heading_err])
        prev = truth
        r = np.array([lib_for_rays.map_calc_range(m, truth[0], truth[1], truth[2] + b, range_max)
                      for b in bear])
        r = np.minimum(r + rng.normal(0.0, noise_sigma, r.size), range_max)
        out.append((odom.copy(), r, truth.copy()))
    lib_for_rays.map_free(m)
    return out, bear
