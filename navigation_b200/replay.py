"""Flat-file replay: what AmclNode::runFromBag
(/root/reference/amcl/src/amcl_node.cpp:676-766) does with a rosbag -- feed the
recorded odometry and scans to laserReceived in time order -- for a plain .npz
log, through navigation_b200.driver.AmclDriver.

Log layout (numpy .npz, all float64 unless noted):
    odom   [T, 3]   odometry pose (x, y, yaw) at each scan
    ranges [T, B]   laser ranges; NaN / inf as the driver would deliver them
    truth  [T, 3]   optional ground truth, for scoring only
    scan   [4]      angle_min, angle_increment, range_min, range_max
    occ    [H, W]   int8 occupancy (-1 free, 0 unknown, +1 occupied), optional
    map    [3]      resolution, origin_x, origin_y of the map CENTRE, optional

    python -m navigation_b200.replay run.npz [--max-particles 5000] [--laser-model likelihood_field]
"""
import argparse
import json
import math

import numpy as np

from . import driver


def save_log(path, bag, bearings, range_max, occ=None, scale=0.05, origin=(0.0, 0.0), range_min=0.0):
    """Write what driver.simulate_run returned (or any recorded run) as a flat file."""
    odom = np.array([b[0] for b in bag], dtype=np.float64)
    ranges = np.array([b[1] for b in bag], dtype=np.float64)
    truth = np.array([b[2] for b in bag], dtype=np.float64)
    inc = float(bearings[1] - bearings[0]) if len(bearings) > 1 else 0.0
    out = dict(odom=odom, ranges=ranges, truth=truth,
               scan=np.array([float(bearings[0]), inc, range_min, range_max], dtype=np.float64))
    if occ is not None:
        out["occ"] = np.ascontiguousarray(occ, dtype=np.int8)
        out["map"] = np.array([scale, origin[0], origin[1]], dtype=np.float64)
    np.savez_compressed(path, **out)


def load_log(path):
    z = np.load(path)
    log = {k: z[k] for k in z.files}
    if log["odom"].shape[0] != log["ranges"].shape[0]:
        raise ValueError("odom and ranges disagree on the number of scans")
    return log


def replay(lib, log, occ=None, scale=None, origin=None, **driver_params):
    """Run every scan of `log` through the per-scan loop.  Returns one dict per
    scan that produced (or kept) an estimate: step, pose, cov, n, truth."""
    if occ is None:
        occ = log["occ"]
        scale, origin = float(log["map"][0]), (float(log["map"][1]), float(log["map"][2]))
    d = driver.AmclDriver(lib, occ, scale if scale is not None else 0.05, origin or (0.0, 0.0),
                          **driver_params)
    a0, inc, rmin, rmax = (float(v) for v in log["scan"])
    out = []
    for t in range(log["odom"].shape[0]):
        r = d.laser_received(log["odom"][t], log["ranges"][t], a0, inc, rmin, rmax)
        if r["pose"] is not None:
            out.append(dict(step=t, pose=r["pose"], cov=r["cov"], n=r["n"], updated=r["updated"],
                            truth=tuple(log["truth"][t]) if "truth" in log else None))
    updates = d.updates
    d.close()
    return out, updates


def score(track):
    """Final position / heading error against the logged truth (basic_localization's criterion)."""
    last = track[-1]
    if last["truth"] is None:
        return None
    p, t = last["pose"], last["truth"]
    return math.hypot(p[0] - t[0], p[1] - t[1]), abs(driver.angle_diff(p[2], t[2]))


def main():
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("log")
    ap.add_argument("--min-particles", type=int, default=100)
    ap.add_argument("--max-particles", type=int, default=5000)
    ap.add_argument("--laser-model", default="likelihood_field",
                    choices=sorted(driver.LASER_MODELS))
    ap.add_argument("--laser-max-beams", type=int, default=30)
    ap.add_argument("--init-pose", type=float, nargs=3, default=None)
    args = ap.parse_args()
    from . import library

    lib = library.load()
    log = load_log(args.log)
    init = tuple(args.init_pose) if args.init_pose else tuple(log["odom"][0])
    track, updates = replay(lib, log, min_particles=args.min_particles, max_particles=args.max_particles,
                            laser_model=args.laser_model, laser_max_beams=args.laser_max_beams,
                            init_pose=init)
    res = dict(scans=int(log["odom"].shape[0]), updates=updates, final_pose=list(track[-1]["pose"]),
               final_n=track[-1]["n"])
    sc = score(track)
    if sc:
        res["final_error_m"], res["final_error_rad"] = sc
    print(json.dumps(res))


if __name__ == "__main__":
    main()
