#!/usr/bin/env python
"""bench.py -- the AMCL hot path on B200, next to the reference's CPU path.

Workload (BASELINE.json configs[4], "K5"): one FULL AMCL UPDATE per step --
diff-drive odometry motion model + likelihood_field laser model (1080 beams)
+ KLD-adaptive resampling + cluster statistics -- on a 4000x4000 map at
0.05 m, global-localisation cloud, every step restarted from the same cloud.
STRONG scaling: --particles (default 16 Mi) is the whole job, sharded over the
ranks; N = 1 runs all of it on one GPU.

metric  = particle-beam evaluations per second over the whole update
          (value), with ms_per_step the full-update latency.
value   = device-resident: cloud, map and scan table already in HBM.
e2e     = the same update driven through the reference-facing API with HOST
          buffers: the scan and odometry come from host memory every step and
          the pose estimate (cluster table, set mean/cov) is read back.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

MAP_SIZE = 4000
SCALE = 0.05
N_BEAMS = 1080
RANGE_MAX = 30.0
MAX_OCC_DIST = 2.0
LASER = dict(z_hit=0.95, z_rand=0.05, sigma_hit=0.2)
ALPHAS = (0.2, 0.2, 0.2, 0.2, 0.2)
ODOM_POSE = (0.3, 0.0, 0.05)
ODOM_DELTA = (0.3, 0.0, 0.05)
METRIC = "particle_beam_evals_per_s"
UNIT = "evals/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--particles", type=int, default=16 * 1024 * 1024)
    ap.add_argument("--map-size", type=int, default=MAP_SIZE)
    ap.add_argument("--beams", type=int, default=N_BEAMS)
    ap.add_argument("--cpu-particles", type=int, default=4000,
                    help="particles per CPU worker for the reference legs")
    ap.add_argument("--cpu-workers", type=int, default=0, help="0 = one per host core (max 32)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


# ------------------------------------------------------------------ workload ----

def build_world(size, beams):
    from navigation_b200 import synth

    occ = synth.rooms_map(size, seed=1)
    pose = synth.room_centre_pose(size, SCALE)
    bear = synth.scan_bearings(beams)
    return occ, pose, bear


def scan_from(lib_side, pose, bear, gpu):
    """Scan(B): ranges from map_calc_range at the true pose."""
    if gpu:
        rays = np.stack([np.full_like(bear, pose[0]), np.full_like(bear, pose[1]), pose[2] + bear], 1).copy()
        out = np.empty(bear.size)
        lib_side.lib.amclb200_map_calc_ranges(lib_side.map, rays.ctypes.data, bear.size, RANGE_MAX,
                                              out.ctypes.data)
    else:
        out = np.array([lib_side.lib.map_calc_range(lib_side.map, pose[0], pose[1], pose[2] + b,
                                                    RANGE_MAX) for b in bear])
    rng = np.minimum(out, RANGE_MAX)
    return np.stack([rng, bear], 1).copy()


# ------------------------------------------------------- reference (CPU) legs ----

def _cpu_worker(args):
    """One single-threaded replica of the reference on its own particle shard."""
    wid, size, beams, n, steps, warmup = args
    from navigation_b200 import abi, synth
    from oracle import oracle
    from tests.scenario import Side

    ref = oracle.ref()
    kind = "reference"
    if ref is None:
        return None
    occ, pose, bear = build_world(size, beams)
    side = Side(ref, occ, scale=SCALE)
    t0 = time.perf_counter()
    side.alloc(100, n, pop_err=0.01, pop_z=0.99)
    side.make_laser(beams, model="lf", **dict(LASER, max_occ_dist=MAX_OCC_DIST))  # runs the wavefront
    t_cspace = time.perf_counter() - t0
    side.make_odom(model=abi.ODOM_MODEL_DIFF, alphas=ALPHAS)
    scan = scan_from(side, pose, bear, gpu=False)
    used = int(np.sum(scan[:, 0] < RANGE_MAX))
    cloud = synth.uniform_cloud(occ, SCALE, n, 100 + wid)
    side.seed(1 + wid)
    times = []
    for it in range(warmup + steps):
        side.set_samples(cloud)
        t0 = time.perf_counter()
        side.update_action(ODOM_POSE, ODOM_DELTA)
        side.update_sensor(scan, RANGE_MAX)
        side.resample()
        tab = side.cluster_table()
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt)
    side.close()
    return dict(kind=kind, seconds=float(np.sum(times)), steps=steps, n=n, used_beams=used,
                cspace_s=t_cspace, clusters=len(tab))


def run_cpu_reference(size, beams, n_per_worker, steps, warmup, workers):
    """`workers` independent single-thread replicas (the reference has no threads
    of its own); returns aggregate evals/s and the per-step wall time."""
    import multiprocessing as mp

    ctx = mp.get_context("fork")
    with ctx.Pool(workers) as pool:
        t0 = time.perf_counter()
        res = pool.map(_cpu_worker, [(w, size, beams, n_per_worker, steps, warmup) for w in range(workers)])
        wall = time.perf_counter() - t0
    if any(r is None for r in res):
        return None
    evals = sum(r["n"] * r["used_beams"] * r["steps"] for r in res)
    slowest = max(r["seconds"] for r in res)
    return dict(value=evals / slowest, ms_per_step=1e3 * slowest / steps, cores=workers,
                kind="reference", wall_s=wall, cspace_s=max(r["cspace_s"] for r in res),
                sample="%d workers x %d particles x %d used beams, %d steps each, %dx%d map"
                       % (workers, n_per_worker, res[0]["used_beams"], steps, size, size))


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    workers = args.cpu_workers if args.cpu_workers > 0 else min(max(1, (os.cpu_count() or 1)), 32)
    r = run_cpu_reference(args.map_size, args.beams, args.cpu_particles, max(1, min(args.steps, 3)),
                          min(args.warmup, 1), workers)
    if r is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libamcl_ref.so not built"}))
        return
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, note="CPU reference, bounded sample"),
        "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
                         "sample": r["sample"]},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def workload_config(args, note=None):
    cfg = {
        "workload": "K5 full AMCL update: diff odom + likelihood_field + KLD resample + cluster stats",
        "particles_total": args.particles, "beams": args.beams,
        "map": "%dx%d rooms @ %.2f m, max_occ_dist %.1f" % (args.map_size, args.map_size, SCALE, MAX_OCC_DIST),
        "cloud": "global (uniform over free space), restored from a device copy every step",
        "rng": "philox4x32-10", "l2": "particle arrays (2 x 32 B x N) exceed L2; map (2 B/cell) is meant to be L2-resident",
        "parallelism": "particles sharded over %d GPU(s)" % args.gpus,
    }
    if note:
        cfg["note"] = note
    return cfg


# --------------------------------------------------------------------- clocks ----

class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.FIELDS,
                                       "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        self.p.wait()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.f.read().splitlines():
            parts = [x.strip() for x in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.f.name)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": float(np.max(mx)) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------ B200 arm ----

def b200_arm(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    # stdout carries exactly one JSON line: anything libraries print while we work
    # (NCCL's version banner, the reference's fprintf's) is sent to stderr instead
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    from navigation_b200 import abi, library, synth
    from tests.scenario import Side

    lib = library.load()
    lib.amclb200_set_rng(library.RNG_PHILOX, 0x5EED)       # one seed: ranks share one filter
    occ, pose, bear = build_world(args.map_size, args.beams)
    side = Side(lib, occ, scale=SCALE)
    n_total = args.particles
    # every rank allocates the GLOBAL filter and owns a contiguous slice of it
    side.alloc(100, n_total, pop_err=0.01, pop_z=0.99)
    coll = None
    if world > 1:
        from navigation_b200 import sharded
        coll = sharded.Collectives(device=torch.device("cuda", local))
        coll.attach(lib, side.pf)
    rng2 = (C.c_int * 2)()
    lib.amclb200_pf_shard_range(side.pf, n_total, rng2)
    own_lo, own_hi = rng2[0], rng2[1]
    n_local = own_hi - own_lo
    lib.amclb200_set_mirror(side.pf, 0)      # deferred host mirror
    t0 = time.perf_counter()
    side.make_laser(args.beams, model="lf", **dict(LASER, max_occ_dist=MAX_OCC_DIST))  # GPU EDT
    t_cspace = time.perf_counter() - t0
    side.make_odom(model=abi.ODOM_MODEL_DIFF, alphas=ALPHAS)
    scan = scan_from(side, pose, bear, gpu=True)
    used_beams = int(np.sum(scan[:, 0] < RANGE_MAX))

    # the same global cloud on every rank (hash of the global index); keep the own slice
    cloud = synth.uniform_cloud(occ, SCALE, n_total, 100)[own_lo:own_hi]
    pristine = torch.from_numpy(np.ascontiguousarray(cloud)).cuda()
    del cloud
    scan_host = torch.from_numpy(scan).pin_memory()
    odom_pose = (C.c_double * 3)(*ODOM_POSE)
    odom_delta = (C.c_double * 3)(*ODOM_DELTA)

    def step():
        lib.amclb200_pf_load_device_shard(side.pf, pristine.data_ptr(), n_total)
        lib.amcl_c_odom_update_action(side.odom, side.pf, odom_pose, odom_delta)
        lib.amcl_c_laser_update_sensor(side.laser, side.pf, scan_host.data_ptr(), args.beams, RANGE_MAX)
        lib.pf_update_resample(side.pf)
        # what the node does next (amcl_node.cpp:1369-1394): arg-max weight cluster
        best = lib.amclb200_pf_best_cluster(side.pf, C.byref(bw), C.byref(bm), C.byref(bc))
        return (best, bw.value, bm.v[0], bm.v[1], bm.v[2]), side.cur.cluster_count

    bw = C.c_double()
    bm = abi.pf_vector_t()
    bc = abi.pf_matrix_t()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    stage_ms = np.zeros(3, dtype=np.float32)
    kstats = (C.c_double * 2)()
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    launches0 = lib.amclb200_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kernel_ms = []
    stages = []
    t_wall0 = time.perf_counter()
    e0.record()
    for _ in range(args.steps):
        best, ncl = step()
        lib.amclb200_pf_sensor_kernel_stats(side.pf, kstats)
        kernel_ms.append(kstats[0])
        lib.amclb200_pf_last_timings(side.pf, stage_ms.ctypes.data_as(C.POINTER(C.c_float)))
        stages.append(stage_ms.copy())
    e1.record()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    launches = lib.amclb200_launch_count() - launches0
    clocks = sampler.stop() if sampler else None
    ms_total = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms_total], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_step = ms_total / args.steps
    evals_step = n_total * used_beams
    value = evals_step / (ms_step * 1e-3)

    # ---- roofline of the dominant kernel (likelihood_field) ------------------
    k_ms = float(np.mean(kernel_ms))
    k_evals = n_local * used_beams
    alg_bytes = 32.0 * k_evals + 40.0 * n_local
    achieved = alg_bytes / (k_ms * 1e-3) / 1e9
    map_bytes = args.map_size * args.map_size * 2
    gather_rate = lib.amclb200_measure_gather_rate(map_bytes, 5)
    peak = gather_rate * 32.0 / 1e9
    peaks_file = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm = json.load(open(peaks_file))["hbm_gbs"] if os.path.exists(peaks_file) else 6650.0
    traffic = None
    l2_sectors = None
    tfile = os.path.join(ROOT, "profiles", "lf_traffic.json")
    if os.path.exists(tfile) and args.beams == N_BEAMS and args.map_size == MAP_SIZE:
        tf = json.load(open(tfile))
        traffic = tf.get(str(n_local))     # measured by ncu for this launch shape
        l2_sectors = tf.get("l2_sectors_%d" % n_local)
    roofline = {
        "kernel": "likelihood_field_kernel<LF>", "bound": "l2_gather",
        "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
        "traffic": traffic,
        "note": ("algorithmic bytes = 32 B per evaluation (one L2 sector per uncorrelated gather) + 40 B per "
                 "particle (SURVEY.md 8d); the map is L2-resident, so HBM is not the binding bound: DRAM "
                 "traffic ('traffic', ncu) is ~64 B per particle.  Neighbouring beams share sectors, which "
                 "is why achieved exceeds the uncorrelated-gather peak; ncu: L1TEX 90 %, issue 61 %, fp64 pipe "
                 "41 %; ablations (profiles/r1_ncu_full_likelihood_field_16M_v2.md) put the arithmetic alone at "
                 "15.2 ms of the 18.3 ms"),
        "evals_per_s": k_evals / (k_ms * 1e-3), "gather_peak_per_s": gather_rate,
        "kernel_ms": k_ms, "share_of_step": k_ms / ms_step,
        "peak_source": "measured live: uncorrelated 2-byte gathers over a %d MB table (one 32-B sector each)" % (map_bytes >> 20),
        "hbm_peak_gbs": hbm, "hbm_peak_source": "MEASURED_PEAKS.json" if os.path.exists(peaks_file) else "fallback",
    }
    # The bound that actually binds (ablations): instruction issue.  One evaluation is 20 SASS
    # instructions per warp of which 7 are fp64 (two dispatch cycles each): 27 issue cycles.
    props = torch.cuda.get_device_properties(local)
    sm_hz = 1e6 * ((clocks or {}).get("sm_mhz") or (clocks or {}).get("sm_max_mhz") or 1965.0)
    issue_peak = props.multi_processor_count * 4 * sm_hz * 32 / 27.0
    if l2_sectors:
        # actual L2 sector traffic of the launch (ncu) against the measured gather peak
        roofline["l2_sectors"] = {"per_launch": l2_sectors, "rate_per_s": l2_sectors / (k_ms * 1e-3),
                                  "frac_of_gather_peak": l2_sectors / (k_ms * 1e-3) / gather_rate}
    roofline["issue"] = {"issue_cycles_per_warp_eval": 27, "peak_evals_per_s": issue_peak,
                         "frac": roofline["evals_per_s"] / issue_peak,
                         "note": "4 schedulers x SMs x SM clock x 32 lanes / 27 cycles; SASS count from cuobjdump"}

    # ---- e2e: host buffers through the reference-facing API ------------------
    # The step() above already IS the reference-facing call sequence with the
    # scan in (pinned) host memory and the cluster table read back on the host;
    # e2e repeats it timed by the wall clock around the API calls.
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    barrier()
    e2e_s = (time.perf_counter() - t0) / args.steps
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    # scan + odometry from the caller, plus the beam table and code->likelihood
    # table the library stages for the kernel each update
    h2d = scan.nbytes + 48 + 16 * used_beams + 8 * (40 * 40 + 3)
    d2h = 80 * max(ncl, 1) + 64
    e2e = {"value": evals_step / e2e_s, "unit": UNIT, "ms_per_step": 1e3 * e2e_s,
           "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
           "note": "deferred host mirror; eager mirror adds a %d MB D2H per API call" % ((32 * n_local) >> 20)}

    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    os.close(saved_stdout)
    if rank != 0:
        return
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args), "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches),
        "roofline": roofline,
        "collectives": (dict(coll.calls, note="counts over warm-up + timed + e2e steps; in-place NCCL via torch.distributed") if coll else None),
        "stage_ms": {"action": float(np.mean([s[0] for s in stages])),
                     "sensor": float(np.mean([s[1] for s in stages])),
                     "resample": float(np.mean([s[2] for s in stages])),
                     "cspace_setup_s": t_cspace},
        "result": {"clusters": int(ncl), "used_beams": used_beams, "n_after_resample": int(side.cur.sample_count),
                   "leaf_count": int(side.cur.kdtree.contents.leaf_count)},
    }
    if not args.no_cpu_baseline:
        # a fresh process: the CPU reference never shares an address space with CUDA
        out = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference",
                              "--steps", "1", "--warmup", "0", "--cpu-workers", "1",
                              "--map-size", str(args.map_size), "--beams", str(args.beams),
                              "--cpu-particles", str(args.cpu_particles)],
                             capture_output=True, text=True)
        cb = None
        for ln in out.stdout.splitlines():
            if ln.startswith("{"):
                cb = json.loads(ln)
        if cb is not None and "cpu_baseline" in cb:
            line["cpu_baseline"] = dict(cb["cpu_baseline"], ms_per_step=cb["ms_per_step"])
        else:
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference",
                                    "sample": "unavailable: " + (out.stderr[-200:] if out.stderr else "oracle/_ref not built")}
    print(json.dumps(line))


def main():
    args = parse()
    if args.impl == "reference":
        reference_arm(args)
    else:
        b200_arm(args)


if __name__ == "__main__":
    main()
