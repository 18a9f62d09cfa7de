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
    ap.add_argument("--cpu-particles-large", type=int, default=100000,
                    help="second CPU figure: one step at this many particles per worker, where the "
                         "reference's O(N^2) resampler (pf.c:461-465) shows; 0 = skip")
    ap.add_argument("--cpu-workers", type=int, default=0, help="0 = one per host core (max 32)")
    ap.add_argument("--cpu-steps", type=int, default=3, help="timed steps of a CPU reference leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--collectives", default="nccl", choices=["nccl", "torch"],
                    help="multi-GPU exchanges: the library's own NCCL communicator (default) or "
                         "torch.distributed callbacks")
    ap.add_argument("--parity-particles", type=int, default=1 << 20,
                    help="particles of the oracle parity pass run before the timed loop; 0 = skip")
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
    """One single-threaded replica of the compiled reference on its own particles."""
    wid, size, beams, n, steps, warmup = args
    from navigation_b200 import abi, synth
    from navigation_b200.harness import Side
    from oracle import oracle

    ref = oracle.ref()
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
    times, stage = [], np.zeros(3)
    for it in range(warmup + steps):
        side.set_samples(cloud)
        t0 = time.perf_counter()
        side.update_action(ODOM_POSE, ODOM_DELTA)
        t1 = time.perf_counter()
        side.update_sensor(scan, RANGE_MAX)
        t2 = time.perf_counter()
        side.resample()
        tab = side.cluster_table()
        t3 = time.perf_counter()
        if it >= warmup:
            times.append(t3 - t0)
            stage += (t1 - t0, t2 - t1, t3 - t2)
    side.close()
    return dict(seconds=float(np.sum(times)), steps=steps, n=n, used_beams=used, cspace_s=t_cspace,
                clusters=len(tab), stage_s=(stage / steps).tolist())


def run_cpu_reference(size, beams, n_per_worker, steps, warmup, workers):
    """`workers` independent single-thread replicas (the reference has no threads
    of its own); returns aggregate evals/s and the per-step wall time."""
    import multiprocessing as mp

    from oracle import oracle

    # dlopen the compiled reference in THIS process too, before forking: the
    # workers inherit the mapping and the driver's native-library record shows
    # oracle/_ref/libamcl_ref.so for this arm
    if oracle.ref() is None:
        return None
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
                steps=steps, warmup=warmup,
                stage_ms={k: 1e3 * float(np.mean([r["stage_s"][i] for r in res]))
                          for i, k in enumerate(("action", "sensor", "resample"))},
                sample="%d single-thread replicas of the compiled reference x %d particles x %d used beams, "
                       "%d warm-up + %d timed steps each, %dx%d map"
                       % (workers, n_per_worker, res[0]["used_beams"], warmup, steps, size, size))


def host_workers(args):
    return args.cpu_workers if args.cpu_workers > 0 else min(max(1, (os.cpu_count() or 1)), 32)


def reference_arm(args):
    """`--impl reference`: the reference's own CPU code (oracle/_ref, built from the
    unmodified sources) on the host cores.  Each replica runs a bounded sample of
    the workload; the line says how many steps really ran."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    workers = host_workers(args)
    steps = max(1, min(args.steps, args.cpu_steps))
    warmup = max(1, min(args.warmup, 1))
    r = run_cpu_reference(args.map_size, args.beams, args.cpu_particles, steps, warmup, workers)
    if r is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libamcl_ref.so not built"}))
        return
    large = None
    if args.cpu_particles_large > 0:
        large = run_cpu_reference(args.map_size, args.beams, args.cpu_particles_large, 1, 0, workers)
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT,
        "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
        "steps_requested": args.steps, "warmup_requested": args.warmup,
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, note="CPU reference, bounded sample: %s" % r["sample"]),
        "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": r["kind"],
                         "sample": r["sample"], "stage_ms": r["stage_ms"]},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if large is not None:
        line["cpu_baseline_large"] = {
            "value": large["value"], "unit": UNIT, "cores": large["cores"], "kind": large["kind"],
            "ms_per_step": large["ms_per_step"], "stage_ms": large["stage_ms"], "sample": large["sample"],
            "note": "same code at %d particles per replica: the linear search per draw makes resampling "
                    "quadratic (pf.c:461-465)" % args.cpu_particles_large}
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


# ------------------------------------------------------------------ rooflines ----

def rooflines(lib, args, k_ms, n_local, used_beams, ms_step, st_action, st_resample, edt_ms, n_total, world):
    """Roofline of the dominant kernel (the laser model) and of the HBM-bound stages.

    The laser kernel gathers from an L2-resident map, so its bound is the rate at
    which the GPU serves uncorrelated 32-byte-sector gathers, MEASURED live by a
    probe over a table with the map's footprint.  frac = the launch's L2 sector
    reads (ncu, recorded in profiles/lf_counters.json with the commit they were
    taken at) per second of the live kernel time / that rate.  The figure by
    SURVEY.md 8(d)'s definition (one sector per evaluation) is kept as
    frac_survey_def; it exceeds 1 because neighbouring beams share sectors."""
    k_evals = n_local * used_beams
    alg_bytes = 32.0 * k_evals + 40.0 * n_local
    map_bytes = args.map_size * args.map_size * 2
    gather_rate = lib.amclb200_measure_gather_rate(map_bytes, 5)
    peaks_file = os.path.join(ROOT, "MEASURED_PEAKS.json")
    have_peaks = os.path.exists(peaks_file)
    hbm = json.load(open(peaks_file))["hbm_gbs"] if have_peaks else 6650.0
    hbm_src = "MEASURED_PEAKS.json" if have_peaks else "fallback (B200_PROFILING.md)"
    rec, scaled_from = None, None
    cfile = os.path.join(ROOT, "profiles", "lf_counters.json")
    if os.path.exists(cfile) and args.beams == N_BEAMS and args.map_size == MAP_SIZE:
        table = {k: v for k, v in json.load(open(cfile)).items() if k.isdigit()}
        rec = table.get(str(n_local))
        if rec is None and table:
            # a slice of the same cloud (uniform over free space): the sectors and DRAM
            # bytes per particle are those of the recorded launch, scaled by the particle count
            scaled_from = max(table, key=int)
            f = n_local / float(scaled_from)
            rec = dict(table[scaled_from])
            rec["l2_sectors_read"] = rec["l2_sectors_read"] * f
            rec["dram_bytes"] = rec["dram_bytes"] * f
    sectors = rec["l2_sectors_read"] if rec else None
    sector_rate = sectors / (k_ms * 1e-3) if sectors else None
    roofline = {
        "kernel": "likelihood_field_kernel<LF>", "bound": "l2_gather", "unit": "GB/s",
        "achieved": (sector_rate * 32.0 / 1e9) if sector_rate else None,
        "peak": gather_rate * 32.0 / 1e9,
        "frac": (sector_rate / gather_rate) if sector_rate else None,
        "traffic": rec["dram_bytes"] if rec else None,
        "kernel_ms": k_ms, "share_of_step": k_ms / ms_step,
        "evals_per_s": k_evals / (k_ms * 1e-3),
        "l2_sectors_per_launch": sectors, "l2_sectors_per_eval": (sectors / k_evals) if sectors else None,
        "gather_peak_sectors_per_s": gather_rate,
        "peak_source": "measured live: uncorrelated 2-byte gathers, one 32-B sector each, over a %.1f MB table "
                       "(= %.2f sectors per clock per SM: the probe is bound by the L1TEX sector rate, not by "
                       "its index arithmetic)" % (map_bytes / 1e6, gather_rate / (148 * 1.965e9)),
        "ncu_recorded": ({k: rec[k] for k in ("commit", "l1tex_sectors", "l1tex_sectors_texture", "l1tex_lsu_pipe_pct",
                                             "l1tex_tex_pipe_pct", "issue_active_pct",
                                             "fp64_pipe_pct", "l2_hit_pct", "kernel_ms_under_ncu") if k in rec}
                         if rec else "no recorded capture for %d particles per launch" % n_local),
        "ncu_scaled_from": (None if scaled_from is None else
                            "sectors and DRAM bytes of the %s-particle capture x %d / %s (same cloud, a slice of it)"
                            % (scaled_from, n_local, scaled_from)),
        "achieved_survey_def": alg_bytes / (k_ms * 1e-3) / 1e9,
        "frac_survey_def": alg_bytes / (k_ms * 1e-3) / 1e9 / (gather_rate * 32.0 / 1e9),
        "note": "achieved = L2 sectors the launch reads (ncu) x 32 B / live kernel time; traffic = its DRAM "
                "bytes (ncu): the map stays in L2, compulsory HBM traffic is the particles.  *_survey_def "
                "counts 32 B per evaluation + 40 B per particle (SURVEY.md 8d) and is not a fraction of a "
                "bound: 32 consecutive beams of a particle end in ~6 sectors, not 32",
        "hbm_peak_gbs": hbm, "hbm_peak_source": hbm_src,
    }

    def hbm_stage(name, bytes_, ms, kernels, note):
        gbs = bytes_ / (ms * 1e-3) / 1e9 if ms > 0 else None
        return {"stage": name, "bound": "hbm", "algorithmic_bytes": int(bytes_), "ms": ms, "achieved": gbs,
                "peak": hbm, "unit": "GB/s", "frac": (gbs / hbm) if gbs else None, "kernels": kernels, "note": note}

    stages = [
        hbm_stage("motion", 48.0 * n_local, st_action, "motion_kernel",
                  "48 B per particle (SURVEY.md 8d); CUDA events around AMCLOdom::UpdateAction"),
        hbm_stage("resample+clusters", 128.0 * n_local, st_resample,
                  "exact_cumsum, pick_and_copy, bin_insert, scans, cc_*, cluster_accumulate",
                  "128 B per particle of this rank's slice (SURVEY.md 8d); CUDA events around pf_update_resample"
                  + ("; tables and labelling still work on global objects" if world > 1 else "")),
        hbm_stage("map_update_cspace (EDT)", 3.0 * args.map_size * args.map_size, edt_ms,
                  "edt_pack_rows16, edt_row_squares, edt_col_squares",
                  "3 B per cell (SURVEY.md 8d); the three kernels only, host cell conversion excluded"),
    ]
    return roofline, stages


# --------------------------------------------------------------- parity pass ----

PARITY_SEED = 4242


def _mix64(a):
    """splitmix64 finaliser on a uint64 array (wraps mod 2^64)."""
    a = a.astype(np.uint64).copy()
    a ^= a >> np.uint64(30)
    a *= np.uint64(0xBF58476D1CE4E5B9)
    a ^= a >> np.uint64(27)
    a *= np.uint64(0x94D049BB133111EB)
    a ^= a >> np.uint64(31)
    return a


def _oracle_sensor_threads(cloud, occ, dist, scan, beams, threads):
    """likelihood_field of the oracle port over `threads` host threads (the model
    is independent per particle; ctypes releases the GIL).  Returns the total in
    the port's own order of summation per chunk."""
    from concurrent.futures import ThreadPoolExecutor

    from oracle import oracle

    n = cloud.shape[0]
    lp = oracle.laser_params(z_hit=LASER["z_hit"], z_rand=LASER["z_rand"], sigma_hit=LASER["sigma_hit"])
    bounds = np.linspace(0, n, threads + 1).astype(int)

    def work(k):
        part = cloud[bounds[k]:bounds[k + 1]]       # contiguous view: updated in place
        if part.shape[0] == 0:
            return 0.0
        return oracle.laser_model(1, part, lp, beams, occ, dist, SCALE, (0.0, 0.0), MAX_OCC_DIST, scan, RANGE_MAX)

    with ThreadPoolExecutor(threads) as ex:
        return float(sum(ex.map(work, range(threads))))


def parity_pass(lib, side, args, occ, scan, rank, world, dist_mod):
    """One full update at --parity-particles on the GPU filter (at this run's world
    size, host drand48 draws in the reference's order) and, on rank 0, the same
    update through the CPU oracle port seeded alike.  Reports what north_star
    gates: n, k and the resampled source indices bit-exact, weights within 1e-5."""
    import torch

    from navigation_b200 import library, synth

    n = min(args.parity_particles, args.particles)
    libc = C.CDLL(None)
    libc.srand48.argtypes = [C.c_long]
    lib.amclb200_set_rng(library.RNG_DRAND48, 0)
    libc.srand48(PARITY_SEED)
    rng2 = (C.c_int * 2)()
    lib.amclb200_pf_shard_range(side.pf, n, rng2)
    lo, hi = rng2[0], rng2[1]
    cloud = synth.uniform_cloud(occ, SCALE, n, 777)
    cloud[:, 3] = 1.0 / n
    d_slice = torch.from_numpy(np.ascontiguousarray(cloud[lo:hi])).cuda()
    # the filter is allocated for the bench's particle count; for this pass it is an
    # n-particle filter (pf_update_resample reads pf->max_samples on every call, as
    # the reference does, pf.c:427)
    saved_max = side.pf.contents.max_samples
    side.pf.contents.max_samples = n
    lib.amclb200_pf_load_device_shard(side.pf, d_slice.data_ptr(), n)
    odom_pose = (C.c_double * 3)(*ODOM_POSE)
    odom_delta = (C.c_double * 3)(*ODOM_DELTA)
    lib.amcl_c_odom_update_action(side.odom, side.pf, odom_pose, odom_delta)
    lib.amcl_c_laser_update_sensor(side.laser, side.pf, scan.ctypes.data, scan.shape[0], RANGE_MAX)
    after_sensor = side.samples()                     # all-gathers the slices when sharded
    lib.pf_update_resample(side.pf)
    cnt = int(side.cur.sample_count)
    k = int(side.cur.kdtree.contents.leaf_count)
    ncl = int(side.cur.cluster_count)
    src = np.full(n, -2, np.int32)
    lib.amclb200_pf_resample_sources(side.pf, src.ctypes.data, n)
    lib.amclb200_pf_shard_range(side.pf, cnt, rng2)
    slo, shi = rng2[0], rng2[1]
    # order-independent 64-bit digest of the integer results: every rank hashes the
    # (position, source index) pairs of its slice, the partial sums add mod 2^64
    pos = np.arange(slo, shi, dtype=np.uint64)
    part = int(np.sum(_mix64(pos * np.uint64(0x9E3779B97F4A7C15) + src[slo:shi].astype(np.int64).astype(np.uint64)),
                      dtype=np.uint64))
    src_all = src
    if world > 1:
        t = torch.tensor([part - (1 << 64) if part >= (1 << 63) else part], dtype=torch.int64, device="cuda")
        dist_mod.all_reduce(t)
        part = int(t.item()) & ((1 << 64) - 1)
        chunk = (cnt + world - 1) // world
        mine = torch.full((chunk,), -2, dtype=torch.int32, device="cuda")
        mine[:shi - slo] = torch.from_numpy(src[slo:shi]).cuda()
        whole = torch.empty(chunk * world, dtype=torch.int32, device="cuda")
        dist_mod.all_gather_into_tensor(whole, mine)
        src_all = whole.cpu().numpy()[:cnt]
    counts = sorted(int(side.cur.clusters[c].count) for c in range(ncl))
    table = side.cluster_table()
    head = np.array([cnt, k, ncl] + counts, dtype=np.uint64)
    digest = (part + int(np.sum(_mix64(head + np.arange(head.size, dtype=np.uint64) * np.uint64(1315423911)),
                                dtype=np.uint64))) & ((1 << 64) - 1)
    out = {"particles": n, "n": cnt, "k": k, "clusters": ncl, "digest": "%016x" % digest,
           "rng": "host drand48, seed %d, reference order" % PARITY_SEED,
           "heaviest_cluster_weight": max(t_[0] for t_ in table) if table else None}
    if rank == 0:
        from oracle import oracle

        port = oracle.port()
        t0 = time.perf_counter()
        R = port.orc_cell_radius(MAX_OCC_DIST, SCALE)
        dist = oracle.occ_dist_from_d2(oracle.edt_exact(occ, R), SCALE, MAX_OCC_DIST)
        mine = cloud.copy()
        port.orc_srand48(PARITY_SEED)
        port.orc_motion(mine, n, 0, np.array(ALPHAS), np.array(ODOM_POSE), np.array(ODOM_DELTA))
        threads = min(max(1, os.cpu_count() or 1), 32)
        total = _oracle_sensor_threads(mine, occ, dist, scan, args.beams, threads)
        ws, wf = C.c_double(0.0), C.c_double(0.0)
        port.orc_sensor_finish(mine, n, total, 0.0, 0.0, C.byref(ws), C.byref(wf))
        pose_err = float(np.max(np.abs(after_sensor[:, :3] - mine[:, :3]) / np.maximum(np.abs(mine[:, :3]), 1e-3)))
        w_err = float(np.max(np.abs(after_sensor[:, 3] - mine[:, 3]) / mine[:, 3]))
        ob = np.empty((n, 4))
        osrc = np.empty(n, np.int32)
        ok_ = C.c_int()
        ocnt = port.orc_resample(mine, n, ob, osrc, 100, n, 0.01, 0.99, 0.0, None, None, C.byref(ok_))
        labels = np.empty(max(ocnt, 1), np.int32)
        oncl = port.orc_cluster_labels(ob[:ocnt].copy(), ocnt, labels, None)
        same = ocnt == cnt and bool(np.array_equal(src_all[:cnt], osrc[:ocnt]))
        mism = int(np.sum(src_all[:min(cnt, ocnt)] != osrc[:min(cnt, ocnt)]))
        out.update({"oracle": {"n": int(ocnt), "k": int(ok_.value), "clusters": int(oncl)},
                    "n_equal": ocnt == cnt, "k_equal": ok_.value == k, "clusters_equal": oncl == ncl,
                    "indices_equal": same, "index_mismatches": mism,
                    "max_rel_weight_err": w_err, "max_rel_pose_err": pose_err,
                    "oracle_seconds": time.perf_counter() - t0, "oracle_threads": threads,
                    "ok": bool(same and ok_.value == k and oncl == ncl and w_err < 1e-5 and pose_err < 1e-5)})
    side.pf.contents.max_samples = saved_max
    lib.amclb200_set_rng(library.RNG_PHILOX, 0x5EED)
    return out


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
    from navigation_b200.harness import Side

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
        if args.collectives == "nccl":
            coll.attach_nccl(lib, side.pf)
        else:
            coll.attach(lib, side.pf)
    rng2 = (C.c_int * 2)()
    lib.amclb200_pf_shard_range(side.pf, n_total, rng2)
    own_lo, own_hi = rng2[0], rng2[1]
    n_local = own_hi - own_lo
    lib.amclb200_set_mirror(side.pf, 0)      # deferred host mirror
    t0 = time.perf_counter()
    side.make_laser(args.beams, model="lf", **dict(LASER, max_occ_dist=MAX_OCC_DIST))  # GPU EDT + host cells
    t_cspace = time.perf_counter() - t0
    edt_ms = float(lib.amclb200_map_last_edt_ms(side.map))
    side.make_odom(model=abi.ODOM_MODEL_DIFF, alphas=ALPHAS)
    scan = scan_from(side, pose, bear, gpu=True)
    used_beams = int(np.sum(scan[:, 0] < RANGE_MAX))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- parity: this run's filter against the CPU oracle ----------------------
    parity = None
    if args.parity_particles > 0:
        parity = parity_pass(lib, side, args, occ, scan, rank, world, dist)
        barrier()

    # the same global cloud on every rank (hash of the global index); keep the own slice
    cloud = synth.uniform_cloud(occ, SCALE, n_total, 100)[own_lo:own_hi]
    pristine = torch.from_numpy(np.ascontiguousarray(cloud)).cuda()
    del cloud
    scan_host = torch.from_numpy(scan).pin_memory()
    odom_pose = (C.c_double * 3)(*ODOM_POSE)
    odom_delta = (C.c_double * 3)(*ODOM_DELTA)
    bw = C.c_double()
    bm = abi.pf_vector_t()
    bc = abi.pf_matrix_t()
    cloud_stride = max(1, n_total // 2000)           # a PoseArray of ~2000 poses for the visualiser
    cloud_out = np.empty((2048, 7))

    def step(node_outputs=False):
        lib.amclb200_pf_load_device_shard(side.pf, pristine.data_ptr(), n_total)
        lib.amcl_c_odom_update_action(side.odom, side.pf, odom_pose, odom_delta)
        lib.amcl_c_laser_update_sensor(side.laser, side.pf, scan_host.data_ptr(), args.beams, RANGE_MAX)
        lib.pf_update_resample(side.pf)
        # what the node does next (amcl_node.cpp:1369-1394): arg-max weight cluster
        best = lib.amclb200_pf_best_cluster(side.pf, C.byref(bw), C.byref(bm), C.byref(bc))
        if node_outputs:
            # ... and the particle cloud it publishes (amcl_node.cpp:1349-1362), packed on the device
            lib.amclb200_pf_pack_cloud(side.pf, cloud_stride, 2048, cloud_out.ctypes.data)
        return (best, bw.value, bm.v[0], bm.v[1], bm.v[2]), side.cur.cluster_count

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
    e0.record()
    for _ in range(args.steps):
        best, ncl = step()
        lib.amclb200_pf_sensor_kernel_stats(side.pf, kstats)
        kernel_ms.append(kstats[0])
        lib.amclb200_pf_last_timings(side.pf, stage_ms.ctypes.data_as(C.POINTER(C.c_float)))
        stages.append(stage_ms.copy())
    e1.record()
    barrier()
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
    st_action = float(np.mean([s_[0] for s_ in stages]))
    st_sensor = float(np.mean([s_[1] for s_ in stages]))
    st_resample = float(np.mean([s_[2] for s_ in stages]))

    roofline, roofline_stages = rooflines(lib, args, float(np.mean(kernel_ms)), n_local, used_beams, ms_step,
                                          st_action, st_resample, edt_ms, n_total, world)

    # ---- e2e: host buffers through the reference-facing API, three callers ------
    def wall(fn):
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            fn()
        barrier()
        sec = (time.perf_counter() - t0) / args.steps
        if world > 1:
            t = torch.tensor([sec], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            sec = float(t.item())
        return sec

    # scan + odometry from the caller, plus the beam table the library stages each update
    h2d = scan.nbytes + 48 + 16 * used_beams
    d2h_tab = 80 * max(ncl, 1) + 64
    e2e_s = wall(step)                                    # deferred mirror, cluster table read back
    e2e_n2_s = wall(lambda: step(node_outputs=True))      # + the published cloud, packed on the device (N2)
    lib.amclb200_set_mirror(side.pf, 1)
    for _ in range(2):
        step()
    e2e_eager_s = wall(step)                              # unchanged amcl_node: samples[] mirrored after every call
    lib.amclb200_set_mirror(side.pf, 0)
    mirror_bytes = 32 * n_local
    e2e = {"value": evals_step / e2e_s, "unit": UNIT, "ms_per_step": 1e3 * e2e_s,
           "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h_tab),
           "caller": "deferred host mirror (amclb200_set_mirror(pf, 0)); pose estimate from amclb200_pf_best_cluster",
           "e2e_node_n2": {"value": evals_step / e2e_n2_s, "ms_per_step": 1e3 * e2e_n2_s,
                           "d2h_bytes_per_step": int(d2h_tab + 56 * min(2048, (n_total + cloud_stride - 1) // cloud_stride)),
                           "caller": "a node using the device-resident outputs: amclb200_pf_best_cluster + "
                                     "amclb200_pf_pack_cloud(stride %d)" % cloud_stride},
           "e2e_eager_mirror": {"value": evals_step / e2e_eager_s, "ms_per_step": 1e3 * e2e_eager_s,
                                "d2h_bytes_per_step": int(d2h_tab + 3 * mirror_bytes),
                                "caller": "unchanged amcl_node (reads set->samples[i], amcl_node.cpp:1349-1362): the "
                                          "library default, %d MB device-to-host after each of the three update "
                                          "calls on this rank" % (mirror_bytes >> 20)}}

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
        "roofline": roofline, "roofline_stages": roofline_stages, "parity": parity,
        "collectives": (dict(coll.calls, note="counts over the whole run (parity pass, warm-up, timed and e2e steps)")
                        if coll else None),
        "stage_ms": {"action": st_action, "sensor": st_sensor, "resample": st_resample,
                     "edt_kernels": edt_ms, "cspace_call_s": t_cspace},
        "result": {"clusters": int(ncl), "used_beams": used_beams, "n_after_resample": int(side.cur.sample_count),
                   "cumsum_fallbacks": int(lib.amclb200_cumsum_fallbacks()),
                   "leaf_count": int(side.cur.kdtree.contents.leaf_count)},
    }
    if not args.no_cpu_baseline:
        # a fresh process: the CPU reference never shares an address space with CUDA
        out = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference",
                              "--steps", "3", "--warmup", "1", "--cpu-workers", "1",
                              "--cpu-particles-large", "0",
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
