"""Times the other BASELINE.json configurations (K1-K4) on one B200 and writes
profiles/configs_r2.json.  bench.py carries the headline (K5); these are the
parity-test shapes, timed here so their kernels have numbers too.

    python profiles/run_configs.py            # on the GPU box
"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from navigation_b200 import abi, library, synth  # noqa: E402
from navigation_b200.harness import Side  # noqa: E402

lib = library.load()
lib.amclb200_set_rng(library.RNG_PHILOX, 7)
HBM = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
    os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0


def gpu_scan(side, pose, beams, range_max):
    bear = synth.scan_bearings(beams)
    rays = np.stack([np.full_like(bear, pose[0]), np.full_like(bear, pose[1]), pose[2] + bear], 1).copy()
    out = np.empty(beams)
    lib.amclb200_map_calc_ranges(side.map, rays.ctypes.data, beams, range_max, out.ctypes.data)
    return np.stack([np.minimum(out, range_max), bear], 1).copy()


def timed(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        ts.append(time.perf_counter() - t0)
    return float(np.median(ts)) * 1e3


def sensor_case(name, size, n, beams, model, cloud_kind="global", resample=False, range_max=30.0):
    occ = synth.rooms_map(size, seed=1)
    side = Side(lib, occ)
    side.alloc(100, n)
    lib.amclb200_set_mirror(side.pf, 0)
    side.make_laser(beams, model=model)
    pose = synth.room_centre_pose(size, 0.05)
    scan = gpu_scan(side, pose, beams, range_max)
    used = int(np.sum(scan[:, 0] < range_max)) if model != "beam" else beams
    cloud = synth.uniform_cloud(occ, 0.05, n, 3) if cloud_kind == "global" else \
        synth.gaussian_cloud(pose, (0.5, 0.5, 0.26), n, 3)
    pristine = torch.from_numpy(cloud).cuda()
    k = (C.c_double * 2)()
    st = np.zeros(3, np.float32)

    def step():
        lib.amclb200_pf_load_device(side.pf, pristine.data_ptr(), n)
        side.update_sensor(scan, range_max)
        if resample:
            side.resample()

    ms = timed(step)
    lib.amclb200_pf_last_timings(side.pf, st.ctypes.data_as(C.POINTER(C.c_float)))
    lib.amclb200_pf_sensor_kernel_stats(side.pf, k)
    out = dict(config=name, particles=n, beams=beams, used_beams=used, map=size, model=model,
               cloud=cloud_kind, ms_per_call=ms, sensor_ms=float(st[1]), resample_ms=float(st[2]) if resample else None,
               evals_per_s=n * used / (float(st[1]) * 1e-3), n_after=int(side.cur.sample_count),
               leaf_count=int(side.cur.kdtree.contents.leaf_count), clusters=int(side.cur.cluster_count))
    if model != "beam":
        out["kernel_ms"] = k[0]
        out["kernel_evals_per_s"] = k[1] / (k[0] * 1e-3)
    side.close()
    return out


def edt_case(size):
    occ = synth.rooms_map(size, seed=1)
    side = Side(lib, occ)
    t_api = timed(lambda: side.update_cspace(2.0), reps=3, warm=1)     # upload + kernels + host cells
    t_dev = lib.amclb200_map_last_edt_ms(side.map)
    lib.amclb200_map_set_mirror(side.map, 0)                            # device-authoritative map
    t_api_dev = timed(lambda: side.update_cspace(2.0), reps=5, warm=1)
    lib.amclb200_map_set_mirror(side.map, 1)
    d_occ = torch.from_numpy(occ).cuda()
    codes = torch.empty((size, size), dtype=torch.int16, device="cuda")
    ok = size * size <= lib.amclb200_edt_static_max_cells()
    t_k = None
    if ok:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        lib.amclb200_edt_codes_static(d_occ.data_ptr(), size, size, 40, codes.data_ptr(), None)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(10):
            lib.amclb200_edt_codes_static(d_occ.data_ptr(), size, size, 40, codes.data_ptr(), None)
        e1.record()
        torch.cuda.synchronize()
        t_k = e0.elapsed_time(e1) / 10
    side.close()
    return dict(config="K3 map_update_cspace", map=size, api_ms=t_api, api_ms_device_authoritative=t_api_dev,
                edt_kernels_ms=t_dev,
                edt_hbm_frac=3 * size * size / (t_dev * 1e-3) / 1e9 / HBM, kernels_ms_static_variant=t_k,
                algorithmic_bytes=3 * size * size,
                hbm_frac=(3 * size * size / (t_k * 1e-3) / 1e9 / HBM) if t_k else None)


def inflation_case(size):
    """N3: InflationLayer::updateCosts over the whole K3 map (costmap defaults: radius 0.55 m, weight 10)."""
    from oracle import oracle
    occ = synth.rooms_map(size, seed=1)
    g = np.where(occ == 1, 254, 0).astype(np.uint8)
    d = torch.from_numpy(g).cuda()
    pristine = d.clone()

    def dev():
        d.copy_(pristine)
        lib.amclb200_inflate_costs(d.data_ptr(), 1, size, size, 0, 0, size, size, 0.05, 0.2, 0.55, 10.0, 0)

    t_dev_call = timed(dev)
    t_k = lib.amclb200_inflate_last_ms()
    h = g.copy()
    t_host_call = timed(lambda: lib.amclb200_inflate_costs(h.ctypes.data, 0, size, size, 0, 0, size, size,
                                                           0.05, 0.2, 0.55, 10.0, 0), reps=3, warm=1)
    sub = 1000
    t0 = time.perf_counter()
    oracle.inflate(g[:sub, :sub], None, 0.05, 0.2, 0.55, 10.0)
    t_cpu = (time.perf_counter() - t0) * 1e3 * (size / sub) ** 2
    return dict(config="N3 inflation %d^2, radius 11 cells" % size, map=size, kernels_ms=t_k,
                device_call_ms=t_dev_call, host_call_ms=t_host_call,
                cpu_sweep_ms_scaled_from_1000sq=t_cpu, algorithmic_bytes=2 * size * size,
                hbm_frac=2 * size * size / (t_k * 1e-3) / 1e9 / HBM)


def main():
    # `python profiles/run_configs.py maps` runs the map-side cases only (K3, N3)
    only_maps = len(sys.argv) > 1 and sys.argv[1] == "maps"
    res = []
    res.append(inflation_case(4000))
    if not only_maps:
        res.append(sensor_case("K1 likelihood_field 5000 x 180, 1000^2", 1000, 5000, 180, "lf", resample=True))
        res.append(sensor_case("K2 beam 100k x 360, 2000^2", 2000, 100000, 360, "beam"))
    res.append(edt_case(2000))
    res.append(edt_case(4000))
    if not only_maps:
        for n in (1 << 20, 1 << 22, 1 << 24):
            res.append(sensor_case("K4 likelihood_field_prob + resample, %d x 1080, 4000^2" % n, 4000, n, 1080,
                                   "lfp", resample=True))
        res.append(sensor_case("K4t tracking cloud lfp + resample 4M x 1080", 4000, 1 << 22, 1080, "lfp",
                               cloud_kind="tracking", resample=True))
    out = os.path.join(ROOT, "gpurun_out", "configs_r2_maps.json" if only_maps else "configs_r2.json")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    json.dump(res, open(out, "w"), indent=1)
    for r in res:
        print(json.dumps(r))


if __name__ == "__main__":
    main()
