"""Worker for the multi-GPU parity check (launched by torchrun, one rank per GPU).

Every rank builds the same world, runs the SHARDED filter (its slice only,
collectives through navigation_b200.sharded.Collectives over NCCL) for a few
full updates, and also runs an ordinary one-GPU filter on the same inputs with
the same Philox seed.  Gates: identical n / leaf_count / cluster count /
resampled source indices (bit-exact), poses bit-exact after the all-gather,
weights and statistics within 1e-9 (only the order of the cross-rank weight
sum differs).  Prints "SHARDED-OK" from rank 0.
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist

    from navigation_b200 import abi, library, sharded, synth
    from tests.scenario import Side

    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = library.load()
    coll = sharded.Collectives(device=torch.device("cuda", local))

    size, n, beams = 600, int(os.environ.get("SHARD_N", "100003")), 90
    occ = synth.rooms_map(size, seed=4)
    pose = synth.room_centre_pose(size, 0.05)
    results = {}
    for mode in ("single", "sharded"):
        lib.amclb200_set_rng(library.RNG_PHILOX, 0xC0FFEE)
        side = Side(lib, occ)
        model = os.environ.get("SHARD_MODEL", "lf")       # lf | lfp | lfp_skip | beam | lf_recover
        recover = model == "lf_recover"
        if recover:
            # recovery injection (pf.c:411-431): a share of the new samples are fresh
            # poses from the caller's host callback; tight cloud, so the KLD bound cuts
            side.alloc(100, n, alpha_slow=0.001, alpha_fast=0.1, pop_err=0.01, pop_z=0.99)
            model = "lf"
        else:
            side.alloc(100, n, pop_err=0.01, pop_z=0.99)
        lib.amclb200_set_mirror(side.pf, 0)
        if model == "lfp_skip":
            side.make_laser(beams, model="lfp", beamskip=(1, 0.5, 0.3, 0.9))
        else:
            side.make_laser(beams, model=model)
        side.make_odom(model=abi.ODOM_MODEL_DIFF)
        bear = synth.scan_bearings(beams)
        rays = np.stack([np.full_like(bear, pose[0]), np.full_like(bear, pose[1]), pose[2] + bear], 1).copy()
        rng = np.empty(beams)
        lib.amclb200_map_calc_ranges(side.map, rays.ctypes.data, beams, 30.0, rng.ctypes.data)
        scan = np.stack([np.minimum(rng, 30.0), bear], 1).copy()
        cloud = synth.uniform_cloud(occ, 0.05, n, 9)
        cloud[: n // 3] = synth.gaussian_cloud(pose, (0.4, 0.4, 0.2), n // 3, 5)
        cloud[:, 3] = 1.0 / n
        if model == "lfp_skip":
            # beam skipping needs a converged set (amcl_laser.cpp:342): a tight cloud
            cloud = synth.gaussian_cloud(pose, (0.05, 0.05, 0.05), n, 5)
            cloud[:, 3] = 1.0 / n
        if recover:
            cloud = synth.gaussian_cloud(pose, (0.3, 0.3, 0.1), n, 5)
            cloud[:, 3] = 1.0 / n
        if mode == "sharded":
            p2p = os.environ.get("SHARD_P2P", "1") == "1"
            if os.environ.get("SHARD_NCCL", "0") == "1":
                coll.attach_nccl(lib, side.pf, peer_access=p2p)     # the library's own communicator
            else:
                coll.attach(lib, side.pf, peer_access=p2p)          # torch.distributed callbacks
        side.set_samples(cloud)            # replicated start on every rank
        log = []
        fallbacks0 = lib.amclb200_cumsum_fallbacks()
        for step in range(3):
            side.update_action((0.1 * step, 0.0, 0.02), (0.1, 0.0, 0.02))
            if model == "lfp_skip":
                side.cur.converged = 1
                side.pf.contents.converged = 1
            side.update_sensor(scan, 30.0)
            neff = side.cur.n_effective
            if recover:
                side.pf.contents.w_slow = 1.0e-3      # w_diff = 0.4
                side.pf.contents.w_fast = 0.6e-3
                side.seed(1000 + step)                # the callback's libc stream; rank 0's poses are used
            side.resample()
            src = np.empty(n, np.int32)
            cnt = lib.amclb200_pf_resample_sources(side.pf, src.ctypes.data, n)
            lo, hi = (0, cnt) if mode == "single" else sharded.shard_range(cnt, rank, world)
            # the decimated PoseArray, packed BEFORE anything replicates the set: every
            # rank packs its own poses and one small all-reduce completes the cloud
            packed = np.zeros((64, 7))
            got = lib.amclb200_pf_pack_cloud(side.pf, max(1, cnt // 50), 64, packed.ctypes.data)
            packed = packed[:got].copy()
            samples = side.samples()       # pulls (all-gathers first when sharded)
            log.append(dict(n=side.cur.sample_count, k=side.cur.kdtree.contents.leaf_count,
                            ncl=side.cur.cluster_count, conv=side.pf.contents.converged, neff=neff,
                            src=src[lo:hi].copy(), lo=lo, hi=hi, samples=samples, packed=packed,
                            mean=np.array(side.cur.mean.tolist()), table=side.cluster_table()))
        # the exact table -- each rank's own slice of it when sharded -- must hold its
        # own recurrence check: a fallback to the replicated walk is correct but is a bug
        assert lib.amclb200_cumsum_fallbacks() == fallbacks0, "cumulative table fell back (%s)" % mode
        results[mode] = log
        side.close()

    for a, b in zip(results["single"], results["sharded"]):
        assert a["n"] == b["n"] and a["k"] == b["k"] and a["ncl"] == b["ncl"] and a["conv"] == b["conv"], \
            (a["n"], b["n"], a["k"], b["k"], a["ncl"], b["ncl"])
        assert abs(a["neff"] - b["neff"]) <= 1e-9 * a["neff"]
        # this rank's slice of the source indices equals the one-GPU run's
        assert np.array_equal(a["src"][b["lo"]:b["hi"]], b["src"])
        assert np.array_equal(a["samples"][:, :3], b["samples"][:, :3])
        assert np.allclose(a["samples"][:, 3], b["samples"][:, 3], rtol=1e-12, atol=0)
        assert np.allclose(a["mean"], b["mean"], rtol=1e-9, atol=1e-12)
        assert a["packed"].shape[0] > 0 and np.array_equal(a["packed"], b["packed"])
        wa = sorted(t[0] for t in a["table"])
        wb = sorted(t[0] for t in b["table"])
        assert np.allclose(wa, wb, rtol=1e-9, atol=1e-15)
    dist.barrier()
    if rank == 0:
        print("SHARDED-OK world=%d n=%d steps=3 collectives=%s" % (world, n, coll.calls))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
