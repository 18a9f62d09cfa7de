"""World-size-2 checks of the multi-GPU plumbing on CPU (gloo backend).

The CUDA library cannot run here, so these tests drive what surrounds it:
  * the slice arithmetic (must match Shard in context.hpp),
  * the two collectives the library calls back into, invoked THROUGH their C
    function pointers on raw host buffers exactly as the library invokes them
    on device buffers,
  * the sharding plan itself, with the CPU oracle standing in for the kernels:
    each rank weights its slice, totals are all-reduced, slices all-gathered,
    and the result must equal the one-process run.
"""
import ctypes as C
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

from navigation_b200 import sharded, synth


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        coll = sharded.Collectives(device=None)
        # --- allreduce through the C function pointer, every dtype / op -------
        ar = C.cast(coll._ar, C.c_void_p).value
        ar_fn = sharded.ALLREDUCE_FN(ar)
        buf = np.array([1.5 + rank, -2.0 * (rank + 1), 10.0], dtype=np.float64)
        ar_fn(buf.ctypes.data, 3, sharded.F64, sharded.SUM, None)
        assert np.array_equal(buf, [1.5 + 2.5, -2.0 - 4.0, 20.0])
        ib = np.array([5 + rank, 9 - rank], dtype=np.int32)
        ar_fn(ib.ctypes.data, 2, sharded.I32, sharded.MIN, None)
        assert np.array_equal(ib, [5, 8])
        ib = np.array([rank, 1 - rank], dtype=np.int32)
        ar_fn(ib.ctypes.data, 2, sharded.I32, sharded.MAX, None)
        assert np.array_equal(ib, [1, 1])
        # --- in-place allgather of 32-byte particle records -------------------
        n = 11
        lo, hi = sharded.shard_range(n, rank, world)
        chunk = sharded.shard_chunk(n, world)
        full = np.full((world * chunk, 4), -1.0)
        truth = np.arange(world * chunk * 4, dtype=np.float64).reshape(-1, 4)
        full[lo:hi] = truth[lo:hi]
        ag_fn = sharded.ALLGATHER_FN(C.cast(coll._ag, C.c_void_p).value)
        ag_fn(full.ctypes.data, chunk * 32, None)
        assert np.array_equal(full[:n], truth[:n])

        # --- the sharding plan with the oracle as the kernel -------------------
        from oracle import oracle
        port_lib = oracle.port()
        occ = synth.rooms_map(200, seed=5, room=50, door=(20, 30))
        R = port_lib.orc_cell_radius(1.0, 0.05)
        dist_map = oracle.occ_dist_from_d2(oracle.edt_exact(occ, R), 0.05, 1.0)
        pose = synth.room_centre_pose(200, 0.05, room=50)
        B, n = 40, 1001
        bear = synth.scan_bearings(B)
        rng = np.array([port_lib.orc_calc_range(occ, 200, 200, 0.0, 0.0, 0.05, pose[0], pose[1],
                                                pose[2] + b, 8.0) for b in bear])
        scan = np.stack([np.minimum(rng, 8.0), bear], 1).copy()
        cloud = synth.gaussian_cloud(pose, (0.3, 0.3, 0.2), n, 3)
        lp = oracle.laser_params()
        # one process
        whole = cloud.copy()
        total = oracle.laser_model(1, whole, lp, B, occ, dist_map, 0.05, (0.0, 0.0), 1.0, scan, 8.0)
        whole[:, 3] /= total
        # two ranks: weight the slice, all-reduce the total, normalise, all-gather
        lo, hi = sharded.shard_range(n, rank, world)
        chunk = sharded.shard_chunk(n, world)
        mine = np.zeros((world * chunk, 4))
        mine[:n] = cloud
        part = np.ascontiguousarray(mine[lo:hi])
        t = np.array([oracle.laser_model(1, part, lp, B, occ, dist_map, 0.05, (0.0, 0.0), 1.0,
                                         scan, 8.0)])
        coll.allreduce(t.ctypes.data, 1, sharded.F64, sharded.SUM)
        part[:, 3] /= t[0]
        mine[lo:hi] = part
        coll.allgather(mine.ctypes.data, chunk * 32)
        assert abs(t[0] - total) <= 1e-12 * total
        assert np.array_equal(mine[:n, :3], whole[:, :3])
        assert np.allclose(mine[:n, 3], whole[:, 3], rtol=1e-12, atol=0)
        assert coll.calls["allreduce"] >= 4 and coll.calls["allgather"] == 2
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, traceback.format_exc()))
        raise e
    finally:
        dist.destroy_process_group()


def test_shard_range_covers_everything():
    for n in (0, 1, 7, 8, 9, 1000, 16 * 1024 * 1024):
        for world in (1, 2, 3, 8):
            got = []
            for r in range(world):
                lo, hi = sharded.shard_range(n, r, world)
                assert 0 <= lo <= hi <= n
                got.extend([lo, hi])
            assert got[0] == 0 and got[-1] == n
            assert all(got[2 * r + 1] == got[2 * r + 2] for r in range(world - 1))
            chunk = sharded.shard_chunk(n, world)
            assert world * chunk <= n + sharded.SHARD_PAD      # the padding pf_ctx_create reserves
            assert world == 1 or chunk % sharded.SHARD_TILE == 0


def test_collectives_and_plan_world_size_2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in out:
        assert msg == "ok", "rank %d: %s" % (rank, msg)
