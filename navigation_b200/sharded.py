"""Multi-GPU plumbing: the collectives libamcl_b200 calls back into.

One process per GPU (torchrun).  Every rank allocates the filter with the
GLOBAL max_samples and issues the same reference-API calls; the library keeps
only the rank's contiguous particle slice current and asks the host program for
two in-place collectives on raw buffers (include/amcl_b200.h,
amclb200_pf_set_shard):

    allreduce(buf, count, dtype, op)      dtype 0 = fp64, 1 = int32, 2 = int64; op sum/min/max
    allgather(buf, bytes_per_rank)        rank r's block already sits at r*bytes_per_rank

`Collectives` implements both on torch.distributed.  On a GPU the buffers are
device pointers wrapped zero-copy through __cuda_array_interface__ and the
backend is NCCL over NVLink; on a CPU (tests, gloo backend) the very same
callbacks are handed host pointers.  Nothing here computes anything.
"""
import ctypes as C

import numpy as np
import torch
import torch.distributed as dist

ALLREDUCE_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p)
ALLGATHER_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_uint64, C.c_void_p)

F64, I32, I64 = 0, 1, 2
SUM, MIN, MAX = 0, 1, 2

_NP = {F64: (np.float64, "<f8", torch.float64, 8), I32: (np.int32, "<i4", torch.int32, 4),
       I64: (np.int64, "<i8", torch.int64, 8)}
_OPS = {SUM: dist.ReduceOp.SUM, MIN: dist.ReduceOp.MIN, MAX: dist.ReduceOp.MAX}


SHARD_TILE, SHARD_MAX_ALIGNED = 1024, 32
SHARD_PAD = SHARD_TILE * (SHARD_MAX_ALIGNED + 1)      # extra records every particle array carries


def shard_chunk(n, world):
    per = (n + world - 1) // world
    align = SHARD_TILE if 1 < world <= SHARD_MAX_ALIGNED else 1
    return (per + align - 1) // align * align


def shard_range(n, rank, world):
    """[lo, hi) of rank's slice of n items (mirrors Shard in context.hpp): chunks of
    ceil(n/world), rounded up to whole 1024-weight tiles of the cumulative table
    for 2..32 ranks."""
    chunk = shard_chunk(n, world)
    lo = min(rank * chunk, n)
    hi = min((rank + 1) * chunk, n)
    return lo, hi


class _DevicePtr:
    """Zero-copy view of `count` elements at a raw CUDA address."""

    def __init__(self, ptr, count, typestr):
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": typestr,
                                         "data": (int(ptr), False), "version": 2}


class Collectives:
    """In-place collectives over torch.distributed for raw device/host buffers."""

    def __init__(self, group=None, device=None):
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.on_gpu = device is not None and torch.device(device).type == "cuda"
        self.device = torch.device(device) if device is not None else torch.device("cpu")
        self.calls = {"allreduce": 0, "allgather": 0, "allgather_bytes": 0}
        # ctypes trampolines must outlive the filter that stores them
        self._ar = ALLREDUCE_FN(self._allreduce)
        self._ag = ALLGATHER_FN(self._allgather)

    # -- pointer -> tensor -----------------------------------------------------
    def _tensor(self, ptr, count, dtype):
        npdt, typestr, tdt, size = _NP[dtype]
        if self.on_gpu:
            return torch.as_tensor(_DevicePtr(ptr, count, typestr), device=self.device)
        buf = (C.c_char * (count * size)).from_address(ptr)
        return torch.from_numpy(np.frombuffer(buf, dtype=npdt, count=count))

    # -- callbacks ---------------------------------------------------------------
    def _allreduce(self, ptr, count, dtype, op, _user):
        t = self._tensor(ptr, count, dtype)
        dist.all_reduce(t, op=_OPS[op], group=self.group)
        self.calls["allreduce"] += 1

    def _allgather(self, ptr, bytes_per_rank, _user):
        # moved as int32 words: every block this library gathers is a multiple of 4 bytes
        words = int(bytes_per_rank) // 4
        assert words * 4 == bytes_per_rank
        whole = self._tensor(ptr, words * self.world, I32)
        mine = whole[self.rank * words:(self.rank + 1) * words]
        if self.on_gpu:
            dist.all_gather_into_tensor(whole, mine, group=self.group)   # in place (NCCL)
        else:
            parts = [torch.empty_like(mine) for _ in range(self.world)]
            dist.all_gather(parts, mine.clone(), group=self.group)
            for r, p in enumerate(parts):
                whole[r * words:(r + 1) * words] = p
        self.calls["allgather"] += 1
        self.calls["allgather_bytes"] += int(bytes_per_rank) * (self.world - 1)

    # -- public --------------------------------------------------------------------
    def allreduce(self, ptr, count, dtype, op):
        self._allreduce(ptr, count, dtype, op, None)

    def allgather(self, ptr, bytes_per_rank):
        self._allgather(ptr, bytes_per_rank, None)

    def attach_nccl(self, lib, pf, peer_access=True):
        """The same, with the exchanges on an NCCL communicator owned by the library
        (amclb200_pf_set_shard_nccl): this object only distributes the 128-byte
        ncclUniqueId; no collective of the filter passes through Python after that."""
        assert self.on_gpu, "the library's own NCCL path needs device buffers"
        ident = (C.c_ubyte * 128)()
        if self.rank == 0:
            assert lib.amclb200_nccl_unique_id(ident) == 0
        t = torch.tensor(list(ident), dtype=torch.uint8, device=self.device)
        dist.broadcast(t, src=0, group=self.group)
        ident = (C.c_ubyte * 128)(*t.cpu().tolist())
        lib.amclb200_pf_set_shard_nccl(pf, self.rank, self.world, ident)
        self.calls["in_library_nccl"] = True
        if peer_access and self.world > 1:
            lib.amclb200_pf_enable_peer_access(pf)

    def attach(self, lib, pf, peer_access=True):
        """Make `pf` this rank's shard of a world-wide filter.  With peer_access
        (GPU only) every rank's particle arrays are mapped into every process, so
        resampling reads picked poses over NVLink instead of all-gathering them."""
        lib.amclb200_pf_set_shard(pf, self.rank, self.world,
                                  C.cast(self._ar, C.c_void_p), C.cast(self._ag, C.c_void_p), None)
        if peer_access and self.on_gpu and self.world > 1:
            lib.amclb200_pf_enable_peer_access(pf)
