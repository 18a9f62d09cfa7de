"""amclb200_edt_codes_static: the distance transform for callers that may not
allocate device memory (no-allocation-after-start-up rules in real-time or
certified robot software, or a process whose GPU memory is owned by another
allocator).  The caller's buffers go in and out; the library's only scratch is a
fixed array inside its image.

This file is named so that pytest collects it last: if it fails under `-x`,
every other test has already run.  It must pass on its own in a fresh process
and after the rest of the suite.

The check: through the C ABI, the static entry point and the ordinary one
(map_update_cspace) must produce the same distance for every cell, bit for bit
(the tolerance of the EDT parity tests), on two different maps in a row -- the
second call reuses the scratch the first one left dirty.
"""
import ctypes as C

import numpy as np
import pytest

from navigation_b200 import synth
from tests.scenario import Side

pytestmark = pytest.mark.gpu


def _static_codes(lib, torch, occ, R, stream=None):
    sy, sx = occ.shape
    d_occ = torch.from_numpy(occ).cuda()         # caller-owned input
    d_codes = torch.full((sy, sx), -1, dtype=torch.int16, device="cuda")   # caller-owned output
    torch.cuda.synchronize()
    rc = lib.amclb200_edt_codes_static(d_occ.data_ptr(), sx, sy, R, d_codes.data_ptr(),
                                       C.c_void_p(stream.cuda_stream) if stream is not None else None)
    assert rc == 0
    torch.cuda.synchronize()
    return d_codes.cpu().numpy().view(np.uint16).astype(np.int64)


def _static_distances(lib, torch, occ, scale, max_d):
    R = int(max_d / scale)                       # map_cspace.cpp:42
    codes = _static_codes(lib, torch, occ, R)
    dist = np.sqrt(codes.astype(np.float64)) * scale
    dist[codes > R * R] = max_d
    return dist


def _reference_entry_distances(lib, occ, scale, max_d):
    side = Side(lib, occ, scale=scale)
    side.update_cspace(max_d)                    # the existing, allocating entry point
    out = side.occ_dist()
    side.close()
    return out


def test_static_edt_matches_map_update_cspace_twice_in_a_row(gpu_lib):
    import torch

    lib = gpu_lib
    first = synth.rooms_map(300, seed=11, room=60, door=(20, 35), clutter_mod=400, unknown_ring=True)[:260]
    second = synth.rooms_map(120, seed=12, room=40, door=(10, 20), clutter_mod=150)[:, :90]
    for occ, scale, max_d in ((np.ascontiguousarray(first), 0.05, 2.0),
                              (np.ascontiguousarray(second), 0.1, 0.75)):
        got = _static_distances(lib, torch, occ, scale, max_d)
        want = _reference_entry_distances(lib, occ, scale, max_d)
        assert np.array_equal(got, want)


def test_static_edt_against_the_oracle_and_without_allocating(gpu_lib, port):
    """Directly against the CPU oracle's exact transform (not only against the other
    CUDA entry point), on shapes that cross the row pass's 512-cell segments, at
    the largest and the smallest radius, on a width that is no multiple of 32 and
    on a caller's own stream; and the property the entry point exists for: the
    free device memory the driver reports is the same before and after a call."""
    import torch

    from oracle import oracle

    lib = gpu_lib
    wide = np.ascontiguousarray(synth.rooms_map(1100, seed=21, room=90, door=(30, 50), clutter_mod=700)[:70])
    odd = np.ascontiguousarray(synth.rooms_map(600, seed=22, room=70, door=(20, 40), clutter_mod=500)[:333, :557])
    cases = [(wide, 254, None), (wide, 0, None), (odd, 40, None), (odd, 17, torch.cuda.Stream())]
    _static_codes(lib, torch, odd, 3)            # set-up call for this GPU, outside the measurement
    for occ, R, stream in cases:
        d2 = oracle.edt_exact(occ, R).astype(np.int64)
        want = np.where((d2 < 0) | (d2 > R * R), R * R + 1, d2)    # the oracle marks "beyond R" with -1
        sy, sx = occ.shape
        d_occ = torch.from_numpy(occ).cuda()
        d_codes = torch.full((sy, sx), -1, dtype=torch.int16, device="cuda")
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info()
        rc = lib.amclb200_edt_codes_static(d_occ.data_ptr(), sx, sy, R, d_codes.data_ptr(),
                                           C.c_void_p(stream.cuda_stream) if stream is not None else None)
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info()
        assert rc == 0
        assert free_after == free_before         # cudaMemGetInfo: nothing was allocated or freed
        got = d_codes.cpu().numpy().view(np.uint16).astype(np.int64)
        assert np.array_equal(got, want), (occ.shape, R)


def test_static_edt_refuses_maps_beyond_its_scratch(gpu_lib):
    import torch

    lib = gpu_lib
    limit = lib.amclb200_edt_static_max_cells()
    assert limit == 4 * 1024 * 1024
    side = 2049                                  # 2049^2 > 4 Mi cells
    d_occ = torch.full((side, side), -1, dtype=torch.int8, device="cuda")
    d_codes = torch.full((side, side), 7, dtype=torch.int16, device="cuda")
    assert lib.amclb200_edt_codes_static(d_occ.data_ptr(), side, side, 10, d_codes.data_ptr(), None) == -2
    assert lib.amclb200_edt_codes_static(d_occ.data_ptr(), 64, 64, 255, d_codes.data_ptr(), None) == -2
    assert lib.amclb200_edt_codes_static(None, 64, 64, 10, d_codes.data_ptr(), None) == -1
    torch.cuda.synchronize()
    assert int((d_codes != 7).sum().item()) == 0   # nothing was written
