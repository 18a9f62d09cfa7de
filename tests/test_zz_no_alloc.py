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


def _static_distances(lib, torch, occ, scale, max_d):
    sy, sx = occ.shape
    R = int(max_d / scale)                       # map_cspace.cpp:42
    d_occ = torch.from_numpy(occ).cuda()         # caller-owned input
    d_codes = torch.full((sy, sx), -1, dtype=torch.int16, device="cuda")   # caller-owned output
    rc = lib.amclb200_edt_codes_static(d_occ.data_ptr(), sx, sy, R, d_codes.data_ptr(), None)
    assert rc == 0
    torch.cuda.synchronize()
    codes = d_codes.cpu().numpy().view(np.uint16).astype(np.int64)
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
