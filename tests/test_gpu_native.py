"""A separately compiled C++ program (tests/native/amcl_node_probe.cpp) that uses
the AMCL libraries the way amcl_node.cpp does -- `new AMCLLaser`, the copy
constructor, UpdateAction / UpdateSensor, reading set->samples[] through the
struct -- is built against include/ + libamcl_b200.so, run on the GPU, and its
output compared with the same program built against the reference's headers
and the compiled reference (oracle/_ref/amcl_node_probe_ref, made by
oracle/Makefile where /root/reference is mounted)."""
import os
import subprocess

import numpy as np
import pytest

from navigation_b200 import library

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "native", "amcl_node_probe.cpp")
REF_EXE = os.path.join(ROOT, "oracle", "_ref", "amcl_node_probe_ref")


def parse(text):
    """{tag: array [n, 4]} plus the trailing scalars."""
    out, scal = {}, {}
    lines = text.strip().splitlines()
    k = 0
    while k < len(lines):
        parts = lines[k].split()
        if parts[0] in ("init", "action", "lf", "beam"):
            n = int(parts[1])
            out[parts[0]] = np.array([[float(v) for v in ln.split()] for ln in lines[k + 1:k + 1 + n]])
            k += 1 + n
        else:
            scal[parts[0]] = [float(v) for v in parts[1:] if v.replace(".", "").replace("-", "").replace("e", "").replace("+", "").isdigit()]
            k += 1
    return out, scal


def test_native_cxx_caller_links_and_matches_reference(gpu_lib, tmp_path):
    if not os.path.exists(REF_EXE):
        pytest.skip("oracle/_ref/amcl_node_probe_ref not built (reference tree not mounted)")
    exe = str(tmp_path / "amcl_node_probe_b200")
    libdir = os.path.dirname(library.LIB_PATH)
    build = subprocess.run(["g++", "-O1", "-std=c++11", "-Wall", "-I" + os.path.join(ROOT, "include"), SRC,
                            "-o", exe, "-L" + libdir, "-lamcl_b200", "-Wl,-rpath," + libdir],
                           capture_output=True, text=True)
    assert build.returncode == 0, build.stderr
    mine = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert mine.returncode == 0, mine.stderr[-2000:]
    want = subprocess.run([REF_EXE], capture_output=True, text=True, timeout=300)
    assert want.returncode == 0, want.stderr[-2000:]
    a, sa = parse(mine.stdout)
    b, sb = parse(want.stdout)
    assert sa["clusters"] == sb["clusters"]
    assert sa["ok"][0] == 1.0 and sa["ok"] == sb["ok"]      # incl. two occ_dist cells written back into map->cells
    assert np.array_equal(a["init"], b["init"])              # callback poses, 1/n weights
    for tag in ("action", "lf", "beam"):
        assert a[tag].shape == b[tag].shape
        # zero-noise odometry: poses agree to rounding of the device sin/cos
        assert np.allclose(a[tag][:, :3], b[tag][:, :3], rtol=1e-12, atol=1e-12)
        rel = np.abs(a[tag][:, 3] - b[tag][:, 3]) / b[tag][:, 3]
        assert rel.max() < 1e-5, (tag, rel.max())
