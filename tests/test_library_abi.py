"""CPU-only checks of the drop-in boundary: the shared library loads without a
GPU, exports every symbol the headers declare, and the ABI headers compile as
C and as C++ with the reference's struct sizes."""
import ctypes as C
import os
import subprocess
import tempfile

import pytest

from navigation_b200 import abi, library

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INC = os.path.join(ROOT, "include")


@pytest.fixture(scope="module")
def so():
    if not os.path.exists(library.LIB_PATH):
        library.build()
    return C.CDLL(library.LIB_PATH)


def test_exports_every_declared_symbol(so):
    missing = [s for s in library.declared_symbols() if not hasattr(so, s)]
    assert missing == []


def test_exports_reference_cxx_symbols():
    out = subprocess.run(["nm", "-D", "--defined-only", "-C", library.LIB_PATH],
                         capture_output=True, text=True, check=True).stdout
    for sym in [
        "amcl::AMCLSensor::AMCLSensor()", "amcl::AMCLSensor::~AMCLSensor()",
        "amcl::AMCLSensor::UpdateAction(", "amcl::AMCLSensor::InitSensor(",
        "amcl::AMCLSensor::UpdateSensor(", "vtable for amcl::AMCLSensor",
        "typeinfo for amcl::AMCLSensor", "amcl::AMCLOdom::AMCLOdom()",
        "amcl::AMCLOdom::SetModelDiff(", "amcl::AMCLOdom::SetModelOmni(",
        "amcl::AMCLOdom::SetModel(", "amcl::AMCLOdom::UpdateAction(",
        "vtable for amcl::AMCLOdom", "amcl::AMCLLaser::AMCLLaser(unsigned long, map_t*)",
        "amcl::AMCLLaser::~AMCLLaser()", "amcl::AMCLLaser::SetModelBeam(",
        "amcl::AMCLLaser::SetModelLikelihoodField(", "amcl::AMCLLaser::SetModelLikelihoodFieldProb(",
        "amcl::AMCLLaser::UpdateSensor(", "amcl::AMCLLaser::BeamModel(",
        "amcl::AMCLLaser::LikelihoodFieldModel(", "amcl::AMCLLaser::LikelihoodFieldModelProb(",
        "amcl::AMCLLaser::reallocTempData(", "vtable for amcl::AMCLLaser",
    ]:
        assert sym in out, sym


def test_struct_sizes_without_gpu(so):
    abi.bind(so, tables=(abi.FACADE_SYMBOLS,))
    assert so.amcl_c_sizeof_pf() == 408
    assert so.amcl_c_sizeof_laser() == 192
    assert so.amcl_c_sizeof_odom() == 96


def test_version_string(so):
    so.amclb200_version.restype = C.c_char_p
    assert b"sm_100a" in so.amclb200_version()


def test_host_helpers_need_no_gpu(so):
    """pf_vector / eigen helpers are host code and must run on a CPU-only box."""
    abi.bind(so, tables=(abi.REFERENCE_SYMBOLS,))
    a = abi.pf_vector_t()
    a.v[:] = [1.0, 0.0, 0.5]
    b = abi.pf_vector_t()
    b.v[:] = [2.0, 3.0, 1.0]
    c = so.pf_vector_coord_add(a, b)
    import math
    assert abs(c.v[0] - (2.0 + math.cos(1.0))) < 1e-15
    assert abs(c.v[2] - 1.5) < 1e-15


C_PROBE = r"""
#include <math.h>
#include "amcl/map/map.h"
#include "amcl/pf/pf.h"
#include "amcl/pf/pf_pdf.h"
#include "amcl/pf/pf_kdtree.h"
#include "amcl/pf/pf_vector.h"
#include "amcl/pf/eig3.h"
_Static_assert(sizeof(pf_t) == 408, "pf_t");
_Static_assert(sizeof(pf_sample_set_t) == 152, "set");
_Static_assert(sizeof(pf_sample_t) == 32, "sample");
_Static_assert(sizeof(pf_cluster_t) == 176, "cluster");
_Static_assert(sizeof(pf_kdtree_node_t) == 72, "node");
_Static_assert(sizeof(map_t) == 48, "map");
_Static_assert(sizeof(map_cell_t) == 16, "cell");
int probe(map_t *map) { return MAP_VALID(map, (int)MAP_GXWX(map, 0.0), 0) ? MAP_INDEX(map, 1, 2) : 0; }
"""

CXX_PROBE = r"""
#include "amcl/map/map.h"
#include "amcl/pf/pf.h"
#include "amcl/sensors/amcl_odom.h"
#include "amcl/sensors/amcl_laser.h"
using namespace amcl;
static_assert(sizeof(AMCLSensor) == 40, "");
static_assert(sizeof(AMCLOdom) == 96, "");
static_assert(sizeof(AMCLOdomData) == 72, "");
static_assert(sizeof(AMCLLaser) == 192, "");
static_assert(sizeof(AMCLLaserData) == 48, "");
// what amcl_node.cpp does with these types (amcl_node.cpp:934-958, 1143-1171, 1257-1329)
bool drive(pf_t *pf, map_t *map) {
  AMCLOdom *odom = new AMCLOdom();
  odom->SetModel(ODOM_MODEL_DIFF, 0.2, 0.2, 0.2, 0.2, 0.2);
  AMCLLaser *laser = new AMCLLaser(30, map);
  laser->SetModelLikelihoodField(0.95, 0.05, 0.2, 2.0);
  AMCLLaser *copy = new AMCLLaser(*laser);
  pf_vector_t lp = pf_vector_zero();
  copy->SetLaserPose(lp);
  AMCLLaserData ldata;
  ldata.sensor = copy;
  ldata.range_count = 2;
  ldata.range_max = 10.0;
  ldata.ranges = new double[2][2];
  AMCLOdomData odata;
  odata.pose = lp; odata.delta = lp;
  bool ok = odom->UpdateAction(pf, (AMCLSensorData*)&odata);
  ok = ok && copy->UpdateSensor(pf, (AMCLSensorData*)&ldata);
  pf->pop_err = 0.01; pf->pop_z = 0.99;
  return ok && pf->sets[pf->current_set].samples[0].pose.v[0] == 0.0;
}
"""


def _compile(src, lang, flags):
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "probe." + ("c" if lang == "c" else "cpp"))
        with open(path, "w") as f:
            f.write(src)
        cc = "gcc" if lang == "c" else "g++"
        out = subprocess.run([cc, "-c", path, "-o", os.path.join(d, "probe.o"), "-I" + INC] + flags,
                             capture_output=True, text=True)
        assert out.returncode == 0, out.stderr


def test_headers_compile_as_c():
    _compile(C_PROBE, "c", ["-std=c11", "-Wall", "-Werror"])


def test_headers_compile_as_cxx_with_node_usage():
    _compile(CXX_PROBE, "cxx", ["-std=c++11", "-Wall"])


@pytest.mark.skipif(not os.path.exists("/root/reference/amcl/include/amcl/pf/pf.h"),
                    reason="reference tree not mounted")
def test_same_probes_compile_against_reference_headers():
    """Source compatibility: identical caller code builds against both header sets."""
    ref_inc = ["-I/root/reference/amcl/include", "-I/root/reference/amcl/src/include"]
    for src, lang, flags in ((CXX_PROBE, "cxx", ["-std=c++11"]),):
        with tempfile.TemporaryDirectory() as d:
            path = os.path.join(d, "probe.cpp")
            with open(path, "w") as f:
                f.write(src)
            out = subprocess.run(["g++", "-c", path, "-o", os.path.join(d, "p.o")] + ref_inc + flags,
                                 capture_output=True, text=True)
            assert out.returncode == 0, out.stderr


# ---- map_load_occ (map_store.c:39-136): host code, compared with the compiled reference ----

def _write_pgm(path, pixels, depth, comment=True):
    h, w = pixels.shape
    with open(path, "wb") as f:
        f.write(b"P5\n")
        if comment:
            f.write(b"# made by tests/test_library_abi.py\n# second comment line\n")
        f.write(b"%d %d\n%d\n" % (w, h, depth))
        f.write(pixels.astype("uint8").tobytes())


def _load_states(lib, path, scale, negate, prealloc=None):
    import numpy as np

    lib.map_alloc.restype = abi.map_p
    lib.map_load_occ.argtypes = [abi.map_p, C.c_char_p, C.c_double, C.c_int]
    lib.map_load_occ.restype = C.c_int
    lib.map_free.argtypes = [abi.map_p]
    m = lib.map_alloc()
    if prealloc is not None:
        libc = C.CDLL(None)
        libc.calloc.restype = C.c_void_p
        libc.calloc.argtypes = [C.c_size_t, C.c_size_t]
        m.contents.size_x, m.contents.size_y = prealloc
        m.contents.cells = C.cast(libc.calloc(prealloc[0] * prealloc[1], 16), C.POINTER(abi.map_cell_t))
    rc = lib.map_load_occ(m, path.encode(), scale, negate)
    out = None
    if rc == 0:
        sx, sy = m.contents.size_x, m.contents.size_y
        raw = np.ctypeslib.as_array(C.cast(m.contents.cells, C.POINTER(C.c_int32)), shape=(sy, sx, 4))
        out = (raw[:, :, 0].copy(), sx, sy, m.contents.scale)
    lib.map_free(m)
    return rc, out


@pytest.mark.parametrize("depth,negate", [(255, 0), (255, 1), (100, 0), (15, 1)])
def test_map_load_occ_matches_reference(so, ref, tmp_path, depth, negate):
    import numpy as np

    rng = np.random.default_rng(depth + negate)
    px = rng.integers(0, depth + 1, size=(37, 53))
    px[0, :] = 0                     # top image row = last map row
    px[:, 0] = depth
    path = str(tmp_path / "m.pgm")
    _write_pgm(path, px, depth)
    rc_a, a = _load_states(ref, path, 0.05, negate)
    rc_b, b = _load_states(so, path, 0.05, negate)
    assert rc_a == rc_b == 0
    assert a[1:] == b[1:] == (53, 37, 0.05)
    assert np.array_equal(a[0], b[0])
    assert set(np.unique(b[0]).tolist()) <= {-1, 0, 1}
    # a map that already has cells must match the file's size (map_store.c:85-92)
    assert _load_states(ref, path, 0.05, negate, prealloc=(53, 36))[0] == -1
    assert _load_states(so, path, 0.05, negate, prealloc=(53, 36))[0] == -1
    rc_c, c = _load_states(so, path, 0.05, negate, prealloc=(53, 37))
    assert rc_c == 0 and np.array_equal(c[0], a[0])


def test_map_load_occ_rejects_what_the_reference_rejects(so, ref, tmp_path):
    bad = str(tmp_path / "bad.pgm")
    with open(bad, "wb") as f:
        f.write(b"P2\n2 2\n255\n0 0 0 0\n")
    for lib in (ref, so):
        assert _load_states(lib, bad, 0.05, 0)[0] == -1
        assert _load_states(lib, str(tmp_path / "missing.pgm"), 0.05, 0)[0] == -1
