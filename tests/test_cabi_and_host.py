"""CPU-side checks of the product: the C-ABI library loads and exports every
symbol include/daala_tdlt.h declares (no compute is called here -- there is no
GPU in this container), the generated lifting code (the very header the kernels
compile) agrees with the reference-pinned oracle when compiled for the host, the
post-filter's multiply-high division is exact, and nothing under daala_b200/
reaches into oracle/."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import oracle_libs as ol
from oracle_libs import ptr

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "daala_tdlt.h")
CSRC = os.path.join(ROOT, "daala_b200", "csrc")


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"\b(od_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def test_library_exports_every_declared_symbol():
    import daala_b200
    L = daala_b200.lib()
    names = _declared_functions()
    assert len(names) >= 30
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing


def test_vtbl_installer_fills_the_dispatch_tables():
    """od_tdlt_opt_vtbl_init_cuda writes the CUDA entry points into tables shaped like
    od_state.opt_vtbl.fdct_2d / idct_2d (reference src/state.h:123-124) and leaves the
    64-point slot alone.  No kernel is launched."""
    import daala_b200
    L = daala_b200.lib()
    fd = (C.c_void_p * 5)(*([0xdead] * 5))
    iv = (C.c_void_p * 5)(*([0xbeef] * 5))
    L.od_tdlt_opt_vtbl_init_cuda.argtypes = [C.c_void_p, C.c_void_p]
    L.od_tdlt_opt_vtbl_init_cuda.restype = None
    L.od_tdlt_opt_vtbl_init_cuda(fd, iv)
    for k, n in enumerate((4, 8, 16, 32)):
        assert fd[k] == C.cast(getattr(L, "od_bin_fdct%dx%d_cuda" % (n, n)), C.c_void_p).value
        assert iv[k] == C.cast(getattr(L, "od_bin_idct%dx%d_cuda" % (n, n)), C.c_void_p).value
    assert fd[4] == 0xdead and iv[4] == 0xbeef


def test_context_calls_fail_loudly_without_a_device():
    """No GPU here: creating a context must report an error, not fall back."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import daala_b200
    with pytest.raises(daala_b200.TdltError):
        daala_b200.TdltContext(352, 288)


def test_product_does_not_touch_the_oracle():
    pat = re.compile(r"oracle_libs|libtdlt_oracle|libdaala_ref|tdlt_oracle\.h|ref_glue")
    bad = []
    for dp, _, files in os.walk(os.path.join(ROOT, "daala_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                if pat.search(open(os.path.join(dp, f), errors="replace").read()):
                    bad.append(f)
    assert not bad, bad


HOST_SHIM = r"""
#include "tdlt_lift_gen.cuh"
extern "C" {
void host_dct(int bs, int inverse, int *v) {
  if (!inverse) { if (bs == 0) tdlt_fdct4(v); else if (bs == 1) tdlt_fdct8(v);
                  else if (bs == 2) tdlt_fdct16(v); else tdlt_fdct32(v); }
  else { if (bs == 0) tdlt_idct4(v); else if (bs == 1) tdlt_idct8(v);
         else if (bs == 2) tdlt_idct16(v); else tdlt_idct32(v); }
}
static inline int mulhi(int a, int b) { return (int)(((long long)a*b) >> 32); }
long long post_div_mismatches(int lim) {
  long long bad = 0;
  for (int t = -lim + 1; t < lim; t++) {
    int a = mulhi(t, (int)3665038760u) + t + (int)((unsigned)t >> 31);
    int b = mulhi(t, (int)3233857729u) + t + (int)((unsigned)t >> 31);
    bad += (a != (t*64)/75) + (b != (t*64)/85);
  }
  return bad;
}
}
"""


@pytest.fixture(scope="module")
def hostlib(tmp_path_factory):
    d = tmp_path_factory.mktemp("hostlift")
    src = d / "shim.cpp"
    src.write_text(HOST_SHIM)
    so = d / "libhostlift.so"
    subprocess.check_call(["g++", "-O1", "-fPIC", "-shared", "-I", CSRC, "-o", str(so), str(src)])
    return C.CDLL(str(so))


@pytest.mark.parametrize("bs", [0, 1, 2, 3])
def test_generated_lifting_code_matches_oracle(hostlib, bs):
    """daala_b200/csrc/tdlt_lift_gen.cuh is what the kernels compile; built for the
    host it must give the oracle's (= the reference's) result for every vector,
    including the d -= x  ==  d += -x - 1 rewriting of the subtracting lifts."""
    P = ol.port()
    n = 4 << bs
    rng = np.random.default_rng(2000 + bs)
    X = np.concatenate([rng.integers(-32768, 32768, size=(3000, n)),
                        rng.integers(-5, 6, size=(500, n)),
                        np.where(rng.integers(0, 2, size=(500, n)) > 0, 32767, -32768)]).astype(np.int32)
    for x in X:
        x = np.ascontiguousarray(x)
        for inverse in (0, 1):
            v = x.copy()
            hostlib.host_dct(bs, inverse, ptr(v))
            y = np.zeros(n, np.int32)
            if inverse:
                P.oc_idct1d(bs, ptr(y), 1, ptr(x))
            else:
                P.oc_fdct1d(bs, ptr(y), ptr(x), 1)
            assert np.array_equal(v, y)


def test_post_filter_division_is_exact(hostlib):
    """(t*64)/75 and (t*64)/85 (reference src/filter.c:233-236, C division) as one
    multiply-high each: exact over the whole range in which t*64 fits an int."""
    hostlib.post_div_mismatches.restype = C.c_longlong
    assert hostlib.post_div_mismatches(1 << 25) == 0


@pytest.mark.parametrize("size", [(3840, 2160), (1920, 1080), (352, 288), (100, 60), (33, 17)])
@pytest.mark.parametrize("rows", [1, 3, 8, 24, 1000])
def test_inverse_region_order_keeps_every_dependency_ahead(size, rows):
    """The inverse kernel's regions wait for the strips of their left / upper / upper-left
    neighbours and draw tickets in the order of this table (frame-interleaved), so the
    table must be a permutation of a frame's regions in which every region comes after
    those three -- whatever the skew (rows of regions in flight)."""
    import daala_b200
    pic_w, pic_h = size
    decs = (0, 1, 1)
    order = daala_b200.region_order(pic_w, pic_h, decs, rows)
    fw, fh = (pic_w + 63) & ~63, (pic_h + 63) & ~63
    planes, reg0 = [], 0
    for d in decs:
        nrx, nry = ((fw >> d) + 127)//128, ((fh >> d) + 63)//64
        planes.append((reg0, nrx, nry))
        reg0 += nrx*nry
    assert sorted(order.tolist()) == list(range(reg0))
    pos = np.empty(reg0, np.int64)
    pos[order] = np.arange(reg0)
    for r0, nrx, nry in planes:
        for ry in range(nry):
            for rx in range(nrx):
                me = pos[r0 + ry*nrx + rx]
                if rx > 0:
                    assert pos[r0 + ry*nrx + rx - 1] < me
                if ry > 0:
                    assert pos[r0 + (ry - 1)*nrx + rx] < me
                if rx > 0 and ry > 0:
                    assert pos[r0 + (ry - 1)*nrx + rx - 1] < me
    if rows == 1:         # raster order within each plane
        assert order.tolist() == list(range(reg0))
