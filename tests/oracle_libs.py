"""ctypes bindings for the two CHECKERS (test infrastructure only):

* ``port``  -- oracle/libtdlt_oracle.so, the plain-C restatement.
* ``ref``   -- oracle/_ref/libdaala_ref.so, the unmodified reference compiled
               from /root/reference by oracle/Makefile (present wherever
               ``__graft_entry__.build()`` ran with the reference mounted; it
               travels to the GPU box as a prebuilt file).

Nothing under daala_b200/ imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
PORT_SO = os.path.join(ORACLE_DIR, "libtdlt_oracle.so")
# DAALA_REF_LIB: another build of the same reference sources (bench.py's cpu_baseline times the
# -O3 -march=native one in a subprocess); the tests never set it.
REF_SO = os.environ.get("DAALA_REF_LIB") or os.path.join(ORACLE_DIR, "_ref", "libdaala_ref.so")

I32P = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
U8P = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")


def _ensure_port():
    if not os.path.exists(PORT_SO) or (
            os.path.getmtime(PORT_SO) < os.path.getmtime(os.path.join(ORACLE_DIR, "tdlt_oracle.c"))):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "port"], stdout=subprocess.DEVNULL)


_port = None
_ref = None


def port():
    """The C restatement (always buildable: needs only gcc)."""
    global _port
    if _port is None:
        _ensure_port()
        _port = C.CDLL(PORT_SO)
    return _port


def have_ref():
    if os.path.exists(REF_SO):
        return True
    if os.path.isdir("/root/reference/src"):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "ref", "-j8"], stdout=subprocess.DEVNULL)
        return os.path.exists(REF_SO)
    return False


def ref():
    """The compiled reference (None-safe: call have_ref() first)."""
    global _ref
    if _ref is None:
        assert have_ref(), "oracle/_ref/libdaala_ref.so is not built"
        _ref = C.CDLL(REF_SO)
        _ref.ref_ctx_create.restype = C.c_void_p
        _ref.ref_ctx_create.argtypes = [C.c_int, C.c_int]
        _ref.ref_ctx_free.argtypes = [C.c_void_p]
        for f in ("ref_ctx_frame_width", "ref_ctx_frame_height", "ref_ctx_nhsb", "ref_ctx_nvsb"):
            getattr(_ref, f).argtypes = [C.c_void_p]
    return _ref


def ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def padded_dims(pic_w, pic_h):
    """Frame dimensions: picture padded to a multiple of 64 (reference
    src/state.c:372-375)."""
    return (pic_w + 63) & ~63, (pic_h + 63) & ~63


# ---------------------------------------------------------------- port API --
def port_fwd_plane(c, bsize, dec, pic_w, pic_h, keyframe, use_haar=0):
    """c: (H, W) int32 spatial plane.  Returns (d, lapped c)."""
    lib = port()
    c = np.ascontiguousarray(c, dtype=np.int32).copy()
    d = np.zeros_like(c)
    H, W = c.shape
    nhsb, nvsb = (W << dec) // 32, (H << dec) // 32
    bsize = np.ascontiguousarray(bsize, dtype=np.uint8)
    lib.oc_tdlt_fwd_plane(ptr(d), ptr(c), nhsb, nvsb, dec, ptr(bsize), bsize.shape[1],
                          pic_w, pic_h, int(keyframe), int(use_haar))
    return d, c


def port_inv_plane(d, bsize, dec, pic_w, pic_h, dc_split, use_haar=0):
    lib = port()
    d = np.ascontiguousarray(d, dtype=np.int32).copy()
    c = np.zeros_like(d)
    H, W = d.shape
    nhsb, nvsb = (W << dec) // 32, (H << dec) // 32
    bsize = np.ascontiguousarray(bsize, dtype=np.uint8)
    lib.oc_tdlt_inv_plane(ptr(c), ptr(d), nhsb, nvsb, dec, ptr(bsize), bsize.shape[1],
                          pic_w, pic_h, int(dc_split), int(use_haar))
    return c


def port_cfl_pred_plane(d_luma, bsize):
    """Chroma-from-luma predictor plane of one frame (restatement)."""
    lib = port()
    d_luma = np.ascontiguousarray(d_luma, dtype=np.int32)
    H, W = d_luma.shape
    bsize = np.ascontiguousarray(bsize, dtype=np.uint8)
    l = np.zeros((H // 2, W // 2), np.int32)
    lib.oc_cfl_pred_plane(ptr(l), ptr(d_luma), W // 32, H // 32, ptr(bsize), bsize.shape[1])
    return l


def ref_cfl_pred_plane(d_luma, bsize):
    """The same plane from the REFERENCE's od_resample_luma_coeffs (src/intra.c:72), called
    once per chroma leaf block with the arguments of its call sites (src/encode.c:1512-1517):
    luma origin (bx << (2 + bsi), by << (2 + bsi)), xdec = ydec = 1, bs = max(obs, 1) - 1,
    chroma_bs = obs."""
    lib = ref()
    d_luma = np.ascontiguousarray(d_luma, dtype=np.int32)
    H, W = d_luma.shape
    cw = W // 2
    l = np.zeros((H // 2, cw), np.int32)
    f = lib.od_resample_luma_coeffs
    f.restype = None
    f.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
    for uy in range(H // 8):
        for ux in range(W // 8):
            obs = int(bsize[uy, ux])
            bsi = max(obs, 1)
            m = 1 << (bsi - 1)
            if (ux & (m - 1)) or (uy & (m - 1)):
                continue
            f(l.ctypes.data + 4 * ((uy * 4) * cw + ux * 4), cw,
              d_luma.ctypes.data + 4 * ((uy * 8) * W + ux * 8), W, 1, 1, bsi - 1, obs)
    return l


def dering_args(planes, decs, bskip):
    """ctypes arrays for the frame-level deringing calls."""
    n = len(planes)
    pa = (C.c_void_p * 4)(*[p.ctypes.data for p in planes])
    da = (C.c_int * 4)(*decs)
    ba = (C.c_void_p * 4)(*[b.ctypes.data for b in bskip])
    return n, pa, da, ba


def port_dering_frame(planes, decs, flags, bskip, q):
    """planes[p]: (H >> dec, W >> dec) int32 reconstructions; flags: (nvsb, nhsb) uint8, one per
    32x32 luma superblock; bskip[p]: (nvsb*8, nhsb*8) uint8 skip flags of plane p (one per 4x4
    samples; chroma planes use the upper-left quarter); q[p]: quantizers.  Returns filtered planes."""
    lib = port()
    planes = [np.ascontiguousarray(p, dtype=np.int32) for p in planes]
    out = [np.zeros_like(p) for p in planes]
    H, W = planes[0].shape
    nhsb, nvsb = W // 32, H // 32
    flags = np.ascontiguousarray(flags, dtype=np.uint8)
    bskip = [np.ascontiguousarray(b, dtype=np.uint8) for b in bskip]
    n, pa, da, ba = dering_args(planes, decs, bskip)
    oa = (C.c_void_p * 4)(*[o.ctypes.data for o in out])
    qa = (C.c_int * 4)(*q)
    lib.oc_dering_frame(oa, pa, n, da, nhsb, nvsb, ptr(flags), ba, bskip[0].shape[1], qa)
    return out


def ref_dering_frame(rc, planes, decs, flags, bskip, q, impl=1):
    """The same frame through the REFERENCE's od_dering, one superblock and plane at a time, in
    the decoder's loop order (src/decode.c:1055-1122): etmp = int16 copy, output into ctmp."""
    lib = rc.lib
    planes = [np.ascontiguousarray(p, dtype=np.int32) for p in planes]
    e = [p.astype(np.int16) for p in planes]
    out = [p.copy() for p in planes]
    H, W = planes[0].shape
    nhsb, nvsb = W // 32, H // 32
    bskip = [np.ascontiguousarray(b, dtype=np.uint8) for b in bskip]
    ss = bskip[0].shape[1]
    for sby in range(nvsb):
        for sbx in range(nhsb):
            if not flags[sby, sbx]:
                continue
            dirs = np.zeros((4, 4), np.int32)
            for p, dec in enumerate(decs):
                ln = 5 - dec
                n = 1 << ln
                w = W >> dec
                buf = np.zeros((n, n), np.int16)
                lib.ref_dering(C.c_void_p(rc.h), ptr(buf), n,
                               C.c_void_p(e[p].ctypes.data + 2 * ((sby << ln) * w + (sbx << ln))), w, ln,
                               sbx, sby, nhsb, nvsb, int(q[p]), dec, ptr(dirs), p,
                               C.c_void_p(bskip[p].ctypes.data + (sby << (3 - dec)) * ss + (sbx << (3 - dec))),
                               ss, impl)
                out[p][sby << ln:(sby + 1) << ln, sbx << ln:(sbx + 1) << ln] = buf
    return out


def ref_dering_block(rc, impl, bsi, y1, y2, ystride, inp, org, ostride, thr, d):
    """One block through the REFERENCE's deringing filters: impl 0 = the C functions
    (od_filter_dering_*_c, src/filter.c:1715-1789), 1 = what a fresh state's dispatch table
    holds (the SSE2 twins, src/x86/x86state.c).  y1 <- direction filter, y2 <- orthogonal."""
    lib = rc.lib
    n = 4 << bsi
    if impl == 0:
        fd = getattr(lib, "od_filter_dering_direction_%dx%d_c" % (n, n))
        fo = getattr(lib, "od_filter_dering_orthogonal_%dx%d_c" % (n, n))
        fd(ptr(y1), ystride, C.c_void_p(inp), thr, d)
        fo(ptr(y2), ystride, C.c_void_p(inp), ptr(org), ostride, thr, d)
    else:
        lib.ref_dering_direction(C.c_void_p(rc.h), bsi, ptr(y1), ystride, C.c_void_p(inp), thr, d)
        lib.ref_dering_orthogonal(C.c_void_p(rc.h), bsi, ptr(y2), ystride, C.c_void_p(inp),
                                  ptr(org), ostride, thr, d)


def dering_test_frame(rng, nhsb, nvsb, decs=(0, 1, 1), skip_prob=0.5, flag_prob=0.7):
    """A frame that exercises the filter: oriented edges (so that the direction search has
    something to find) plus noise, skip flags in patches, superblock decisions."""
    H, W = nvsb * 32, nhsb * 32
    yy, xx = np.mgrid[0:H, 0:W]
    ang = rng.uniform(0, np.pi, size=(nvsb, nhsb)).repeat(32, 0).repeat(32, 1)
    per = rng.uniform(6, 40)
    luma = 900 * np.sign(np.sin((xx * np.cos(ang) + yy * np.sin(ang)) * 2 * np.pi / per))
    luma = luma + rng.normal(0, 40, size=(H, W))
    planes = [np.clip(luma, -2040, 2040).astype(np.int32)]      # 8-bit video in od_coeff scale
    for dec in decs[1:]:
        c = planes[0][::1 << dec, ::1 << dec] // 2 + rng.integers(-300, 300, size=(H >> dec, W >> dec))
        planes.append(np.ascontiguousarray(c).astype(np.int32))
    flags = (rng.random((nvsb, nhsb)) < flag_prob).astype(np.uint8)
    bskip = []
    for dec in decs:
        coarse = rng.random((nvsb * 2, nhsb * 2)) < skip_prob
        b = np.zeros((nvsb * 8, nhsb * 8), np.uint8)
        fine = coarse.repeat(4 >> dec, 0).repeat(4 >> dec, 1) & (rng.random((nvsb * 8 >> dec, nhsb * 8 >> dec)) < 0.9)
        b[:nvsb * 8 >> dec, :nhsb * 8 >> dec] = fine
        bskip.append(b)
    return planes, flags, bskip


def port_smooth_frame(planes, decs, bsize, q):
    """Keyframe bilinear smoothing of the 32x32-coded superblocks (restatement); returns new planes."""
    lib = port()
    out = [np.ascontiguousarray(p, dtype=np.int32).copy() for p in planes]
    H, W = out[0].shape
    bsize = np.ascontiguousarray(bsize, dtype=np.uint8)
    pa = (C.c_void_p * 4)(*[o.ctypes.data for o in out])
    lib.oc_smooth_frame(pa, len(out), (C.c_int * 4)(*decs), W // 32, H // 32, ptr(bsize), bsize.shape[1],
                        (C.c_int * 4)(*q))
    return out


def ref_smooth_frame(planes, decs, bsize, q):
    """The same through the REFERENCE's od_smooth_recursive, called per superblock and plane
    with the decoder's arguments (src/decode.c:1139-1143)."""
    lib = ref()
    out = [np.ascontiguousarray(p, dtype=np.int32).copy() for p in planes]
    H, W = out[0].shape
    bsize = np.ascontiguousarray(bsize, dtype=np.uint8)
    f = lib.od_smooth_recursive
    f.restype = None
    f.argtypes = [C.c_void_p, C.c_void_p] + [C.c_int] * 10
    for p, dec in enumerate(decs):
        for sby in range(H // 32):
            for sbx in range(W // 32):
                f(ptr(out[p]), ptr(bsize), bsize.shape[1], sbx, sby, 3, W >> dec, dec, dec, 3, int(q[p]), p)
    return out


SCAN_TABLE_LEN = (15, 48, 192, 768)      # entries of OD_ZIGZAG4/8/16/32
# how many of them the layouts order: OD_LAYOUT*.band_offsets[nb_bands] (src/partition.c:43-73);
# the 32x32 layout stops after two bands of 128
SCAN_LEN = (15, 48, 192, 256)


def scan_tables():
    """The reference's scan tables OD_ZIGZAG4/8/16/32 (src/zigzag*.c) as four (len, 2) uint8 arrays
    of (x, y): read from the compiled reference when it is here, else from the golden fixture
    tests/golden/make_golden.py took from it."""
    if have_ref():
        lib = ref()
        return [np.ctypeslib.as_array((C.c_ubyte * (2 * n)).in_dll(lib, "OD_ZIGZAG%d" % (4 << k))).reshape(n, 2).copy()
                for k, n in enumerate(SCAN_TABLE_LEN)]
    G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "tdlt_golden.npz"))
    return [G["scan_table_%d" % (4 << k)] for k in range(4)]


def _scan_arg(tables):
    tables = [np.ascontiguousarray(t, dtype=np.uint8) for t in tables]
    return tables, (C.c_void_p * 4)(*[t.ctypes.data for t in tables]), (C.c_int * 4)(*SCAN_LEN)


def port_coding_order_plane(src, bsize, dec, tables, to_coding=True, out=None):
    """to_coding: raster plane d -> block-linear plane v of every block's coding-order vector;
    else the inverse.  Entries the scans do not cover (half of every 32x32 block) keep what `out`
    held (zeros when it is not given).  Returns the produced plane."""
    lib = port()
    src = np.ascontiguousarray(src, dtype=np.int32).copy()
    out = np.zeros_like(src) if out is None else out
    H, W = src.shape
    bsize = np.ascontiguousarray(bsize, dtype=np.uint8)
    keep, sa, la = _scan_arg(tables)
    v, d = (out, src) if to_coding else (src, out)
    lib.oc_coding_order_plane(ptr(v), ptr(d), (W << dec) // 32, (H << dec) // 32, dec, ptr(bsize),
                              bsize.shape[1], sa, la, 1 if to_coding else 0)
    return out


def ref_coding_order_plane(src, bsize, dec, to_coding=True, out=None):
    """The same planes from the REFERENCE's od_raster_to_coding_order / od_coding_order_to_raster
    (src/partition.c:144-200) called block by block over the map, as its call sites do
    (src/encode.c:1160-1161, src/decode.c:630); only the block-linear placement of the vectors
    (oc_block_linear_offset, this library's own layout) comes from the port."""
    R, P = ref(), port()
    R.od_raster_to_coding_order.restype = None
    R.od_coding_order_to_raster.restype = None
    P.oc_block_linear_offset.restype = C.c_longlong
    src = np.ascontiguousarray(src, dtype=np.int32)
    out = np.zeros_like(src) if out is None else out
    H, W = src.shape
    flat_in, flat_out = src.ravel(), out.reshape(-1)
    for uy in range(bsize.shape[0]):
        for ux in range(bsize.shape[1]):
            obs = int(bsize[uy, ux])
            lu = max(obs - 1, 0)
            if (ux & ((1 << lu) - 1)) or (uy & ((1 << lu) - 1)):
                continue
            n = max(4, (4 << obs) >> dec)
            xu, yu = (ux * 8) >> dec, (uy * 8) >> dec
            # an 8x8 luma unit coded as 4x4 blocks holds four of them (one in 4:2:0 chroma)
            subs = ((0, 0), (4, 0), (0, 4), (4, 4)) if (obs == 0 and dec == 0) else ((0, 0),)
            for sx, sy in subs:
              x0, y0 = xu + sx, yu + sy
              off = P.oc_block_linear_offset(x0, y0, W, 32 >> dec)
              if to_coding:
                  blk = np.ascontiguousarray(src[y0:y0 + n, x0:x0 + n])
                  vec = np.ascontiguousarray(flat_out[off:off + n * n])
                  ncoded = min(n * n, 512)
                  tmp = np.zeros(n * n, np.int32)
                  R.od_raster_to_coding_order(ptr(tmp), n, ptr(blk), n)
                  vec[:ncoded] = tmp[:ncoded]
                  flat_out[off:off + n * n] = vec
              else:
                  vec = np.ascontiguousarray(flat_in[off:off + n * n])
                  blk = np.ascontiguousarray(out[y0:y0 + n, x0:x0 + n])
                  R.od_coding_order_to_raster(ptr(blk), n, ptr(vec), n)
                  out[y0:y0 + n, x0:x0 + n] = blk
    return out


# ----------------------------------------------------------------- ref API --
class RefCtx:
    """An instance of the reference encoder state for one picture size."""

    def __init__(self, pic_w, pic_h):
        self.lib = ref()
        self.h = self.lib.ref_ctx_create(pic_w, pic_h)
        assert self.h, "daala_encode_create failed"
        self.pic_w, self.pic_h = pic_w, pic_h
        self.fw = self.lib.ref_ctx_frame_width(self.h)
        self.fh = self.lib.ref_ctx_frame_height(self.h)
        self.nhsb = self.lib.ref_ctx_nhsb(self.h)
        self.nvsb = self.lib.ref_ctx_nvsb(self.h)

    def close(self):
        if self.h:
            self.lib.ref_ctx_free(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def fwd_plane(self, c, bsize, dec, keyframe, use_haar=0, impl=0):
        c = np.ascontiguousarray(c, dtype=np.int32).copy()
        assert c.shape == (self.fh >> dec, self.fw >> dec)
        d = np.zeros_like(c)
        bsize = np.ascontiguousarray(bsize, dtype=np.uint8)
        self.lib.ref_tdlt_fwd_plane(C.c_void_p(self.h), ptr(d), ptr(c), dec, ptr(bsize),
                                    bsize.shape[1], int(keyframe), int(use_haar), impl)
        return d, c

    def inv_plane(self, d, bsize, dec, dc_split, use_haar=0, impl=0):
        d = np.ascontiguousarray(d, dtype=np.int32).copy()
        c = np.zeros_like(d)
        bsize = np.ascontiguousarray(bsize, dtype=np.uint8)
        self.lib.ref_tdlt_inv_plane(C.c_void_p(self.h), ptr(c), ptr(d), dec, ptr(bsize),
                                    bsize.shape[1], int(dc_split), int(use_haar), impl)
        return c

    def encode_keyframe_bsize(self, y, u, v, quality=20, complexity=7):
        out = np.zeros((self.nvsb * 4, self.nhsb * 4), dtype=np.uint8)
        y, u, v = (np.ascontiguousarray(p, dtype=np.uint8) for p in (y, u, v))
        rc = self.lib.ref_encode_keyframe_bsize(C.c_void_p(self.h), ptr(y), ptr(u), ptr(v),
                                                quality, complexity, ptr(out))
        assert rc == 0, rc
        return out


def dct2d_blocks(which, bs, inverse, X, rc=None, impl=0):
    """nblocks tightly packed blocks through a checker: which = "port" or "ref" (ref needs a
    RefCtx; impl 0 = the C tables, 1 = the x86 dispatch)."""
    X = np.ascontiguousarray(X, dtype=np.int32)
    out = np.zeros_like(X)
    if which == "ref":
        rc.lib.ref_dct2d_blocks(C.c_void_p(rc.h), bs, int(inverse), impl, ptr(out), ptr(X), X.shape[0])
    else:
        port().oc_dct2d_blocks(bs, int(inverse), ptr(out), ptr(X), X.shape[0])
    return out


def dct1d_vectors(which, bs, inverse, X):
    X = np.ascontiguousarray(X, dtype=np.int32)
    out = np.zeros_like(X)
    f = ref().ref_dct1d_vectors if which == "ref" else port().oc_dct1d_vectors
    f(bs, int(inverse), ptr(out), ptr(X), X.shape[0])
    return out


# ------------------------------------------------------------- generators --
def lcg_stream(seed, n):
    """The IEEE-1180 style LCG the reference's self-test seeds its vectors with
    (x = x*1103515245 + 12345, reference src/dct.c:2385)."""
    out = np.empty(n, dtype=np.int64)
    x = seed
    for i in range(n):
        x = (x * 1103515245 + 12345) & 0xFFFFFFFF
        out[i] = x
    return out


def lcg_stream_fast(seed, n):
    """lcg_stream without the Python loop: the n-step affine map is built by doubling
    (x -> a*x + c composed with itself), all in uint32 arithmetic.  Returns (values, next
    seed) with values uint32, bit-identical to lcg_stream."""
    a = np.uint64(1103515245)
    c = np.uint64(12345)
    m = np.uint64(0xFFFFFFFF)
    out = np.empty(max(n, 1), dtype=np.uint64)
    out[0] = (np.uint64(seed & 0xFFFFFFFF) * a + c) & m
    have = 1
    A, Cc = a, c                      # the `have`-step map
    while have < n:
        take = min(have, n - have)
        out[have:have + take] = (out[:take] * A + Cc) & m
        Cc = (A * Cc + Cc) & m
        A = (A * A) & m
        have += take
    out = out[:n]
    return out.astype(np.uint32), (int(out[-1]) if n else seed)


def ieee1180_rand(seed, n, l, h):
    """n values of the reference self-test's ieee1180_rand(l, h) (src/dct.c:2378-2389),
    starting from generator state `seed`; returns (int32 values, next state)."""
    x, nxt = lcg_stream_fast(seed, n)
    i = (x & np.uint32(0x7FFFFFFE)).astype(np.float64)
    v = (i / float(0x7FFFFFFF) * (l + h + 1)).astype(np.int64) - l
    return v.astype(np.int32), nxt


def random_quadtree_bsize(nhsb, nvsb, rng, p32=0.8, p16=0.6, p8=0.5):
    """Random block-size map: one byte per 8x8 luma block."""
    b = np.full((nvsb * 4, nhsb * 4), 3, dtype=np.uint8)
    for sy in range(nvsb):
        for sx in range(nhsb):
            if rng.random() >= p32:
                continue
            for qy in range(2):
                for qx in range(2):
                    y0, x0 = sy * 4 + qy * 2, sx * 4 + qx * 2
                    if rng.random() >= p16:
                        b[y0:y0 + 2, x0:x0 + 2] = 2
                        continue
                    for ry in range(2):
                        for rx in range(2):
                            b[y0 + ry, x0 + rx] = 0 if rng.random() < p8 else 1
    return b


def ar1_plane(h, w, rng, rho=0.95, sigma=40.0, mean=0.0, scale=16, lo=-2048, hi=2047):
    """Separable AR(1) field (the reference's own image model, src/dct.c:2209-2222),
    returned as int32 coefficients-domain samples."""
    e = rng.standard_normal((h, w)) * sigma * (1 - rho * rho)
    a = np.empty_like(e)
    a[0] = e[0]
    for i in range(1, h):
        a[i] = rho * a[i - 1] + e[i]
    for j in range(1, w):
        a[:, j] = rho * a[:, j - 1] + a[:, j]
    v = np.clip(np.rint(a + mean) * scale, lo, hi)
    return v.astype(np.int32)


# ---- the same reference loops in C (oracle/ref_glue.c), for the TIMED CPU baselines ----------
def ref_dering_frame_c(rc, planes, decs, flags, bskip, q, impl=1):
    """ref_dering_frame with the superblock loop in C (what bench.py / tools/bench_dering.py time)."""
    planes = [np.ascontiguousarray(p, dtype=np.int32) for p in planes]
    out = [np.zeros_like(p) for p in planes]
    flags = np.ascontiguousarray(flags, dtype=np.uint8)
    bskip = [np.ascontiguousarray(b, dtype=np.uint8) for b in bskip]
    n, pa, da, ba = dering_args(planes, decs, bskip)
    oa = (C.c_void_p * 4)(*[o.ctypes.data for o in out])
    rc.lib.ref_dering_frame(C.c_void_p(rc.h), oa, pa, n, da, ptr(flags), ba, bskip[0].shape[1],
                            (C.c_int * 4)(*q), impl)
    return out


def ref_smooth_frame_c(rc, planes, decs, bsize, q):
    out = [np.ascontiguousarray(p, dtype=np.int32).copy() for p in planes]
    bsize = np.ascontiguousarray(bsize, dtype=np.uint8)
    pa = (C.c_void_p * 4)(*[o.ctypes.data for o in out])
    rc.lib.ref_smooth_frame(C.c_void_p(rc.h), pa, len(out), (C.c_int * 4)(*decs), ptr(bsize),
                            bsize.shape[1], (C.c_int * 4)(*q))
    return out


def ref_cfl_pred_plane_c(d_luma, bsize):
    d_luma = np.ascontiguousarray(d_luma, dtype=np.int32)
    H, W = d_luma.shape
    bsize = np.ascontiguousarray(bsize, dtype=np.uint8)
    l = np.zeros((H // 2, W // 2), np.int32)
    ref().ref_cfl_pred_frame(ptr(l), ptr(d_luma), W, H, ptr(bsize), bsize.shape[1])
    return l


def ref_coding_order_plane_c(src, bsize, dec, to_coding=True, out=None):
    src = np.ascontiguousarray(src, dtype=np.int32).copy()
    out = np.zeros_like(src) if out is None else out
    H, W = src.shape
    bsize = np.ascontiguousarray(bsize, dtype=np.uint8)
    v, d = (out, src) if to_coding else (src, out)
    ref().ref_coding_order_plane(ptr(v), ptr(d), W, H, dec, ptr(bsize), bsize.shape[1], 1 if to_coding else 0)
    return out


# ---- the whole codec through the reference's call sites (oracle/ref_glue.c) -------------------
REF_CUDA_SO = os.path.join(ORACLE_DIR, "_ref", "libdaala_ref_cuda.so")


def codec_roundtrip(lib, frames, quality, impl, complexity=7, keyframe_rate=1):
    """frames: list of (y, u, v) uint8 planes (4:2:0).  Runs daala_encode_img_in ->
    daala_encode_packet_out -> daala_decode_packet_in -> daala_decode_img_out in `lib` (the
    compiled reference, or its CUDA-redirected twin).  Returns (packet bytes, per-frame packet
    sizes, list of decoded (y, u, v))."""
    h, w = frames[0][0].shape
    cw, ch = (w + 1) // 2, (h + 1) // 2
    per = w * h + 2 * cw * ch
    yuv = np.concatenate([np.concatenate([p.ravel() for p in f]) for f in frames]).astype(np.uint8)
    cap = per * len(frames) * 2 + (1 << 20)
    packets = np.zeros(cap, np.uint8)
    sizes = np.zeros(len(frames), np.int32)
    recon = np.zeros(per * len(frames), np.uint8)
    lib.ref_codec_roundtrip.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                        C.c_int, C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p]
    n = lib.ref_codec_roundtrip(w, h, len(frames), ptr(yuv), quality, complexity, keyframe_rate, impl,
                                ptr(packets), cap, ptr(sizes), ptr(recon))
    assert n == len(frames), "ref_codec_roundtrip returned %d" % n
    out = []
    for f in range(len(frames)):
        b = recon[f * per:(f + 1) * per]
        out.append((b[:w * h].reshape(h, w).copy(), b[w * h:w * h + cw * ch].reshape(ch, cw).copy(),
                    b[w * h + cw * ch:].reshape(ch, cw).copy()))
    return packets[:int(sizes.sum())].copy(), sizes, out


# ---- the encoder's input side (od_img_copy_pad) ------------------------------------------------
def _img_pad_buffers(pic, plane_w, plane_h, border, dst_bits):
    ddt = np.uint8 if dst_bits == 8 else np.int16
    buf = np.full((plane_h + 2 * border, plane_w + 2 * border), 77, ddt)
    return buf, buf[border:, border:]


def port_img_copy_pad_plane(pic, src_bits, plane_w, plane_h, border, dst_bits):
    """pic: (pic_h, pic_w) uint8 / int16 picture plane.  Returns the bordered plane buffer
    (plane_h + 2 border, plane_w + 2 border) of the restatement."""
    lib = port()
    pic = np.ascontiguousarray(pic)
    buf, vis = _img_pad_buffers(pic, plane_w, plane_h, border, dst_bits)
    lib.oc_img_copy_pad_plane.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                          C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
    lib.oc_img_copy_pad_plane(C.c_void_p(vis.ctypes.data), buf.itemsize, buf.strides[0], dst_bits, plane_w,
                              plane_h, border, ptr(pic), pic.itemsize, pic.strides[0], src_bits,
                              pic.shape[1], pic.shape[0])
    return buf


def ref_img_copy_pad_plane(pic, src_bits, plane_w, plane_h, dec, dst_bits, edge_ext=True):
    """The same through the REFERENCE's od_img_plane_copy_pad + od_img_edge_ext; the border is
    the reference's OD_BUFFER_PADDING >> dec."""
    lib = ref()
    pic = np.ascontiguousarray(pic)
    border = lib.ref_buffer_padding() >> dec
    buf, vis = _img_pad_buffers(pic, plane_w, plane_h, border, dst_bits)
    lib.ref_img_copy_pad_plane.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                           C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
    lib.ref_img_copy_pad_plane(C.c_void_p(vis.ctypes.data), buf.itemsize, buf.strides[0], dst_bits, plane_w,
                               plane_h, dec, ptr(pic), pic.itemsize, pic.strides[0], src_bits,
                               pic.shape[1], pic.shape[0], 1 if edge_ext else 0)
    return buf
