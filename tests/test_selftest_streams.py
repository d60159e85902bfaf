"""The reference self-test's own input streams at their full volume
(/root/reference/src/dct.c): ieee1180_test = 2 signs x 3 ranges x 10 000 blocks per size from
ieee1180_srand(1) (:3439-3469, block fill :3334-3338, generator :2378-2389), and check_bias =
10 000 000 1-D vectors per size (:3565-3598).  The reference checks perfect reconstruction
on them; here every output is compared bit for bit: the CUDA library against the compiled
reference (C tables and x86 dispatch) where it is present, else against the restatement --
and, on the CPU, the restatement against the compiled reference.
"""
import numpy as np
import pytest

import oracle_libs as ol

RANGES = ((256, 255), (5, 5), (300, 300))      # IEEE1180_L / IEEE1180_H, src/dct.c:2406-2407
NBLOCKS = 10000                                 # IEEE1180_NBLOCKS, src/dct.c:2404
COEFF_SCALE = 16                                # OD_COEFF_SCALE, src/internal.h:120


def ieee1180_blocks(bs):
    """The six block sets of ieee1180_test(bs), in the order the reference draws them."""
    n = 4 << bs
    sets = []
    for sign in (1, -1):
        state = 1                               # ieee1180_srand(1) before each sign
        for l, h in RANGES:
            v, state = ol.ieee1180_rand(state, NBLOCKS * n * n, l, h)
            sets.append((v * (sign * COEFF_SCALE)).astype(np.int32).reshape(NBLOCKS, n, n))
    return sets


def bias_vectors(bs, state, nvec):
    n = 4 << bs
    v, state = ol.ieee1180_rand(state, nvec * n, 255, 255)
    return (v * COEFF_SCALE).astype(np.int32).reshape(nvec, n), state


def _checkers():
    """[(name, 2-D function, 1-D function)]: the compiled reference when it is here."""
    out = []
    if ol.have_ref():
        rc = ol.RefCtx(64, 64)
        for impl in (0, 1):
            out.append(("reference impl=%d" % impl,
                        lambda bs, inv, X, impl=impl: ol.dct2d_blocks("ref", bs, inv, X, rc, impl),
                        lambda bs, inv, X: ol.dct1d_vectors("ref", bs, inv, X)))
    else:
        out.append(("port", lambda bs, inv, X: ol.dct2d_blocks("port", bs, inv, X),
                    lambda bs, inv, X: ol.dct1d_vectors("port", bs, inv, X)))
    return out


# ------------------------------------------------------------------ CPU half
@pytest.mark.parametrize("bs", [0, 1, 2, 3])
def test_port_matches_reference_on_ieee1180_streams(bs):
    if not ol.have_ref():
        pytest.skip("compiled reference not present")
    with ol.RefCtx(64, 64) as rc:
        for X in ieee1180_blocks(bs):
            F = ol.dct2d_blocks("port", bs, 0, X)
            assert np.array_equal(F, ol.dct2d_blocks("ref", bs, 0, X, rc, 0))
            assert np.array_equal(F, ol.dct2d_blocks("ref", bs, 0, X, rc, 1))
            R = ol.dct2d_blocks("port", bs, 1, F)
            assert np.array_equal(R, ol.dct2d_blocks("ref", bs, 1, F, rc, 1))
            assert np.array_equal(R, X)         # the reference's own check (src/dct.c:3365)


@pytest.mark.parametrize("bs", [0, 1, 2, 3])
def test_port_matches_reference_on_a_slice_of_the_bias_stream(bs):
    if not ol.have_ref():
        pytest.skip("compiled reference not present")
    X, _ = bias_vectors(bs, 1, 500000)
    Y = ol.dct1d_vectors("port", bs, 0, X)
    assert np.array_equal(Y, ol.dct1d_vectors("ref", bs, 0, X))
    assert np.array_equal(ol.dct1d_vectors("port", bs, 1, Y), X)


# ------------------------------------------------------------------ GPU half
@pytest.mark.gpu
@pytest.mark.parametrize("bs", [0, 1, 2, 3])
def test_cuda_ieee1180_streams(bs):
    import daala_b200
    L = daala_b200.lib()
    n = 4 << bs
    chk = _checkers()
    for X in ieee1180_blocks(bs):
        F = np.zeros_like(X)
        R = np.zeros_like(X)
        assert L.od_tdlt_dct2d_blocks(bs, 0, F.ctypes.data, n, n * n, X.ctypes.data, n, n * n, NBLOCKS) == 0
        assert L.od_tdlt_dct2d_blocks(bs, 1, R.ctypes.data, n, n * n, F.ctypes.data, n, n * n, NBLOCKS) == 0
        assert np.array_equal(R, X)
        for name, f2, _ in chk:
            assert np.array_equal(F, f2(bs, 0, X)), name
            assert np.array_equal(R, f2(bs, 1, F)), name


@pytest.mark.gpu
@pytest.mark.parametrize("bs", [0, 1, 2, 3])
def test_cuda_check_bias_stream(bs):
    import daala_b200
    L = daala_b200.lib()
    chk = _checkers()
    state, left, chunk = 1, 10000000, 1000000
    while left:
        X, state = bias_vectors(bs, state, chunk)
        Y = np.zeros_like(X)
        Z = np.zeros_like(X)
        assert L.od_tdlt_dct1d_vectors(bs, 0, Y.ctypes.data, X.ctypes.data, chunk) == 0
        assert L.od_tdlt_dct1d_vectors(bs, 1, Z.ctypes.data, Y.ctypes.data, chunk) == 0
        assert np.array_equal(Z, X)             # src/dct.c:3583-3596
        name, _, f1 = chk[0]
        assert np.array_equal(Y, f1(bs, 0, X)), name
        assert np.array_equal(Z, f1(bs, 1, Y)), name
        left -= chunk
