#!/usr/bin/env python3
"""Step tables for Daala's reversible integer lifting DCTs, and the emitters
that turn them into (a) straight-line CUDA device code for the product and
(b) plain-C data tables for the CPU oracle's interpreter.

Every transform is written ONCE, as the forward sequence of elementary,
individually invertible integer steps.  The inverse is derived mechanically by
running the steps backwards with each step inverted; because every step is a
bijection on Z^n the inverse obtained this way is THE inverse function, so it
is bit-identical to the reference's hand-written od_bin_idct* (this is checked
against the compiled reference on random *coefficient* inputs, not only on
round trips, in tests/test_oracle_vs_ref.py).

Step vocabulary (d = destination register, s = source register, d != s):
  ADD   d += s                 SUB   d -= s               RSUB  d = s - d
  ADDH  d += half(s)           SUBH  d -= half(s)         RSUBH d = half(s) - d
  NEG   d = -d
  LADD  d += (s*K + 2^(S-1)) >> S                         LSUB  d -= (same)
half(a) is division by two rounding toward zero, the reference's
OD_DCT_RSHIFT(a, 1) (reference src/filter.h:40-43).  The reference caches
half(x) in "xh" temporaries; here half() is written at each use and the
compiler's CSE recovers the sharing (each cached value in the reference is
only used while its source register is unchanged, which was checked by hand
while transcribing and is re-checked numerically against the reference).

Sources followed (algorithm, not text): reference src/dct.c
  4-pt   :74-112     8-pt :151-269     16-pt :349-640
  32-pt  :1638-1701, 1757-1857 with the embedded kernels at :790-1401.

Usage:  python tools/gen_lifting.py      (rewrites the two generated headers)
"""
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

ADD, SUB, RSUB, ADDH, SUBH, RSUBH, NEG, LADD, LSUB = range(9)
OPNAMES = ["ADD", "SUB", "RSUB", "ADDH", "SUBH", "RSUBH", "NEG", "LADD", "LSUB"]
INVERSE = {ADD: SUB, SUB: ADD, RSUB: RSUB, ADDH: SUBH, SUBH: ADDH,
           RSUBH: RSUBH, NEG: NEG, LADD: LSUB, LSUB: LADD}


class Prog:
    """A forward transform: register names, input placement, and steps."""

    def __init__(self, name, n, in_regs):
        # in_regs[k] = name of the register that receives x[k].
        # Output y[r] is register number r (names sort as base-32 digits).
        self.name = name
        self.n = n
        self.in_regs = list(in_regs)
        self.steps = []
        assert sorted(self.in_regs) == sorted(set(self.in_regs))
        assert len(self.in_regs) == n

    def _e(self, op, d, s=None, k=0, sh=0):
        assert d != s
        self.steps.append((op, d, s, k, sh))

    def add(self, d, s): self._e(ADD, d, s)
    def sub(self, d, s): self._e(SUB, d, s)
    def rsub(self, d, s): self._e(RSUB, d, s)
    def addh(self, d, s): self._e(ADDH, d, s)
    def subh(self, d, s): self._e(SUBH, d, s)
    def rsubh(self, d, s): self._e(RSUBH, d, s)
    def neg(self, *ds):
        for d in ds:
            self._e(NEG, d)
    def ladd(self, d, s, k, sh): self._e(LADD, d, s, k, sh)
    def lsub(self, d, s, k, sh): self._e(LSUB, d, s, k, sh)

    def rot(self, a, b, spec):
        """Three lifting steps on the pair (a, b), alternating a, b, a.
        spec = ((sign, K, S), (sign, K, S), (sign, K, S))."""
        tgt = (a, b, a)
        src = (b, a, b)
        for (sg, k, sh), d, s in zip(spec, tgt, src):
            (self.ladd if sg > 0 else self.lsub)(d, s, k, sh)


def regidx(name):
    return int(name[1:], 32)


def regname(i):
    return "t" + "0123456789abcdefghijklmnopqrstuv"[i]


# --------------------------------------------------------------------------
# 4-point (reference src/dct.c:74-112)
# --------------------------------------------------------------------------
def build_fdct4():
    p = Prog("dct4", 4, ["t0", "t2", "t1", "t3"])
    p.rsub("t3", "t0")
    p.add("t2", "t1")
    p.rsubh("t1", "t2")
    p.subh("t0", "t3")
    p.addh("t0", "t2")
    p.rsub("t2", "t0")
    p.lsub("t3", "t1", 23013, 15)
    p.ladd("t1", "t3", 21407, 15)
    p.lsub("t3", "t1", 18293, 14)
    return p


# --------------------------------------------------------------------------
# 8-point (reference src/dct.c:151-269)
# --------------------------------------------------------------------------
def build_fdct8():
    p = Prog("dct8", 8, ["t0", "t4", "t2", "t6", "t7", "t3", "t5", "t1"])
    p.rsub("t1", "t0")
    p.subh("t0", "t1")
    p.add("t4", "t5")
    p.subh("t5", "t4")
    p.rsub("t3", "t2")
    p.subh("t2", "t3")
    p.add("t6", "t7")
    p.rsubh("t7", "t6")
    # embedded 4-point DCT
    p.addh("t0", "t6")
    p.rsub("t6", "t0")
    p.rsubh("t2", "t4")
    p.rsub("t4", "t2")
    p.rot("t0", "t4", ((-1, 13573, 15), (+1, 11585, 14), (-1, 13573, 15)))
    p.rot("t6", "t2", ((-1, 21895, 15), (+1, 15137, 14), (-1, 21895, 15)))
    # embedded 4-point DST
    p.rot("t3", "t5", ((+1, 19195, 15), (+1, 11585, 14), (-1, 7489, 13)))
    p.rsubh("t7", "t5")
    p.sub("t5", "t7")
    p.rsubh("t3", "t1")
    p.sub("t1", "t3")
    p.rot("t7", "t1", ((+1, 3227, 15), (-1, 6393, 15), (+1, 3227, 15)))
    p.rot("t5", "t3", ((+1, 2485, 13), (-1, 18205, 15), (+1, 2485, 13)))
    return p


# --------------------------------------------------------------------------
# 16-point (reference src/dct.c:349-640)
# --------------------------------------------------------------------------
def build_fdct16():
    p = Prog("dct16", 16,
             ["t0", "t8", "t4", "tc", "te", "ta", "t6", "t2",
              "t3", "td", "t9", "tf", "t1", "t7", "tb", "t5"])
    p.rsub("t5", "t0")
    p.add("t8", "tb")
    p.rsub("t7", "t4")
    p.add("tc", "t1")
    p.rsub("tf", "te")
    p.add("ta", "t9")
    p.rsub("td", "t6")
    p.add("t2", "t3")
    p.subh("t0", "t5")
    p.rsubh("tb", "t8")
    p.subh("t4", "t7")
    p.rsubh("t1", "tc")
    p.subh("te", "tf")
    p.rsubh("t9", "ta")
    p.subh("t6", "td")
    p.rsubh("t3", "t2")
    p.addh("t0", "t2")
    p.rsubh("t6", "t8")
    p.addh("t4", "ta")
    p.rsubh("te", "tc")
    p.rsub("t2", "t0")
    p.sub("t8", "t6")
    p.rsub("ta", "t4")
    p.sub("tc", "te")
    p.rsub("tc", "t0")
    p.add("t8", "t4")
    p.rsubh("t4", "t8")
    p.subh("t0", "tc")
    p.addh("t0", "t8")
    p.rsub("t8", "t0")
    p.rot("tc", "t4", ((-1, 23013, 15), (+1, 10703, 14), (-1, 9147, 13)))
    p.rot("t6", "ta", ((+1, 13573, 15), (-1, 11585, 14), (+1, 13573, 15)))
    p.add("ta", "te")
    p.add("t2", "t6")
    p.rsubh("te", "ta")
    p.rsubh("t6", "t2")
    p.rot("te", "t2", ((+1, 2275, 11), (-1, 9041, 15), (-1, 2873, 11)))
    p.rot("t6", "ta", ((-1, 8593, 14), (+1, 12873, 14), (+1, 7335, 15)))
    p.rot("t3", "t5", ((+1, 1035, 11), (-1, 14699, 14), (-1, 851, 13)))
    p.rot("tb", "td", ((+1, 17515, 15), (-1, 20435, 14), (+1, 4379, 14)))
    p.rot("t9", "t7", ((+1, 12905, 14), (-1, 3363, 13), (-1, 14101, 14)))
    p.rot("t1", "tf", ((+1, 5417, 13), (-1, 23059, 14), (+1, 20055, 15)))
    p.rsub("tf", "t3")
    p.add("td", "t9")
    p.subh("t3", "tf")
    p.rsubh("t9", "td")
    p.add("t1", "t5")
    p.rsub("tb", "t7")
    p.rsubh("t5", "t1")
    p.subh("t7", "tb")
    p.addh("t3", "tb")
    p.rsubh("t5", "td")
    p.addh("t9", "tf")
    p.rsubh("t7", "t1")
    p.sub("tb", "t3")
    p.sub("td", "t5")
    p.rsub("tf", "t9")
    p.sub("t1", "t7")
    p.rot("t5", "tb", ((-1, 10947, 14), (+1, 15137, 14), (-1, 10947, 14)))
    p.rot("td", "t3", ((+1, 21895, 15), (-1, 15137, 14), (+1, 10947, 14)))
    p.rot("t1", "tf", ((-1, 13573, 15), (+1, 11585, 14), (-1, 13573, 15)))
    return p


# --------------------------------------------------------------------------
# 32-point: embedded kernels (reference src/dct.c:790-1401, 1638-1701).
# Each helper appends to p and takes register names; "xh" arguments of the
# reference's macros are dropped (half() is taken at the point of use).
# --------------------------------------------------------------------------
def k_fdct2(p, t0, t1):
    p.rot(t0, t1, ((-1, 13573, 15), (+1, 5793, 13), (-1, 3393, 13)))


def k_fdst2(p, t0, t1):
    p.rot(t0, t1, ((-1, 10947, 14), (+1, 473, 9), (-1, 10947, 14)))


def k_fdct4_asym(p, t0, t2, t1, t3):
    p.addh(t0, t3)
    p.rsub(t3, t0)
    p.rsubh(t1, t2)
    p.rsub(t2, t1)
    k_fdct2(p, t0, t2)
    k_fdst2(p, t3, t1)


def k_fdst4_asym(p, t0, t2, t1, t3):
    p.rot(t2, t1, ((-1, 7489, 13), (+1, 11585, 14), (+1, 19195, 15)))
    p.addh(t3, t2)
    p.sub(t2, t3)
    p.rsubh(t1, t0)
    p.sub(t0, t1)
    p.rot(t3, t0, ((+1, 6723, 13), (-1, 8035, 13), (+1, 6723, 13)))
    p.rot(t2, t1, ((+1, 8757, 14), (-1, 6811, 13), (+1, 8757, 14)))


def k_fdct8(p, t0, t4, t2, t6, t1, t5, t3, t7):
    p.rsub(t7, t0)
    p.subh(t0, t7)
    p.add(t4, t3)
    p.rsubh(t3, t4)
    p.rsub(t5, t2)
    p.subh(t2, t5)
    p.add(t6, t1)
    p.rsubh(t1, t6)
    k_fdct4_asym(p, t0, t4, t2, t6)
    k_fdst4_asym(p, t7, t3, t5, t1)


def k_fdst8(p, t0, t4, t2, t6, t1, t5, t3, t7):
    p.rot(t6, t1, ((-1, 13573, 15), (+1, 11585, 14), (-1, 13573, 15)))
    p.rot(t5, t2, ((-1, 21895, 15), (+1, 15137, 14), (-1, 10947, 14)))
    p.rot(t4, t3, ((-1, 3259, 14), (+1, 3135, 13), (-1, 3259, 14)))
    p.add(t7, t1)
    p.subh(t1, t7)
    p.rsub(t2, t3)
    p.subh(t3, t2)
    p.sub(t0, t6)
    p.addh(t6, t0)
    p.rsub(t5, t4)
    p.subh(t4, t5)
    p.addh(t1, t5)
    p.rsub(t5, t1)
    p.addh(t4, t0)
    p.sub(t0, t4)
    p.subh(t6, t2)
    p.add(t2, t6)
    p.subh(t3, t7)
    p.add(t7, t3)
    p.neg(t7)
    p.rot(t0, t7, ((-1, 7425, 13), (+1, 8153, 13), (-1, 7425, 13)))
    p.rot(t6, t1, ((-1, 4861, 15), (+1, 1189, 12), (-1, 4861, 15)))
    p.rot(t2, t5, ((-1, 2455, 12), (+1, 7225, 13), (-1, 2455, 12)))
    p.rot(t4, t3, ((-1, 11725, 15), (+1, 5197, 13), (-1, 11725, 15)))


def k_fdct16_asym(p, t0, t8, t4, tc, t2, ta, t6, te,
                  t1, t9, t5, td, t3, tb, t7, tf):
    p.addh(t0, tf)
    p.rsub(tf, t0)
    p.subh(t1, te)
    p.add(te, t1)
    p.addh(t2, td)
    p.rsub(td, t2)
    p.subh(t3, tc)
    p.add(tc, t3)
    p.addh(t4, tb)
    p.rsub(tb, t4)
    p.subh(t5, ta)
    p.add(ta, t5)
    p.addh(t6, t9)
    p.rsub(t9, t6)
    p.subh(t7, t8)
    p.add(t8, t7)
    k_fdct8(p, t0, t8, t4, tc, t2, ta, t6, te)
    k_fdst8(p, tf, t7, tb, t3, td, t5, t9, t1)


def k_fdst16_asym(p, t0, t8, t4, tc, t2, ta, t6, te,
                  t1, t9, t5, td, t3, tb, t7, tf):
    p.neg(t8, t9, ta, tb, td)
    p.rot(t1, te, ((-1, 13573, 14), (+1, 11585, 15), (-1, 13573, 14)))
    p.rot(t2, td, ((+1, 4161, 14), (-1, 15137, 14), (+1, 14341, 14)))
    p.rot(tc, t3, ((-1, 14341, 14), (+1, 15137, 14), (-1, 4161, 14)))
    p.rsubh(te, t0)
    p.sub(t0, te)
    p.rsubh(tf, t1)
    p.sub(t1, tf)
    p.neg(tc)
    p.rsubh(t2, tc)
    p.sub(tc, t2)
    p.rsubh(t3, td)
    p.rsub(td, t3)
    p.rot(t9, t6, ((-1, 7489, 13), (+1, 11585, 14), (+1, 19195, 15)))
    p.addh(t8, t9)
    p.sub(t9, t8)
    p.rsubh(t6, t7)
    p.sub(t7, t6)
    p.rot(t8, t7, ((+1, 6723, 13), (-1, 16069, 14), (+1, 6723, 13)))
    p.rot(t9, t6, ((+1, 17515, 15), (-1, 13623, 14), (+1, 17515, 15)))
    p.rot(t5, ta, ((+1, 13573, 14), (-1, 11585, 15), (+1, 13573, 14)))
    p.addh(tb, t5)
    p.rsub(t5, tb)
    p.addh(ta, t4)
    p.sub(t4, ta)
    p.rot(ta, t5, ((+1, 2485, 13), (-1, 18205, 15), (+1, 2485, 13)))
    p.rot(tb, t4, ((-1, 6723, 13), (+1, 16069, 14), (-1, 6723, 13)))
    p.neg(t5)
    p.sub(tc, tf)
    p.addh(tf, tc)
    p.add(t3, t0)
    p.subh(t0, t3)
    p.sub(td, t1)
    p.addh(t1, td)
    p.add(t2, te)
    p.subh(te, t2)
    p.add(t8, t4)
    p.rsubh(t4, t8)
    p.rsub(t7, tb)
    p.rsubh(tb, t7)
    p.sub(t6, ta)
    p.addh(ta, t6)
    p.rsub(t9, t5)
    p.subh(t5, t9)
    p.subh(t0, t7)
    p.add(t7, t0)
    p.addh(tf, t8)
    p.sub(t8, tf)
    p.subh(te, t6)
    p.add(t6, te)
    p.addh(t1, t9)
    p.sub(t9, t1)
    p.subh(tb, tc)
    p.add(tc, tb)
    p.addh(t4, t3)
    p.sub(t3, t4)
    p.subh(ta, td)
    p.add(td, ta)
    p.rsubh(t5, t2)
    p.sub(t2, t5)
    p.neg(t8, t9, ta, tb, tc, td, tf)
    p.rot(t0, tf, ((-1, 7799, 13), (+1, 4091, 12), (-1, 7799, 13)))
    p.rot(t1, te, ((+1, 2417, 15), (-1, 601, 12), (+1, 2417, 15)))
    p.rot(t7, t8, ((-1, 14525, 15), (+1, 3035, 12), (-1, 7263, 14)))
    p.rot(t2, td, ((-1, 6393, 13), (+1, 3973, 12), (-1, 6393, 13)))
    p.rot(t5, ta, ((-1, 9281, 14), (+1, 7027, 13), (-1, 9281, 14)))
    p.rot(t3, tc, ((-1, 11539, 14), (+1, 7713, 13), (-1, 11539, 14)))
    p.rot(t4, tb, ((-1, 10375, 14), (+1, 7405, 13), (-1, 10375, 14)))
    p.rot(t6, t9, ((-1, 8247, 14), (+1, 1645, 11), (-1, 8247, 14)))


def build_fdct32():
    # x[k] lands in the register whose number is the 5-bit reversal of k
    # (reference src/dct.c:1791-1822).
    def bitrev5(k):
        return int("{:05b}".format(k)[::-1], 2)
    p = Prog("dct32", 32, [regname(bitrev5(k)) for k in range(32)])
    lo = [regname(i) for i in range(16)]           # t0..tf
    hi = [regname(31 - i) for i in range(16)]      # tv..tg
    for i in range(16):
        a, b = lo[i], hi[i]
        if i % 2 == 0:
            p.rsub(b, a)
            p.subh(a, b)
        else:
            p.add(b, a)
            p.rsubh(a, b)
    k_fdct16_asym(p, "t0", "tg", "t8", "to", "t4", "tk", "tc", "ts",
                  "t2", "ti", "ta", "tq", "t6", "tm", "te", "tu")
    k_fdst16_asym(p, "tv", "tf", "tn", "t7", "tr", "tb", "tj", "t3",
                  "tt", "td", "tl", "t5", "tp", "t9", "th", "t1")
    return p


# --------------------------------------------------------------------------
# Interpreters (used to self-check the tables before emitting anything)
# --------------------------------------------------------------------------
def _half(a):
    return int(a / 2) if a % 2 == 0 or a >= 0 else -((-a) // 2)


def _apply(op, regs, d, s, k, sh):
    if op == ADD: regs[d] += regs[s]
    elif op == SUB: regs[d] -= regs[s]
    elif op == RSUB: regs[d] = regs[s] - regs[d]
    elif op == ADDH: regs[d] += _half(regs[s])
    elif op == SUBH: regs[d] -= _half(regs[s])
    elif op == RSUBH: regs[d] = _half(regs[s]) - regs[d]
    elif op == NEG: regs[d] = -regs[d]
    elif op == LADD: regs[d] += (regs[s] * k + (1 << (sh - 1))) >> sh
    elif op == LSUB: regs[d] -= (regs[s] * k + (1 << (sh - 1))) >> sh


def run_forward(p, x):
    regs = {r: x[k] for k, r in enumerate(p.in_regs)}
    for op, d, s, k, sh in p.steps:
        _apply(op, regs, d, s, k, sh)
    return [regs[regname(i)] for i in range(p.n)]


def run_inverse(p, y):
    regs = {regname(i): y[i] for i in range(p.n)}
    for op, d, s, k, sh in reversed(p.steps):
        _apply(INVERSE[op], regs, d, s, k, sh)
    return [regs[r] for r in p.in_regs]


# --------------------------------------------------------------------------
# Emitters
# --------------------------------------------------------------------------
def _c_step(op, d, s, k, sh, half="tdlt_half", cuda=False):
    if cuda and op == LSUB:
        # d - ((s*K + R) >> S) == d + ((s*-K + R - 1) >> S) for R = 2^(S-1)
        # (-(x >> S) == (-x + 2^S - 1) >> S): one IMAD + one LEA.HI.SX32 instead
        # of IMAD + SHF + IADD.
        return "%s += (%s*%d + %d) >> %d;" % (d, s, -k, (1 << (sh - 1)) - 1, sh)
    if op == ADD: return "%s += %s;" % (d, s)
    if op == SUB: return "%s -= %s;" % (d, s)
    if op == RSUB: return "%s = %s - %s;" % (d, s, d)
    if op == ADDH: return "%s += %s(%s);" % (d, half, s)
    if op == SUBH: return "%s -= %s(%s);" % (d, half, s)
    if op == RSUBH: return "%s = %s(%s) - %s;" % (d, half, s, d)
    if op == NEG: return "%s = -%s;" % (d, d)
    if op == LADD: return "%s += (%s*%d + %d) >> %d;" % (d, s, k, 1 << (sh - 1), sh)
    if op == LSUB: return "%s -= (%s*%d + %d) >> %d;" % (d, s, k, 1 << (sh - 1), sh)
    raise ValueError(op)


CUDA_HEADER = """// GENERATED by tools/gen_lifting.py -- do not edit; edit the step tables there.
// Straight-line device code for Daala's reversible integer lifting DCTs.
// v[] holds the line in natural order: samples in, coefficients out (forward),
// coefficients in, samples out (inverse).  All indices are compile-time, so
// after inlining v[] lives in registers (v may point into a larger array).  Parity targets: reference src/dct.c od_bin_fdct{4,8,
// 16,32} / od_bin_idct{4,8,16,32}.
#pragma once

#ifndef TDLT_HD
# ifdef __CUDACC__
#  define TDLT_HD __host__ __device__ __forceinline__
# else
#  define TDLT_HD static inline
# endif
#endif

// Division by two rounding toward zero (reference OD_DCT_RSHIFT(a, 1),
// src/filter.h:40-43).
TDLT_HD int tdlt_half(int a) {
  return (int)(((unsigned)a >> 31) + (unsigned)a) >> 1;
}

"""


def emit_cuda(progs):
    out = [CUDA_HEADER]
    for p in progs:
        names = [regname(i) for i in range(p.n)]
        for direction in ("f", "i"):
            out.append("TDLT_HD void tdlt_%s%s(int *v) {  // v[0..%d]\n"
                       % (direction, p.name, p.n - 1))
            if direction == "f":
                for k, r in enumerate(p.in_regs):
                    out.append("  int %s = v[%d];\n" % (r, k))
                for st in p.steps:
                    out.append("  " + _c_step(*st, cuda=True) + "\n")
                for i, r in enumerate(names):
                    out.append("  v[%d] = %s;\n" % (i, r))
            else:
                for i, r in enumerate(names):
                    out.append("  int %s = v[%d];\n" % (r, i))
                for op, d, s, k, sh in reversed(p.steps):
                    out.append("  " + _c_step(INVERSE[op], d, s, k, sh, cuda=True) + "\n")
                for k, r in enumerate(p.in_regs):
                    out.append("  v[%d] = %s;\n" % (k, r))
            out.append("}\n\n")
    return "".join(out)


def emit_tables(progs):
    out = ["/* GENERATED by tools/gen_lifting.py -- do not edit.\n"
           "   Step tables for the oracle's lifting interpreter (oracle/tdlt_oracle.c).\n"
           "   Each row: {op, dst, src, K, shift}; registers are numbered so that\n"
           "   output coefficient r lives in register r. */\n"
           "#ifndef TDLT_LIFT_TABLES_GEN_H\n#define TDLT_LIFT_TABLES_GEN_H\n\n"
           "enum {" + ", ".join("TL_%s = %d" % (n, i) for i, n in enumerate(OPNAMES))
           + "};\n\n"
           "typedef struct { unsigned char op, dst, src, shift; int k; } tl_step;\n\n"]
    for p in progs:
        out.append("static const unsigned char TL_%s_IN[%d] = {%s};\n"
                   % (p.name.upper(), p.n,
                      ", ".join(str(regidx(r)) for r in p.in_regs)))
        out.append("static const tl_step TL_%s_STEPS[%d] = {\n"
                   % (p.name.upper(), len(p.steps)))
        for op, d, s, k, sh in p.steps:
            out.append("  {TL_%s, %d, %d, %d, %d},\n"
                       % (OPNAMES[op], regidx(d), regidx(s) if s else 0, sh, k))
        out.append("};\n\n")
    out.append("#endif\n")
    return "".join(out)


def census(p):
    c = {}
    for op, *_ in p.steps:
        c[OPNAMES[op]] = c.get(OPNAMES[op], 0) + 1
    return c


def main():
    import random
    progs = [build_fdct4(), build_fdct8(), build_fdct16(), build_fdct32()]
    rnd = random.Random(1)
    for p in progs:
        for _ in range(200):
            x = [rnd.randint(-4096, 4095) for _ in range(p.n)]
            assert run_inverse(p, run_forward(p, x)) == x, p.name
        print(p.name, len(p.steps), "steps", census(p))
    cu = os.path.join(ROOT, "daala_b200", "csrc", "tdlt_lift_gen.cuh")
    tb = os.path.join(ROOT, "oracle", "lift_tables_gen.h")
    for path, text in ((cu, emit_cuda(progs)), (tb, emit_tables(progs))):
        os.makedirs(os.path.dirname(path), exist_ok=True)
        with open(path, "w") as f:
            f.write(text)
        print("wrote", os.path.relpath(path, ROOT))


if __name__ == "__main__":
    main()
