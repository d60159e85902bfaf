#!/usr/bin/env python3
"""Dynamic SASS opcode histogram of one kernel from an ncu report (source page).
usage: ncu_opcodes.py report.ncu-rep kernel-regex"""
import collections, csv, io, subprocess, sys
rep, kre = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "sass", "--csv",
                      "--kernel-name", "regex:" + kre], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = None
cnt = collections.Counter(); smp = collections.Counter(); nstatic = collections.Counter()
FMA = ("IMAD", "FFMA", "FMUL", "HFMA2", "FADD")
for r in rows:
    if r and r[0] == "Address": hdr = r; continue
    if hdr is None or len(r) < 8 or not r[0].startswith("0x"): continue
    src = r[1].strip()
    toks = src.split()
    op = toks[1] if toks[0].startswith("@") else toks[0]
    base = op.split(".")[0]
    key = op if base in ("IMAD", "LEA", "SHF", "LDS", "STS", "LDG", "STG", "BAR", "ATOMS", "SHFL", "IADD3", "LOP3", "ISETP", "BRA", "MOV", "SEL", "UTMALDG", "UTMASTG") else base
    if base in ("LEA", "SHF", "IMAD", "LDS", "STS"): key = ".".join(op.split(".")[:2])
    n = int(r[hdr.index("Instructions Executed")] or 0)
    cnt[key] += n; nstatic[key] += 1
    smp[key] += int(r[hdr.index("# Samples")] or 0)
T = sum(cnt.values()); S = sum(smp.values())
print("total %d warp-instr, %d static instructions" % (T, sum(nstatic.values())))
fma = sum(v for k, v in cnt.items() if k.split(".")[0] in FMA)
print("fma-pipe %.1f%%" % (100.0*fma/T))
for k, v in cnt.most_common(40):
    print("%-16s %6.2f%%  samp %6.2f%%  static %d" % (k, 100.0*v/T, 100.0*smp[k]/max(S, 1), nstatic[k]))
