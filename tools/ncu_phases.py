#!/usr/bin/env python3
"""Splits a kernel's SASS (address order) at BAR.SYNC / named markers and reports dynamic
warp-instructions, stall samples and code bytes per segment.
usage: ncu_phases.py report.ncu-rep kernel-regex [ncta]"""
import csv, io, subprocess, sys
rep, kre = sys.argv[1], sys.argv[2]
ncta = float(sys.argv[3]) if len(sys.argv) > 3 else None
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "sass", "--csv",
                      "--kernel-name", "regex:" + kre], capture_output=True, text=True).stdout
import collections
hdr = None
segs = [[0, 0, 0, 0, "start", collections.Counter()]]
for r in csv.reader(io.StringIO(out)):
    if r and r[0] == "Address": hdr = r; continue
    if hdr is None or len(r) < 8 or not r[0].startswith("0x"): continue
    src = r[1].strip()
    n = int(r[hdr.index("Instructions Executed")] or 0)
    s = int(r[hdr.index("# Samples")] or 0)
    seg = segs[-1]
    seg[0] += n; seg[1] += s; seg[2] += 16
    for i, h in enumerate(hdr):
        if h.startswith("stall_") and "Not Issued" not in h:
            try: seg[5][h[6:]] += int(r[i] or 0)
            except ValueError: pass
    if "BAR.SYNC" in src or "UTMASTG" in src or "SYNCS.PHASECHK" in src or "ATOMS" in src:
        seg[3] = n
        segs.append([0, 0, 0, 0, src[:40], collections.Counter()])
T = sum(s[0] for s in segs); S = sum(s[1] for s in segs)
print("total %d warp-instr%s, %d samples" % (T, " (%.0f per CTA)" % (T/ncta) if ncta else "", S))
for i, s in enumerate(segs):
    tops = " ".join("%s %.0f%%" % (k, 100.0*v/max(sum(s[5].values()), 1)) for k, v in s[5].most_common(4))
    print("seg %2d after %-28s inst %6.2f%% %s samp %6.2f%% code %6d B | %s" % (
        i, s[4][:28], 100.0*s[0]/T, "(%6.0f/CTA)" % (s[0]/ncta) if ncta else "", 100.0*s[1]/max(S, 1), s[2], tops))
