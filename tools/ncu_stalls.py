#!/usr/bin/env python3
"""Stall-reason totals and smem-conflict hot spots from an ncu report's source page.
usage: ncu_stalls.py report.ncu-rep kernel-regex"""
import csv, subprocess, sys, io, collections
rep, kre = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv",
                      "--kernel-name", "regex:" + kre], capture_output=True, text=True).stdout
fpath = None; hdr = None
stall = collections.Counter(); conf = collections.Counter(); wf = collections.Counter(); ideal = collections.Counter(); text = {}
for row in csv.reader(io.StringIO(out)):
    if not row: continue
    if row[0] == "File Path": fpath = row[1].split("/")[-1]; continue
    if row[0] == "Line No": hdr = row; continue
    if hdr is None or len(row) < 10 or row[2] != "-": continue
    try: ln = int(row[0])
    except ValueError: continue
    key = (fpath, ln); text[key] = row[1].strip()[:80]
    for i, h in enumerate(hdr):
        if h.startswith("stall_") and "Not Issued" not in h:
            try: stall[h] += int(row[i] or 0)
            except ValueError: pass
    wf[key] += int(row[hdr.index("L1 Wavefronts Shared")] or 0)
    ideal[key] += int(row[hdr.index("L1 Wavefronts Shared Ideal")] or 0)
    conf[key] += int(row[hdr.index("L1 Wavefronts Shared Excessive")] or 0)
T = sum(stall.values())
print("stall samples:", T)
for k, v in stall.most_common(12): print("  %-26s %6.2f%%" % (k, 100.0*v/T))
W = sum(wf.values()); E = sum(conf.values())
print("smem wavefronts %d  ideal %d  excessive %d (%.1f%%)" % (W, sum(ideal.values()), E, 100.0*E/max(W,1)))
for k, v in wf.most_common(16):
    print("  %-20s %4d wf %6.2f%%  excess %6.2f%% of its wf   %s" % (k[0], k[1], 100.0*v/W, 100.0*conf[k]/max(v,1), text[k]))
