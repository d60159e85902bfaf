#!/usr/bin/env python3
"""Groups an ncu source page by named line ranges. usage: ncu_groups.py rep kernel-regex groups.txt
groups.txt lines: name file first last"""
import csv, subprocess, sys, io, collections
rep, kre, gfile = sys.argv[1:4]
groups = [l.split() for l in open(gfile) if l.strip() and not l.startswith('#')]
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv",
                      "--kernel-name", "regex:" + kre], capture_output=True, text=True).stdout
fpath = None; hdr = None
tot = collections.Counter(); thr = collections.Counter(); smp = collections.Counter()
for row in csv.reader(io.StringIO(out)):
    if not row: continue
    if row[0] == "File Path": fpath = row[1].split("/")[-1]; continue
    if row[0] == "Line No": hdr = row; continue
    if hdr is None or len(row) < 10 or row[2] != "-": continue
    try: ln = int(row[0])
    except ValueError: continue
    name = "other:" + fpath
    for g in groups:
        if g[1] == fpath and int(g[2]) <= ln <= int(g[3]): name = g[0]; break
    tot[name] += int(row[hdr.index("Instructions Executed")] or 0)
    thr[name] += int(row[hdr.index("Thread Instructions Executed")] or 0)
    smp[name] += int(row[hdr.index("# Samples")] or 0)
T = sum(tot.values()); S = sum(smp.values())
nsamp = float(sys.argv[4]) if len(sys.argv) > 4 else None
print("%-28s %7s %7s %6s %s" % ("group", "inst%", "samp%", "thr/w", "lane-instr/sample" if nsamp else ""))
for k, v in tot.most_common():
    print("%-28s %6.2f%% %6.2f%% %6.1f %s" % (k, 100.0*v/T, 100.0*smp[k]/max(S,1), thr[k]/max(v,1), "%.2f" % (32.0*v/nsamp) if nsamp else ""))
print("total %d warp-instr %s" % (T, "%.1f lane-instr/sample" % (32.0*T/nsamp) if nsamp else ""))
