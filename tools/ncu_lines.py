#!/usr/bin/env python3
"""Aggregates an ncu report's source page per CUDA source line.
usage: ncu_lines.py report.ncu-rep kernel-regex [top]"""
import csv, subprocess, sys, io, collections
rep, kre = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv",
                      "--kernel-name", "regex:" + kre], capture_output=True, text=True).stdout
fpath = None
hdr = None
tot = collections.Counter(); thr = collections.Counter(); samp = collections.Counter(); text = {}
seen_kernel = 0
for row in csv.reader(io.StringIO(out)):
    if not row: continue
    if row[0] == "File Path": fpath = row[1].split("/")[-1]; continue
    if row[0] == "Function Name":
        continue
    if row[0] == "Line No": hdr = row; continue
    if hdr is None or len(row) < 10: continue
    if row[2] != "-": continue   # SASS row
    try:
        ln = int(row[0])
    except ValueError:
        continue
    ie = hdr.index("Instructions Executed"); te = hdr.index("Thread Instructions Executed"); se = hdr.index("# Samples")
    key = (fpath, ln)
    tot[key] += int(row[ie] or 0); thr[key] += int(row[te] or 0); samp[key] += int(row[se] or 0)
    text[key] = row[1].strip()[:90]
T = sum(tot.values()); S = sum(samp.values())
print("total warp-instructions %d  thread-instr %d  samples %d" % (T, sum(thr.values()), S))
byfile = collections.Counter()
for k, v in tot.items(): byfile[k[0]] += v
for f, v in byfile.most_common(): print("  %-24s %6.2f%%" % (f, 100.0*v/T))
print("%-22s %5s %7s %7s %6s  %s" % ("file", "line", "inst%", "samp%", "thr/w", "source"))
for k, v in tot.most_common(top):
    print("%-22s %5d %6.2f%% %6.2f%% %6.1f  %s" % (k[0], k[1], 100.0*v/T, 100.0*samp[k]/max(S,1), thr[k]/max(v,1), text[k]))
