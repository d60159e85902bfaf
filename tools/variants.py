#!/usr/bin/env python3
"""Builds experiment variants of the library: tools/variants.py name=-DFLAG[,-DFLAG..] ...
-> daala_b200/variants/lib_<name>.so (git-ignored; they travel to the GPU box)."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from daala_b200 import build as b
out = os.path.join(ROOT, "daala_b200", "variants")
os.makedirs(out, exist_ok=True)
procs = []
for spec in sys.argv[1:]:
    name, _, flags = spec.partition("=")
    lib = os.path.join(out, "lib_%s.so" % name)
    cmd = ["/usr/local/cuda/bin/nvcc"] + b.NVCC_FLAGS + [f for f in flags.split(",") if f] + \
          ["-shared", "-o", lib] + [os.path.join(b.CSRC, f) for f in b.SOURCES] + ["-lcudart"]
    procs.append((name, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
for name, p in procs:
    o = p.communicate()[0].decode()
    print(name, "rc=%d" % p.returncode, o[-300:] if p.returncode else "")
