"""Builds libdaala_tdlt_cuda.so (sm_100a) in-tree with nvcc.  No JIT cache: the
.so lies next to the sources so it travels with the repo snapshot."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libdaala_tdlt_cuda.so")
SOURCES = ["tdlt_plan.cu", "tdlt_fwd_region.cu", "tdlt_inv_region.cu", "tdlt_aux_kernels.cu", "tdlt_dering.cu", "tdlt_input.cu", "tdlt_capi.cu", "tdlt_static.cu"]
HEADERS = ["tdlt_device.cuh", "tdlt_region.cuh", "tdlt_tasks.cuh", "tdlt_lift_gen.cuh", "tdlt_kernels.h",
           "tdlt_aux_kernels.h", os.path.join("..", "..", "include", "daala_tdlt.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC"]


def source_hash():
    """SHA-256 over every file the library is built from (all of csrc/ and include/) and the
    compiler flags: the library is rebuilt whenever it differs from the sidecar written next
    to the .so, so a stale binary cannot survive a header-only edit."""
    import hashlib
    h = hashlib.sha256(" ".join(NVCC_FLAGS).encode())
    inc = os.path.join(HERE, "..", "include")
    files = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cu", ".cuh", ".h"))]
    files += [os.path.join(inc, f) for f in sorted(os.listdir(inc)) if f.endswith(".h")]
    for f in files:
        h.update(os.path.basename(f).encode())
        h.update(open(f, "rb").read())
    return h.hexdigest()


STAMP = LIB + ".sha256"


def _stale():
    if not os.path.exists(LIB) or not os.path.exists(STAMP):
        return True
    return open(STAMP).read().strip() != source_hash()


def build(force=False, verbose=False, trace=False):
    """trace=True builds libdaala_tdlt_cuda_trace.so with -DTDLT_TRACE -rdc (phase
    timestamps for tools/trace_phases.py); never used by tests or bench."""
    if trace:
        return _build_trace()
    if not force and not _stale():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objs = []
    procs = []
    for src in SOURCES:
        obj = os.path.join(CSRC, src.replace(".cu", ".o"))
        objs.append(obj)
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
              ["-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for cmd, p in procs:
        out = p.communicate()[0].decode()
        if verbose or p.returncode:
            sys.stderr.write(out)
        if p.returncode:
            raise RuntimeError("nvcc failed: " + " ".join(cmd))
    cmd = [nvcc, "-shared", "-o", LIB] + objs + ["-lcudart_static" if False else "-lcudart"]
    subprocess.check_call(cmd)
    with open(STAMP, "w") as f:
        f.write(source_hash() + "\n")
    return LIB


def _build_trace():
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    lib = os.path.join(HERE, "libdaala_tdlt_cuda_trace.so")
    cmd = [nvcc] + NVCC_FLAGS + ["-DTDLT_TRACE", "-rdc=true", "-shared", "-o", lib] + \
          [os.path.join(CSRC, f) for f in SOURCES] + ["-lcudart"]
    subprocess.check_call(cmd)
    return lib


if __name__ == "__main__":
    if "--trace" in sys.argv:
        print(build(trace=True))
        sys.exit(0)
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
