"""Build libnsb200.so (CUDA kernels + C ABI) in-tree with nvcc for sm_100a.

    python -m navier_stokes_solver_b200.build        (or __graft_entry__.build())

Each .cu is compiled to an object in parallel and linked into ``libnsb200.so`` next to this file, so
the library travels with the source tree (no JIT cache).  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import concurrent.futures
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "_obj")
LIB = os.path.join(HERE, "libnsb200.so")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
         "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--expt-relaxed-constexpr"]
FLAGS += os.environ.get("NSB_NVCC_EXTRA", "").split()   # tuning sweeps: e.g. -DNSB_SELL_UNROLL=4


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _headers_digest():
    h = hashlib.sha1()
    inc = os.path.join(os.path.dirname(HERE), "include", "nsb200.h")
    for p in sorted([os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))] + [inc]):
        with open(p, "rb") as f:
            h.update(f.read())
    h.update(" ".join(FLAGS).encode())
    return h.hexdigest()


def _compile(src, digest, verbose, extra):
    obj = os.path.join(OBJ, src[:-3] + ".o")
    stamp = obj + ".stamp"
    with open(os.path.join(CSRC, src), "rb") as f:
        key = hashlib.sha1(f.read() + digest.encode() + " ".join(extra).encode()).hexdigest()
    if os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == key:
        return obj, ""
    cmd = [NVCC] + FLAGS + extra + ["-c", os.path.join(CSRC, src), "-o", obj]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, res.stdout, res.stderr))
    with open(stamp, "w") as f:
        f.write(key)
    return obj, res.stderr if verbose else ""


def build(verbose=False, ptxas_info=False):
    os.makedirs(OBJ, exist_ok=True)
    digest = _headers_digest()
    extra = ["-Xptxas", "-v"] if ptxas_info else []
    srcs = _sources()
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        results = list(ex.map(lambda s: _compile(s, digest, verbose or ptxas_info, extra), srcs))
    objs = [o for o, _ in results]
    for _, log in results:
        if log:
            print(log, file=sys.stderr)
    newest = max(os.path.getmtime(o) for o in objs)
    if not os.path.exists(LIB) or os.path.getmtime(LIB) < newest:
        cmd = [NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs + ["-ldl"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (res.stdout, res.stderr))
    return LIB


if __name__ == "__main__":
    print(build(verbose=True, ptxas_info="-v" in sys.argv))
