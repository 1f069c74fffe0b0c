"""In-tree build of ``libctcg_b200.so`` (the C-ABI library) with nvcc for sm_100a.

``python pytorch-asr_b200/build.py`` or ``__graft_entry__.build()``.  The
shared object is written next to this file so that it travels to the GPU box
with the repository snapshot (it is git-ignored, not gpurun-ignored).
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libctcg_b200.so")
STAMP = os.path.join(HERE, "build", "sources.sha256")

SOURCES = ["graph.cpp", "graph_abi.cpp", "abi.cu", "kernels_path.cu", "kernels_rows.cu",
           "kernels_aux.cu", "kernels_bisweep.cu", "kernels_num.cu", "kernels_post.cu"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
]

NVCC_FLAGS += os.environ.get("CTCG_EXTRA_NVCC", "").split()   # experiments (profiles/micro): -D overrides of kernel constants

# CTCG_TRACE_BUILD=1: compile the per-role cycle counters of bisweep_kernel in
# (profiles/micro/run_tim.sh, CTCG_SWEEP_DEBUG=256); off in the product build.
if os.environ.get("CTCG_TRACE_BUILD") == "1":
    NVCC_FLAGS.append("-DCTCG_SWEEP_TIMERS")  # per-role cycle counters of bisweep_kernel (CTCG_SWEEP_DEBUG=256)


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _digest():
    h = hashlib.sha256()
    names = sorted(os.listdir(CSRC)) + ["../../include/ctcg_b200.h"]
    for name in names:
        path = os.path.join(CSRC, name)
        if os.path.isfile(path):
            h.update(name.encode())
            with open(path, "rb") as f:
                h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build_lib(force=False, verbose=False):
    """Compile every source for sm_100a and link the shared library."""
    digest = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(STAMP):
        with open(STAMP) as f:
            if f.read().strip() == digest:
                return LIB
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    nvcc = _nvcc()
    objs = []
    log = []
    for src in SOURCES:
        path = os.path.join(CSRC, src)
        if not os.path.exists(path):
            continue
        obj = os.path.join(HERE, "build", src + ".o")
        cmd = [nvcc] + NVCC_FLAGS + ["-x", "cu", "-c", path, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log.append("$ " + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        if r.returncode != 0:
            sys.stderr.write(log[-1])
            raise RuntimeError("nvcc failed on %s" % src)
        objs.append(obj)
    cmd = [nvcc, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    log.append("$ " + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    if r.returncode != 0:
        sys.stderr.write(log[-1])
        raise RuntimeError("link failed")
    with open(os.path.join(HERE, "build", "build.log"), "w") as f:
        f.write("\n".join(log))
    with open(STAMP, "w") as f:
        f.write(digest)
    if verbose:
        sys.stdout.write("\n".join(log))
    return LIB


if __name__ == "__main__":
    build_lib(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
