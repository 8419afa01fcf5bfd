"""Build libanglican_pop.so (sm_100a) in-tree with nvcc.

nvcc cross-compiles without a GPU; the .so is git-ignored but travels to the
GPU box with the gpurun snapshot.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libanglican_pop.so")
SOURCES = ["ctx.cu", "comm.cu", "resample.cu", "dists.cu", "ssm.cu", "mvn.cu", "stats.cu", "dists2.cu", "dists3.cu"]
HEADERS = ["apop_internal.cuh", "exact_fold.cuh", "rng.cuh", "fabric.cuh", "fastmath.cuh", "fastmath_coef.inc", "gamma_draw.cuh",
           os.path.join("..", "..", "include", "anglican_pop.h")]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
         "--fmad=false",            # a*b+c must never fuse: exact-fold and parity arithmetic
         "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default",
         "-Xcudafe", "--diag_suppress=177"]
FLAGS += os.environ.get("APOP_NVCC_DEFS", "").split()      # tuning experiments (-DAPOP_...=..)


def _srcs():
    return [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def needs_build() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = _srcs() + [os.path.join(CSRC, h) for h in HEADERS]
    return any(os.path.getmtime(p) > t for p in deps if os.path.exists(p))


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return OUT
    objs = []
    bdir = os.path.join(HERE, "build")
    os.makedirs(bdir, exist_ok=True)
    procs = []
    for src in _srcs():
        obj = os.path.join(bdir, os.path.basename(src) + ".o")
        objs.append(obj)
        cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
        procs.append((cmd, subprocess.Popen(cmd)))
    for cmd, p in procs:
        if p.wait() != 0:
            raise RuntimeError("nvcc failed: " + " ".join(cmd))
    link = [NVCC, "-shared", "-o", OUT + ".tmp"] + objs + ["-ldl", "-cudart", "static"]
    subprocess.check_call(link)
    os.replace(OUT + ".tmp", OUT)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
