"""Build the oracle's C restatement into oracle/_build/liboracle.so.

TEST INFRASTRUCTURE: building the checker is not using it.  Called from
__graft_entry__.build() and lazily from oracle.pyoracle.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT_DIR = os.path.join(HERE, "_build")
OUT = os.path.join(OUT_DIR, "liboracle.so")
SRC = [os.path.join(HERE, "anglican_oracle.c")]
HDR = [os.path.join(HERE, "anglican_oracle.h")]


def needs_build() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(p) > t for p in SRC + HDR)


def build(force: bool = False) -> str:
    if not force and not needs_build():
        return OUT
    os.makedirs(OUT_DIR, exist_ok=True)
    cmd = ["gcc", "-O2", "-std=c11", "-ffp-contract=off", "-fno-fast-math", "-fopenmp",
           "-fPIC", "-shared", "-Wall", "-Wextra", "-Wno-unused-parameter",
           "-o", OUT + ".tmp"] + SRC + ["-lm"]
    subprocess.check_call(cmd)
    os.replace(OUT + ".tmp", OUT)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
