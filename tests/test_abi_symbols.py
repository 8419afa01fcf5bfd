"""The C-ABI library loads on a machine without a GPU and exports every symbol
include/anglican_pop.h declares; compute entry points refuse to run without a
device (no CPU fallback)."""
import ctypes as C
import os
import re

from anglican_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    hdr = open(os.path.join(ROOT, "include", "anglican_pop.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(apop_[a-z0-9_]+)\s*\(", hdr)))


def test_every_declared_symbol_is_exported_and_bound():
    lib = L.load()
    names = _declared()
    assert len(names) >= 40
    for n in names:
        assert hasattr(lib, n), f"{n} declared in anglican_pop.h but not exported"
    assert sorted(L.DECLARED_SYMBOLS) == names, "python binding out of sync with the header"


def test_no_device_means_loud_failure():
    lib = L.load()
    assert lib.apop_version() >= 100
    if lib.apop_device_count() == 0:
        ctx = C.c_void_p()
        rc = lib.apop_create(C.byref(ctx), 1024, 1, 0, 0, 1, None)
        assert rc == -2 and not ctx.value          # APOP_ERR_CUDA, never a CPU fallback


def test_host_scalar_entry_points():
    lib = L.load()
    assert lib.apop_pimh_accept(-5.0, float("-inf"), 0.99) == 1     # lmh.cljc:146-151
    assert lib.apop_pimh_accept(-3.0, -2.0, 0.5) == 0
    import numpy as np
    from oracle import pyoracle as orc
    lz = np.array([-3, -4, -1, -8, -10, -5, -4, -3.4])
    rng = np.random.default_rng(5)
    for _ in range(200):
        u = rng.random(4)
        cs = np.empty(4, np.int32)
        zs = np.empty(8)
        rc = lib.apop_ipmcmc_gibbs(lz.ctypes.data_as(L.c_double_p), 8, 4,
                                   u.ctypes.data_as(L.c_double_p),
                                   cs.ctypes.data_as(L.c_int32_p), zs.ctypes.data_as(L.c_double_p))
        assert rc == 0
        rcs, rzs = orc.ipmcmc_gibbs(lz, 4, u)
        assert np.array_equal(cs, rcs) and np.array_equal(zs, rzs)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "anglican_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in src.replace("anglican_oracle", "oracle") or \
                    not re.search(r"(import\s+oracle|from\s+oracle|oracle/|#include.*oracle)", src), f
