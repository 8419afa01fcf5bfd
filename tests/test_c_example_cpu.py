"""The drop-in boundary is a C header: it must parse as plain C99 (what jextract / a JNI stub
sees) and as C++, and a C program written against it alone must link against the library.
Without a CUDA device that program has to fail loudly (APOP_ERR_CUDA), never compute."""
import os
import subprocess

import pytest

from anglican_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INC = os.path.join(ROOT, "include")
LIBDIR = os.path.join(ROOT, "anglican_b200")


def build_example(tmpdir):
    L.load()                                      # builds the library if it is stale
    exe = os.path.join(str(tmpdir), "sv_sweep")
    subprocess.check_call(["gcc", "-std=c99", "-pedantic", "-Wall", "-Wextra", "-Werror", "-I", INC,
                           os.path.join(ROOT, "examples", "sv_sweep.c"), "-o", exe,
                           "-L", LIBDIR, "-langlican_pop", "-Wl,-rpath," + LIBDIR, "-lm"])
    return exe


@pytest.mark.parametrize("cc,std,lang", [("gcc", "-std=c99", "c"), ("gcc", "-std=c11", "c"),
                                         ("g++", "-std=c++11", "c++"), ("g++", "-std=c++17", "c++")])
def test_header_is_plain_c_and_cxx(tmp_path, cc, std, lang):
    src = tmp_path / "probe.c"
    src.write_text('#include "anglican_pop.h"\n'
                   "int probe(void) { apop_sweep_opts o; apop_arg a; a.col = -1; a.scalar = 0.0;\n"
                   "  o.ess_threshold = a.scalar; return (int)sizeof(o) + APOP_SSM_SV + APOP_F64; }\n")
    subprocess.check_call([cc, std, "-pedantic", "-Wall", "-Wextra", "-Werror", "-I", INC,
                           "-x", lang, "-fsyntax-only", str(src)])


def test_c_example_links_and_fails_loudly_without_a_device(tmp_path):
    exe = build_example(tmp_path)
    lib = L.load()
    out = subprocess.run([exe, "10", "5"], capture_output=True, text=True, timeout=120)
    if lib.apop_device_count() == 0:
        assert out.returncode == 3, (out.returncode, out.stderr)          # APOP_ERR_CUDA
        assert "apop_create" in out.stderr and "log Z" not in out.stdout
    else:
        assert out.returncode == 0 and "log Z" in out.stdout, out.stderr
    bad = subprocess.run([exe, "40", "5"], capture_output=True, text=True, timeout=120)
    assert bad.returncode == 2 and "usage" in bad.stderr
