"""The reference-side sources a maintainer would add (clj/anglican/gpu/*.clj):
they cannot be compiled here (no JVM), so what CAN be checked is checked - the generated Panama
binding is in sync with include/anglican_pop.h and binds every exported symbol, every handle the
hand-written namespaces use exists, the struct offsets they hard-code are the C ABI's, the files
say that they are unverified, and their brackets balance."""
import ctypes as C
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
CLJ = os.path.join(ROOT, "clj", "anglican", "gpu")
FILES = ["ffi.clj", "runtime.clj", "smc.clj", "pimh.clj", "pgibbs.clj", "ipmcmc.clj"]


def _read(name):
    return open(os.path.join(CLJ, name)).read()


def test_generated_ffi_is_in_sync_and_binds_every_symbol():
    import gen_clj_ffi as g
    from anglican_b200 import _lib as L
    text = _read("ffi.clj")
    assert text == g.render(), "run `python tools/gen_clj_ffi.py` after changing the header"
    bound = re.findall(r'\(downcall "(apop_[a-z0-9_]+)"', text)
    assert sorted(bound) == sorted(L.DECLARED_SYMBOLS)
    assert len(set(bound)) == len(bound)
    # argument counts agree with the ctypes binding, symbol by symbol
    sigs = {name: args for name, _, args in L._SIGS}
    for name, _ret, params in g.declarations():
        assert len(params) == len(sigs[name]), name


def test_hand_written_namespaces_use_only_existing_handles():
    defined = set(re.findall(r"^\(def(?:n-?)? ([a-zA-Z\-\*\?!]+)", _read("ffi.clj"), flags=re.M))
    for f in FILES[1:]:
        text = _read(f)
        assert "UNVERIFIED: no JVM" in text.splitlines()[0], f
        for used in set(re.findall(r"\bffi/([a-zA-Z\-\*\?!]+)", text)):
            assert used in defined, (f, used)
    # the infer methods of the five algorithms behind inference.cljc:25-28
    methods = "".join(_read(f) for f in FILES[2:])
    for alg in (":smc", ":pimh", ":pgibbs", ":pgas", ":ipmcmc"):
        assert f"(defmethod infer {alg} " in methods, alg


def test_struct_offsets_in_smc_clj_are_the_c_abi():
    from anglican_b200 import _lib as L
    text = _read("smc.clj")
    o = L.ApopSweepOpts
    field = dict(re.findall(r":([a-z\-]+) (\d+)", re.search(r"field \{([^}]*)\}", text).group(1)))
    for k, off in field.items():
        assert getattr(o, k.replace("-", "_")).offset == int(off), k
    assert o.want_final.offset == 20 and "(.set o ValueLayout/JAVA_INT 20 (int 1))" in text
    assert L.ApopSweepResult.x_final.offset == 32 and "(.set r ValueLayout/ADDRESS 32 xs)" in text
    assert L.ApopSSM.m0.offset == 8 and L.ApopSSM.hmm_init.offset == 80
    assert C.sizeof(L.ApopArg) == 16 and L.ApopArg.scalar.offset == 8


def test_brackets_balance():
    for f in FILES:
        text = re.sub(r'\\[()\[\]{}]', "", _read(f))             # character literals
        text = re.sub(r'"(?:[^"\\]|\\.)*"', '""', text, flags=re.S)  # strings
        text = re.sub(r";.*", "", text)                          # comments
        stack = []
        pairs = {")": "(", "]": "[", "}": "{"}
        for ch in text:
            if ch in "([{":
                stack.append(ch)
            elif ch in pairs:
                assert stack and stack.pop() == pairs[ch], f
        assert not stack, f
