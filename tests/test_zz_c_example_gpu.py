"""examples/sv_sweep.c - a C99 program against include/anglican_pop.h alone - runs one sweep on
the GPU and reports the same log Z, to the last bit, as the Python mirror of `infer :smc` with
the same seed and observations: both are thin callers of the one C ABI."""
import re
import subprocess

import numpy as np
import pytest

from anglican_b200 import Population
from anglican_b200.models import StochasticVolatility
from tests.test_c_example_cpu import build_example

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("lg,T", [(12, 7), (20, 40)])
def test_c_program_and_python_mirror_agree_bit_for_bit(tmp_path, lg, T):
    exe = build_example(tmp_path)
    y = np.random.default_rng(20240602).normal(0.0, 0.7, T)
    yfile = tmp_path / "y.f64"
    y.astype("<f8").tofile(str(yfile))
    out = subprocess.run([exe, str(lg), str(T), str(yfile)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    m = re.search(r"log Z = (\S+), ESS at the last observe = (\S+), mean of the final states = (\S+), "
                  r"kernels launched = (\d+)", out.stdout)
    assert m, out.stdout
    with Population(1 << lg, seed=0xA11CE) as pop:
        res = pop.sweep(StochasticVolatility(), y, want_final=True, want_filt=False)
    assert float(m.group(1)) == res.log_z, (m.group(1), res.log_z)
    assert abs(float(m.group(2)) - res.ess[-1]) <= 0.06                   # printed with one decimal
    assert abs(float(m.group(3)) - res.x_final.mean()) < 2e-6             # printed with six decimals
    assert int(m.group(4)) > 0                                            # our kernels ran
