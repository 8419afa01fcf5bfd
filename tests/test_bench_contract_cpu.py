"""bench.py's reference arm runs without a GPU: its JSON line must carry the
keys the driver reads (same metric / unit / config as our arm)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*extra):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference",
                          "--steps", "1", "--warmup", "0", *extra],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1                      # ONE JSON line
    return json.loads(lines[0])


def test_reference_arm_line_sv():
    d = _run("--T", "12")
    assert d["impl"] == "reference" and d["metric"] == "smc_particle_steps_per_sec"
    assert d["unit"] == "particle-steps/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["n_gpus"] == 1 and d["steps"] == 1 and d["warmup"] == 0
    assert d["vs_baseline"] is None and d["dtype"] == "f64" and d["data"] == "synthetic"
    assert "configs[2]" in d["config"]["workload"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    e = d["e2e"]
    assert e["value"] == d["value"] and e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0


def test_reference_arm_line_lgmvn64():
    d = _run("--workload", "lgmvn64", "--T", "3")
    assert "configs[4]" in d["config"]["workload"] and d["value"] > 0


def test_workload_generators_are_deterministic():
    sys.path.insert(0, ROOT)
    import bench
    import numpy as np
    assert np.array_equal(bench.sv_observations(50), bench.sv_observations(50))
    a, b = bench.lgmvn_workload(4, 8), bench.lgmvn_workload(4, 8)
    assert all(np.array_equal(x, y) for x, y in zip(a, b))
    A, Qc, R, y = a
    assert y.shape == (32,) and np.allclose(A @ A.T, 0.81 * np.eye(8))       # A = 0.9 Q, Q orthogonal


def test_reference_arm_line_lgssm1d_and_kalman_helper():
    d = _run("--workload", "lgssm1d", "--T", "10")
    assert "configs[1]" in d["config"]["workload"] and d["value"] > 0
    sys.path.insert(0, ROOT)
    import bench
    from oracle import pyoracle as orc
    y = bench.lgssm_observations(200)
    m = orc.Model(orc.SSM_LGSSM1D, m0=0.0, s0=1.0, a=0.9, b=0.0, q=1.0, c=1.0, d=0.0, r=1.0)
    assert abs(bench.kalman_logz(y) - orc.kalman1d(m, y)[0]) < 1e-9


def test_recorded_log_z_file_is_well_formed():
    ref = json.load(open(os.path.join(ROOT, "profiles", "logz_reference.json")))
    assert "sv/n134217728/T2000" in ref
    for key, vals in ref.items():
        parts = key.split("/")
        wl, n, T = parts[:3]
        assert wl in ("sv", "lgssm1d", "lgmvn64") and n.startswith("n") and T.startswith("T")
        if len(parts) == 3:                # device-timed leg: log Z per population and sweep
            assert all(k.startswith("f") and isinstance(v, float) for k, v in vals.items())
        else:                              # e2e leg: log Z and the final-state checksum per sweep
            assert parts[3:] == ["e2e"]
            for k, v in vals.items():
                if k.startswith("logz_f"):
                    assert isinstance(v, float)
                else:
                    assert k.startswith("xsum_f") and isinstance(v, int) and 0 <= v < 2 ** 64
                    assert "logz_" + k[5:] in vals
    # the default configuration has every timed sweep of a 20-step run on file, both legs
    assert len(ref["sv/n134217728/T2000"]) >= 60
    assert len(ref["sv/n134217728/T2000/e2e"]) >= 120
