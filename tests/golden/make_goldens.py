"""Writes tests/golden/reference_goldens.json: the golden vectors the
reference's own tests hold for the hot path, transcribed with their source
file:line (values are literals from the reference's test files), plus exact
truths computed with numpy/scipy.  Run from the repo root:
    python tests/golden/make_goldens.py
"""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

g = {}
# test/anglican/runtime_test.cljc:41-46
g["mvn"] = [
    {"mean": [0, 0, 0], "cov": [[2, 0, 0], [0, 1, 0], [0, 0, 3]], "x": [3, 4, 5], "logpdf": -18.0694},
    {"mean": [0, 0, 0], "cov": [[2, 0, 0], [0, 1, 0], [0, 0, 3]], "x": [0, 0, 0], "logpdf": -3.6527},
    {"mean": [0, 0, 0], "cov": [[2, 0, 0], [0, 1, 0], [0, 0, 3]], "x": [10, 20, 30], "logpdf": -378.6527},
]
# test/anglican/runtime_chi_squared_test.cljc:76-85 (scipy.stats.chi2.pdf values quoted there)
g["chi_squared"] = [
    {"nu": 3.5, "x": 1.0, "pdf": 0.1962028031081274},
    {"nu": 3.5, "x": 5.31, "pdf": 0.079546575897983585},
    {"nu": 5.5, "x": 5.31, "pdf": 0.12068351943379794},
]
# test/anglican/runtime_test.cljc:18-24: (categorical '((x 1) (y 2)))
g["categorical"] = {"weights": [1.0, 2.0], "x_logp": float(np.log(1.0 / 3.0))}
# test/anglican/ipmcmc_test.clj:33-36
g["ipmcmc"] = {"log_zs": [-3, -4, -1, -8, -10, -5, -4, -3.4], "P": 4, "tolerance": 0.005,
               "true_p_sample": [0.22461, 0.11994, 0.24693, 0.00132, 0.00044, 0.06128, 0.14521,
                                 0.20026]}
# test/anglican/algorithm_test.cljc:102-151 (HMM benchmark)
g["hmm"] = {
    "observations": [0.9, 0.8, 0.7, 0.0, -0.025, -5.0, -2.0, -0.1, 0.0, 0.13, 0.45, 6, 0.2, 0.3, -1, -1],
    "init": [1.0, 1.0, 1.0],
    "trans": [[0.1, 0.5, 0.4], [0.2, 0.2, 0.6], [0.15, 0.15, 0.7]],
    "mu": [-1, 1, 0], "sd": [1, 1, 1],
    "posterior": [[0.3775, 0.3092, 0.3133], [0.0416, 0.4045, 0.5539], [0.0541, 0.2552, 0.6907],
                  [0.0455, 0.2301, 0.7244], [0.1062, 0.1217, 0.7721], [0.0714, 0.1732, 0.7554],
                  [0.9300, 0.0001, 0.0699], [0.4577, 0.0452, 0.4971], [0.0926, 0.2169, 0.6905],
                  [0.1014, 0.1359, 0.7626], [0.0985, 0.1575, 0.744], [0.1781, 0.2198, 0.6022],
                  [0.0000, 0.9848, 0.0152], [0.1130, 0.1674, 0.7195], [0.0557, 0.1848, 0.7595],
                  [0.2017, 0.0472, 0.7511], [0.2545, 0.0611, 0.6844]],
    "rms_threshold": 0.1,
    "exact_log_z": -44.42446961347636,     # forward algorithm
}
# test/anglican/algorithm_test.cljc:71-94 (gaussian benchmark)
g["gaussian"] = {"observations": [9.0, 8.0], "sigma": float(np.sqrt(2.0)), "mu0": 1.0,
                 "sigma0": float(np.sqrt(5.0)), "post_mean": 7.25,
                 "post_sd": float(np.sqrt(1.0 / 1.2)), "kl_threshold": 0.10}
# hand-checked 12-particle resample: literal walk of smc.cljc:105-137
w12 = [0.5, 0.0, 1.0, 0.25, 0.125, 0.0, 0.75, 1.0, 0.0625, 0.3, 0.9, 0.1]
u12 = 0.62
total = 0.0
for v in w12:
    total = total + v
step = total / 12.0
x, acc, i, out = total * u12, 0.0, 0, []
while len(out) < 12:
    nxt = acc + w12[i]
    if x < nxt:
        out.insert(0, i)          # conj onto a list prepends
        x = x + step
    else:
        acc = nxt
        i = (i + 1) % 12
g["resample12"] = {"weights": w12, "u": u12, "anc_ref": out, "total": total}

with open(os.path.join(HERE, "reference_goldens.json"), "w") as f:
    json.dump(g, f, indent=1)
print("wrote", os.path.join(HERE, "reference_goldens.json"))
