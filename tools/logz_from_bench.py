"""Merge the per-seed log Z values of 1-GPU bench lines into profiles/logz_reference.json
(bench.py compares every later run, at any number of GPUs, against them).

    python tools/logz_from_bench.py gpurun_out/rNN/bench_default.json [more lines ...]
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles", "logz_reference.json")


def main():
    try:
        ref = json.load(open(OUT))
    except Exception:
        ref = {}
    for path in sys.argv[1:]:
        d = json.loads(open(path).read().strip().splitlines()[-1])
        if d.get("n_gpus") != 1 or "log_zs" not in d:
            print("skipped", path)
            continue
        c = d["config"]
        wl = "lgmvn64" if "64-dim" in c["model"] else "lgssm1d" if "1-D linear" in c["model"] else "sv"
        key = "%s/n%d/T%d" % (wl, c["particles_total"], c["T"])
        ref.setdefault(key, {}).update(d["log_zs"])
        print(key, len(d["log_zs"]), "values")
    json.dump(ref, open(OUT, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
