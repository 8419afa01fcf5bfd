"""profiles/ncu_traffic.json from an ncu summary (tools/ncu_summary.py output):
dram__bytes_read.sum + dram__bytes_write.sum per launch and per particle, the figure
bench.py reports as roofline.traffic.

    python tools/ncu_traffic.py gpurun_out/rNN/summary.csv 134217728 "label" > profiles/ncu_traffic.json
"""
import csv
import json
import sys

SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    n = int(sys.argv[2])
    hdr, units = rows[0], rows[1]
    ir, iw, it = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("gpu__time_duration.sum")
    out = {}
    for r in rows[2:]:
        name = r[0].replace("void ", "").split("<")[0].strip()
        rd = float(r[ir]) * SCALE[units[ir]]
        wr = float(r[iw]) * SCALE[units[iw]]
        e = out.setdefault(name, {"launches": 0, "dram_bytes": 0.0, "read": 0.0, "write": 0.0})
        e["launches"] += 1
        e["dram_bytes"] += rd + wr
        e["read"] += rd
        e["write"] += wr
    kernels = {k: {"dram_bytes_per_particle": v["dram_bytes"] / v["launches"] / n,
                   "dram_read_bytes_per_launch": v["read"] / v["launches"],
                   "dram_write_bytes_per_launch": v["write"] / v["launches"],
                   "launches_captured": v["launches"]} for k, v in out.items()}
    json.dump({"source": sys.argv[3] if len(sys.argv) > 3 else sys.argv[1], "particles": n,
               "kernels": kernels}, sys.stdout, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
