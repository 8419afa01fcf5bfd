"""Per-source-line executed-instruction / stall totals from `ncu --page source --csv --print-source cuda,sass`."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
per = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0      # particles (x32) per launch for normalisation
cur = None; line = None; agg = {}; txt = {}
for r in rows:
    if len(r) == 2 and r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if len(r) < 8: continue
    if r[0].strip().isdigit():
        line = (cur, int(r[0])); txt[line] = r[1].strip()[:70]
    elif r[0] == '' and r[2].startswith('0x') and line:
        try:
            a = agg.setdefault(line, [0, 0]); a[0] += int(r[7]); a[1] += int(r[4])
        except ValueError: pass
tot = sum(v[0] for v in agg.values()); ts = sum(v[1] for v in agg.values())
print('total warp-inst', tot, 'per particle', tot / per if per > 1 else '')
files = {}
for k, v in agg.items():
    f = files.setdefault(k[0], [0, 0]); f[0] += v[0]; f[1] += v[1]
for f, v in sorted(files.items(), key=lambda kv: -kv[1][0]): print(f'  {f:28s} inst {100*v[0]/tot:5.1f}% stall {100*v[1]/ts:5.1f}%')
n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:n]:
    print(f'{100*v[0]/tot:5.1f}% st{100*v[1]/ts:5.1f}% {k[0]}:{k[1]} {txt[k]}')
