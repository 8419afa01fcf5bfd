"""Summarise an `ncu --page source --csv` dump: stall samples by opcode and the hottest SASS lines."""
import csv, sys
from collections import Counter
rows = list(csv.reader(open(sys.argv[1])))
data = [r for r in rows if len(r) > 5 and r[2].isdigit()]
tot = sum(int(r[2]) for r in data)
print('total samples', tot, 'instrs', len(data))
c = Counter(); ex = Counter()
for r in data:
    toks = r[1].split()
    op = toks[1] if toks[0].startswith('@') else toks[0]
    c[op.split('.')[0]] += int(r[2]); ex[op.split('.')[0]] += int(r[5])
te = sum(ex.values())
for k, v in c.most_common(18):
    print(f'{k:10s} stall {100*v/tot:5.1f}%   executed {100*ex[k]/te:5.1f}%')
print()
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
top = sorted(range(len(data)), key=lambda i: -int(data[i][2]))[:n]
for i in sorted(top):
    print(i, f'{100*int(data[i][2])/tot:5.1f}%', data[i][1].strip()[:90], 'exec', data[i][5])
