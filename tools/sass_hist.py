import csv,sys,collections,re
f=sys.argv[1]
rows=list(csv.reader(open(f)))
hdr=rows[1]; ia=hdr.index("Source"); ie=hdr.index("Instructions Executed")
tot=0; h=collections.Counter()
for r in rows[2:]:
    if len(r)<=ie: continue
    s=r[ia].strip(); n=int(float(r[ie] or 0))
    m=re.match(r'(@!?U?P\d+\s+)?([A-Z0-9_.]+)',s)
    op=m.group(2) if m else s
    op='.'.join(op.split('.')[:2]) if op.startswith(('I2F','F2I','MUFU')) else op.split('.')[0]
    h[op]+=n; tot+=n
print('total',tot)
for k,v in h.most_common(40): print(f'{k:12s} {v:12d} {100*v/tot:5.1f}%')
