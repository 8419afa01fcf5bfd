import csv,sys
rows=list(csv.reader(open(sys.argv[1])))
cur=None; out=[]
for r in rows:
    if len(r)==2 and r[0]=='File Path': cur=r[1].split('/')[-1]; continue
    if len(r)>7 and r[0].strip().isdigit():
        try: out.append((int(float(r[7] or 0)),cur,r[0],r[1].strip()[:90]))
        except ValueError: pass
tot=sum(o[0] for o in out)
print('total',tot)
for n,f,l,s in sorted(out,reverse=True)[:int(sys.argv[2]) if len(sys.argv)>2 else 40]:
    print(f'{100*n/tot:5.1f}% {n:11d} {f}:{l} {s}')
