"""Static SASS opcode histogram per kernel of a .so: python tools/sass_static.py lib.so pattern"""
import subprocess,sys,re,collections
out=subprocess.run(["cuobjdump","-sass",sys.argv[1]],capture_output=True,text=True).stdout
cur=None; fn={}
for line in out.splitlines():
    m=re.search(r"Function : (\S+)",line)
    if m: cur=m.group(1); fn[cur]=[]; continue
    m=re.match(r"\s+/\*[0-9a-f]+\*/\s+(.*?);",line)
    if m and cur: fn[cur].append(m.group(1).strip())
for k,v in fn.items():
    if len(sys.argv)>2 and sys.argv[2] not in k: continue
    h=collections.Counter()
    for ins in v:
        op=re.sub(r"^@!?U?P\w+\s+","",ins).split()[0].split('.')[0]
        h[op]+=1
    print(k[:70],len(v)); print('  ',', '.join(f"{a}:{b}" for a,b in h.most_common(18)))
    if len(sys.argv)>3:
        for ins in v:
            if sys.argv[3] in ins: print('     ',ins)
