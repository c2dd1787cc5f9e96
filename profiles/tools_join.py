import csv, re, sys, collections
rep=sys.argv[1]; kern=sys.argv[2]
import subprocess
out=subprocess.run(["ncu","-i",rep,"--page","source","--csv"],capture_output=True,text=True).stdout.splitlines()
rows=list(csv.reader(out))
hi=[i for i,r in enumerate(rows) if r and r[0]=="Address"][0]
h=rows[hi]; ci=h.index("Instructions Executed"); ti=h.index("Thread Instructions Executed")
data=[r for r in rows[hi+1:] if len(r)>=len(h)-2 and r[0].startswith("0x")]
# disassembly with line info
dis=open('/tmp/cub/dis.txt').read().splitlines()
start=[i for i,l in enumerate(dis) if l.startswith('.text.') and kern in l][0]
cur=None; lines=[]
for l in dis[start+1:]:
    if l.startswith('//---------------------'): break
    m=re.search(r'//## File "([^"]+)", line (\d+)(.*)',l)
    if m:
        cur=(m.group(1).split('/')[-1],int(m.group(2)),m.group(3)); continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/',l):
        lines.append(cur)
print("sass instrs: ncu",len(data),"nvdisasm",len(lines))
agg=collections.Counter(); aggt=collections.Counter()
tot=0
for r,src in zip(data,lines):
    c=float(r[ci] or 0); t=float(r[ti] or 0)
    key=(src[0],src[1]) if src else ("?",0)
    agg[key]+=c; aggt[key]+=t; tot+=c
print("total warp instrs",tot)
srcs={}
def line(f,n):
    import os
    for p in ('/root/repo/cron_operator_b200/csrc/'+f,):
        if os.path.exists(p):
            if p not in srcs: srcs[p]=open(p).read().splitlines()
            return srcs[p][n-1].strip()[:100] if n-1<len(srcs[p]) else ''
    return ''
for (f,n),c in agg.most_common(50):
    print(f"{c/tot*100:5.1f}% thr/inst={aggt[(f,n)]/max(c,1):5.1f} {f}:{n}  {line(f,n)}")
