import csv, re, sys, collections, subprocess
rep=sys.argv[1]; kern=sys.argv[2]
out=subprocess.run(["ncu","-i",rep,"--page","source","--csv"],capture_output=True,text=True).stdout.splitlines()
rows=list(csv.reader(out))
hi=[i for i,r in enumerate(rows) if r and r[0]=="Address"][0]
h=rows[hi]; ci=h.index("Instructions Executed"); ti=h.index("Thread Instructions Executed")
data=[r for r in rows[hi+1:] if r and r[0].startswith("0x")]
dis=open('/tmp/cub/dis.txt').read().splitlines()
start=[i for i,l in enumerate(dis) if l.startswith('.text.') and kern in l][0]
cur=None; lines=[]
for l in dis[start+1:]:
    if l.startswith('//---------------------'): break
    m=re.search(r'//## File "([^"]+)", line (\d+)',l)
    if m: cur=(m.group(1).split('/')[-1],int(m.group(2))); continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/',l): lines.append(cur)
# function ranges from source
import bisect
def ranges(path):
    src=open(path).read().splitlines(); marks=[]
    for i,l in enumerate(src,1):
        m=re.match(r'\s*(?:CW_HD|__device__|__global__|SlowResult|template).*?\b([A-Za-z_0-9]+)\s*\(',l)
        if m and not l.strip().startswith('//') and ('{' in l or l.rstrip().endswith(',') or l.rstrip().endswith(')')):
            marks.append((i,m.group(1)))
    return marks
R={'cron_walk.cuh':ranges('cron_operator_b200/csrc/cron_walk.cuh'),'unmet_walk.cuh':ranges('cron_operator_b200/csrc/unmet_walk.cuh'),'kernels.cuh':ranges('cron_operator_b200/csrc/kernels.cuh')}
def fn(f,n):
    if f not in R: return f
    ms=R[f]; i=bisect.bisect_right([m[0] for m in ms],n)-1
    return f.split('.')[0]+':'+(ms[i][1] if i>=0 else '?')
agg=collections.Counter(); aggt=collections.Counter(); tot=0
for r,src in zip(data,lines):
    c=float(r[ci] or 0); t=float(r[ti] or 0)
    k=fn(*src) if src else '?'
    agg[k]+=c; aggt[k]+=t; tot+=c
print("total",tot)
for k,c in agg.most_common(25): print(f"{c/tot*100:5.1f}%  thr/inst={aggt[k]/max(c,1):5.1f}  {k}")
