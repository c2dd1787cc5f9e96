import csv, re, sys, subprocess
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
# print in SASS order, compressing runs with same exec count
prev=None
for idx,(r,src) in enumerate(zip(data,lines)):
    c=int(float(r[ci] or 0)); t=float(r[ti] or 0)
    key=(c, round(t/max(c,1)))
    if key!=prev:
        print(f"{idx:5d} exec={c:9d} thr/inst={t/max(c,1):5.1f}  {src}  {r[1][:60]}")
        prev=key
