import csv, collections, sys, subprocess
rep, kern, skip = sys.argv[1], sys.argv[2], sys.argv[3]
out = subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","cuda,sass","--kernel-name","regex:"+kern,"--launch-skip",skip,"--launch-count","1"],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
cur=None; agg=collections.defaultdict(int); hdr=None
for r in rows:
    if r and r[0]=="File Path": cur=r[1]; continue
    if r and r[0]=="Line No": hdr=r; continue
    if r and r[0]=="Function Name": continue
    if cur and hdr and len(r)==len(hdr):
        d=dict(zip(hdr,r))
        try: s=int(d["# Samples"])
        except: continue
        if d["Line No"]: agg[(cur.split('/')[-1],int(d["Line No"]))]+=s
tot=sum(agg.values()); print("total samples",tot)
src={}
for (f,l),s in sorted(agg.items(), key=lambda x:-x[1])[:int(sys.argv[4]) if len(sys.argv)>4 else 25]:
    if f not in src:
        try: src[f]=open('/root/repo/dsp-slam-rgbd_b200/csrc/'+f).read().split('\n')
        except: src[f]=[]
    line=src[f][l-1].strip()[:110] if src[f] and l-1<len(src[f]) else ''
    print("%5.1f%% %s:%d  %s"%(100*s/tot,f,l,line))
