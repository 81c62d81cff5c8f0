import csv, subprocess, sys, io, collections
rep = sys.argv[1]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
open('/tmp/src_page.csv','w').write(txt)
rows = list(csv.reader(io.StringIO(txt)))
fname=None; hdr=None
smp=collections.Counter(); inst=collections.Counter(); thr=collections.Counter()
lines=[]
for r in rows:
    if len(r) >= 2 and r[0] in ("File Name", "File Path"): fname = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No": hdr = {n: i for i, n in enumerate(r)}; continue
    if not hdr or len(r) < 10 or not r[0].isdigit(): continue
    try:
        ie=float(r[hdr["Instructions Executed"]] or 0); te=float(r[hdr["Thread Instructions Executed"]] or 0); s=float(r[hdr["# Samples"]] or 0)
    except ValueError: continue
    lines.append((fname,int(r[0]),ie,te,s, {k: float(r[i] or 0) for k,i in hdr.items() if k.startswith('stall_')}))
tot_s=sum(l[4] for l in lines); tot_i=sum(l[2] for l in lines)
print(sorted(hdr.keys()))
byf=collections.defaultdict(lambda:[0,0,0])
for f,ln,ie,te,s,st in lines:
    byf[f][0]+=s; byf[f][1]+=ie; byf[f][2]+=te
for f,(s,ie,te) in sorted(byf.items(), key=lambda x:-x[1][0]):
    print("%-28s smp %5.1f%% inst %5.1f%% thr/inst %.1f"%(f,100*s/tot_s,100*ie/tot_i,te/max(ie,1)))
regions = {}   # optional: name -> [(file, first line, last line)]
for name,rs in regions.items():
    s=ie=te=0; st=collections.Counter()
    for f,ln,i,t,sm,sd in lines:
        for (rf,a,b) in rs:
            if f==rf and a<=ln<=b: s+=sm; ie+=i; te+=t; st.update(sd)
    top=", ".join("%s %.0f%%"%(k.replace('stall_',''),100*v/max(s,1)) for k,v in st.most_common(3))
    print("%-60s smp %5.1f%% inst %5.1f%% thr/inst %4.1f  [%s]"%(name,100*s/tot_s,100*ie/tot_i,te/max(ie,1),top))
