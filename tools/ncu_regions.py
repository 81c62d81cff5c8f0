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
regions = {
 'locate loop (kernels.cu 421-431)': [('gw_kernels.cu',421,431)],
 'suffix_link+rank1+rankN (fm 56-92)': [('gw_fm.cuh',56,92)],
 'rank4/halves/masks (fm 31-54)': [('gw_fm.cuh',31,54)],
 'load_block (fm 13-18)': [('gw_fm.cuh',13,18)],
 'dp columns (verify 45-130)': [('gw_verify.cuh',28,141)],
 'traceback (verify 143-214)': [('gw_verify.cuh',142,223)],
 'fused: begin_dp/job (fused 66-205)': [('gw_fused.cuh',1,230)],
 'queue ops (search 43-110)': [('gw_search.cuh',43,110)],
 'score/numsplit (search 24-42)': [('gw_search.cuh',24,42)],
 'next_si (search 158-206)': [('gw_search.cuh',158,206)],
 'next_state (search 208-314)': [('gw_search.cuh',208,314)],
 'split/update/hits (search 316-430)': [('gw_search.cuh',315,430)],
 'search_poll (automaton 227-265)': [('gw_automaton.cuh',227,265)],
 'search_try (automaton 290-365)': [('gw_automaton.cuh',290,365)],
 'resume/report (automaton 1-226)': [('gw_automaton.cuh',1,226)],
 'search_split (automaton 267-288)': [('gw_automaton.cuh',266,289)],
 'state.cuh accessors 114-136': [('gw_state.cuh',97,137)],
 'state.cuh alloc/promote 150-205': [('gw_state.cuh',138,206)],
 'state.cuh substring/match_window/pattern/staircase 207-268': [('gw_state.cuh',207,268)],
 'kernels.cu flow loop': [('gw_kernels.cu',390,547)],
}
for name,rs in regions.items():
    s=ie=te=0; st=collections.Counter()
    for f,ln,i,t,sm,sd in lines:
        for (rf,a,b) in rs:
            if f==rf and a<=ln<=b: s+=sm; ie+=i; te+=t; st.update(sd)
    top=", ".join("%s %.0f%%"%(k.replace('stall_',''),100*v/max(s,1)) for k,v in st.most_common(3))
    print("%-60s smp %5.1f%% inst %5.1f%% thr/inst %4.1f  [%s]"%(name,100*s/tot_s,100*ie/tot_i,te/max(ie,1),top))
