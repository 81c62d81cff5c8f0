"""Stall-sample share per source file and top lines of a kernel, from an .ncu-rep captured with --import-source on.
usage: python tools/ncu_samples.py report.ncu-rep [top_n]"""
import csv, subprocess, io, collections, sys
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 30
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
fname = None; hdr = None; out = []; per = collections.Counter()
for r in rows:
    if len(r) >= 2 and r[0] in ("File Name", "File Path"): fname = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No": hdr = {n: i for i, n in enumerate(r)}; continue
    if not hdr or len(r) < 10 or not r[0].isdigit(): continue
    try: smp = float(r[hdr["# Samples"]] or 0)
    except ValueError: continue
    out.append((smp, fname, int(r[0]), r[1].strip()[:100])); per[fname] += smp
tot = sum(o[0] for o in out)
for f, v in per.most_common(10): print("%-24s %5.1f%%" % (f, 100 * v / tot))
out.sort(reverse=True)
for smp, f, ln, src in out[:topn]: print("%5.1f%% %s:%d %s" % (100 * smp / tot, f, ln, src))
