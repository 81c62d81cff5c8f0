"""Top source lines of a kernel by executed warp instructions, from an .ncu-rep captured with --import-source on.
usage: python tools/ncu_lines.py report.ncu-rep [top_n]"""
import csv, subprocess, sys, io, collections
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
fname = None; hdr = None; out = []; per_file = collections.Counter(); thr_file = collections.Counter()
for r in rows:
    if len(r) >= 2 and r[0] in ("File Name", "File Path"): fname = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No": hdr = {n: i for i, n in enumerate(r)}; continue
    if not hdr or len(r) < 10 or not r[0].isdigit(): continue
    try:
        ie = float(r[hdr["Instructions Executed"]] or 0); te = float(r[hdr["Thread Instructions Executed"]] or 0)
    except ValueError:
        continue
    smp = float(r[hdr["# Samples"]] or 0); noi = float(r[hdr["stall_no_inst"]] or 0); lsb = float(r[hdr["stall_long_sb"]] or 0)
    out.append((ie, te, smp, noi, lsb, fname, int(r[0]), r[1].strip()[:110]))
    per_file[fname] += ie; thr_file[fname] += te
tot = sum(o[0] for o in out); tots = sum(o[2] for o in out)
print("total warp inst %.3fG, thread inst %.3fG, threads/inst %.2f, samples %d" % (tot / 1e9, sum(o[1] for o in out) / 1e9, sum(o[1] for o in out) / max(tot, 1), tots))
for f, v in per_file.most_common(): print("  %-22s %5.1f%% inst  thr/inst %.1f" % (f, 100 * v / tot, thr_file[f] / max(v, 1)))
if len(sys.argv) > 3:   # list one file in line order: python tools/ncu_lines.py rep N file.cuh [first last]
    lo = int(sys.argv[4]) if len(sys.argv) > 4 else 0; hi = int(sys.argv[5]) if len(sys.argv) > 5 else 10**9
    for ie, te, smp, noi, lsb, f, ln, src in sorted(out, key=lambda o: o[6]):
        if f == sys.argv[3] and lo <= ln <= hi and ie > 0:
            print("%4d %5.2f%% thr %4.1f smp %4.2f%%  %s" % (ln, 100 * ie / tot, te / max(ie, 1), 100 * smp / max(tots, 1), src[:90]))
    sys.exit(0)
out.sort(reverse=True)
for ie, te, smp, noi, lsb, f, ln, src in out[:topn]:
    print("%5.1f%% thr %4.1f smp %4.1f%% (noinst %4.1f%% lsb %4.1f%%) %s:%d  %s" % (100 * ie / tot, te / max(ie, 1), 100 * smp / max(tots, 1), 100 * noi / max(smp, 1), 100 * lsb / max(smp, 1), f, ln, src))
