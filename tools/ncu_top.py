"""Top source lines of a kernel by warp-stall samples, from an .ncu-rep captured with --import-source on.
usage: python tools/ncu_top.py report.ncu-rep [top_n]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
fname = None; hdr = None; out = []
for r in rows:
    if len(r) >= 2 and r[0] in ("File Name", "File Path"): fname = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No": hdr = {n: i for i, n in enumerate(r)}; continue
    if not hdr or len(r) < 10 or not r[0].isdigit(): continue
    def g(k):
        try: return float(r[hdr[k]] or 0)
        except ValueError: return 0.0
    out.append((g('# Samples'), g('stall_long_sb'), g('stall_no_inst'), g('Instructions Executed'), g('Thread Instructions Executed'), fname, int(r[0]), r[1].strip()[:95]))
ts = sum(o[0] for o in out)
for o in sorted(out, reverse=True)[:topn]:
    print("smp %4.1f%% (lsb %3.0f%% noi %3.0f%%) thr %4.1f %s:%d %s" % (100 * o[0] / ts, 100 * o[1] / max(o[0], 1), 100 * o[2] / max(o[0], 1), o[4] / max(o[3], 1), o[5], o[6], o[7]))
