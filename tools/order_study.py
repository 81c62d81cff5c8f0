"""How well does the quick scan predict the search work of a read? Mean / quantiles of numFMIndexSearches and numSW per read by
(mismatching stretches on the cheaper strand, on the other strand, width of the last interval), on a sample of a bench config.
usage (GPU box): python tools/order_study.py [--config 2] [--reads 200000]"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import genome_weaver_align_b200 as gw  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--config", type=int, default=2)
ap.add_argument("--reads", type=int, default=200_000)
a = ap.parse_args()
args = argparse.Namespace(config=a.config, genome_bp=0, reads=a.reads, ref_sample=0, cache="/tmp")
spec = bench.spec_of(args)
contigs, genome = bench.make_genome(args, spec)
fa, _, _ = bench.ensure_index(args, spec, contigs, use_gpu=True)
idx = gw.FMIndexOnGenome.load(fa, 0)
groups = bench.make_reads(spec, genome, 0)
reads = [r.tobytes().decode() for g in groups for r in g]
cfg = gw.AlignmentConfig(k=spec["k"], reportType=spec["report"], extended=spec["extended"])
bsf = gw.BidirectionalSuffixFilter(idx, cfg)
b = bsf.align(reads)
rc = np.asarray(b.read_counters, dtype=np.int64)
s0, s1 = idx.quickScan(reads, 0), idx.quickScan(reads, 1)
k = spec["k"]
mm0, mm1 = s0[:, 0], s1[:, 0]
lo, hi = np.minimum(mm0, mm1), np.maximum(mm0, mm1)
fm, sw = rc[:, 0], rc[:, 1]
print("reads %d, total fm %d, sw %d" % (len(reads), fm.sum(), sw.sum()))
print("%-28s %8s %9s %9s %9s %9s %8s %7s" % ("class", "reads", "fm mean", "fm p50", "fm p90", "fm p99", "sw mean", "fm share"))
for l in range(0, k + 2):
    for h in ("<=k", ">k"):
        sel = (np.minimum(lo, k + 1) == l) & ((hi <= k) if h == "<=k" else (hi > k))
        if sel.sum() == 0:
            continue
        f = fm[sel]
        print("%-28s %8d %9.1f %9.0f %9.0f %9.0f %8.2f %6.1f%%" % ("min mm %d, other strand %s" % (l, h), sel.sum(), f.mean(), np.median(f),
                                                              np.percentile(f, 90), np.percentile(f, 99), sw[sel].mean(), 100.0 * f.sum() / max(fm.sum(), 1)))
# within the heaviest classes: does the length of the longest exact stretch (scan_le - scan_ls) of the cheaper strand matter?
best = np.where(mm0 <= mm1, 0, 1)
ll = np.where(best == 0, s0[:, 2] - s0[:, 1], s1[:, 2] - s1[:, 1])
for l in range(1, k + 1):
    sel = (lo == l) & (hi > k)
    if sel.sum() < 100:
        continue
    q = np.percentile(ll[sel], [25, 50, 75])
    for name, s2 in (("longest < p25", ll < q[0]), ("p25..p50", (ll >= q[0]) & (ll < q[1])), ("p50..p75", (ll >= q[1]) & (ll < q[2])), (">= p75", ll >= q[2])):
        f = fm[sel & s2]
        if len(f):
            print("min mm %d, longest stretch %-14s (%3d..) reads %7d fm mean %8.1f p90 %7.0f" % (l, name, int(ll[sel & s2].min()), len(f), f.mean(), np.percentile(f, 90)))
bsf.close()
idx.close()
