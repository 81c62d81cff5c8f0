"""Runs every kernel of the path once on the bench workload (config 2 index) in a form that `ncu --set full` can capture kernel by kernel:
the quick scan, locate with the .ssa samples only and with the dense table, the DP verifier alone (gw_verify_pairs on fragment-shaped
pairs), the gather micro-benchmark, and the flow kernel for k = 3 (gw_flow_kernel<4>) and k = 6 / 150 bp (gw_flow_kernel<8>).
usage (GPU box):  ncu --set full --clock-control none --import-source on -k regex:'gw_' -o gpurun_out/kernels python tools/profile_kernels.py
       (plain `python tools/profile_kernels.py` first: prints CUDA-event times)"""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import genome_weaver_align_b200 as gw  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--reads", type=int, default=200_000)
ap.add_argument("--genome-bp", type=int, default=0)
a = ap.parse_args()

args = argparse.Namespace(config=2, genome_bp=a.genome_bp, reads=a.reads, ref_sample=0, cache="/tmp")
spec = bench.spec_of(args)
contigs, genome = bench.make_genome(args, spec)
fa, _, _ = bench.ensure_index(args, spec, contigs, use_gpu=True)
idx = gw.FMIndexOnGenome.load(fa, 0)
N = idx.textSize()
from genome_weaver_align_b200 import synth  # noqa: E402

reads100 = [r.tobytes().decode() for r in synth.simulate_reads_fast(genome, a.reads, 100, 3, 12)]
reads150 = [r.tobytes().decode() for r in synth.simulate_reads_fast(genome, a.reads // 4, 150, 6, 13)]


def timed(name, fn):
    t = time.time()
    out = fn()
    print("%-46s %.1f ms (wall, incl. copies)" % (name, 1000 * (time.time() - t)))
    return out


# 1. quick scan (kernel 1), both forms
timed("quick scan, 4 lanes per task", lambda: idx.quickScan(reads100, 0, 1))
timed("quick scan, 1 thread per task", lambda: idx.quickScan(reads100, 0, 0))
# 2. locate (kernel 3): the reference's walk over the .ssa samples, and the dense table
rows = np.random.default_rng(1).integers(0, N, size=1_000_000).astype(np.uint32)
idx.locateTable(32)
timed("locate 1M rows, .ssa samples (LF walks)", lambda: idx.locate(0, rows))
idx.locateTable(1)
timed("locate 1M rows, dense table", lambda: idx.locate(0, rows))
# 3. verifier alone (kernel 4): fragment-shaped pairs (window = read +- k, 0..3 edits)
g = genome
rng = np.random.default_rng(2)
starts = rng.integers(10, len(g) - 200, size=a.reads)
pairs = []
for i, s in enumerate(starts):
    q = reads100[i]
    pairs.append((g[int(s) - 3:int(s) + 103].tobytes().decode(), q))
timed("verify pairs (DP + traceback)", lambda: gw.alignBlockDetailed(pairs, 31))
# 4. flow kernel: search + locate + verify
bsf = gw.BidirectionalSuffixFilter(idx, gw.AlignmentConfig(k=3))
bsf.align(reads100[:20000])
b = timed("align batch k=3 100 bp (flow<4>)", lambda: bsf.align(reads100))
print("   timing ms [scan, flow, -, -, emit, all]:", [round(x, 2) for x in b.timing], b.counters)
bsf.close()
bsf = gw.BidirectionalSuffixFilter(idx, gw.AlignmentConfig(k=6))
bsf.align(reads150[:5000])
b = timed("align batch k=6 150 bp (flow<8>)", lambda: bsf.align(reads150))
print("   timing ms:", [round(x, 2) for x in b.timing], b.counters)
bsf.close()
idx.close()
