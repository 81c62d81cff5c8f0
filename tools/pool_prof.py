"""Phase accounting of gw_pool_kernel (diagnostic build -DGW_POOL_PROF, see tools/sweep_var.sh): share of thread-cycles per phase.
usage: python tools/pool_prof.py [--opt name=value ...]   (expects genome_weaver_align_b200/libgwalign.so built with GW_POOL_PROF)"""
import argparse, ctypes as C, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
ap = argparse.ArgumentParser(); ap.add_argument("--opt", action="append", default=[]); ap.add_argument("--reads", type=int, default=1_000_000)
a = ap.parse_args()
args = argparse.Namespace(genome_bp=100_000_000, reads=a.reads, read_len=100, k=3, cache="/tmp")
import genome_weaver_align_b200 as gw
from genome_weaver_align_b200 import _lib as L
genome = bench.make_genome(args)
fa, _, _ = bench.ensure_index(args, genome, use_gpu=True)
idx = gw.FMIndexOnGenome.load(fa, 0)
reads = bench.make_reads(args, genome, 0)
n, m = reads.shape
off = (np.arange(n + 1, dtype=np.uint64) * np.uint64(m))
bsf = gw.BidirectionalSuffixFilter(idx, gw.AlignmentConfig(k=3))
for o in a.opt:
    k, v = o.split("="); bsf.set_option(k, int(v))
lib = L.lib()
hits = np.zeros(2 * n + 16, dtype=L.HIT_DTYPE); cig = np.zeros(16 * n + 64, dtype=np.uint32)
prof = (C.c_ulonglong * 8)()
names = ["S work", "S wait", "L work", "L wait", "D work", "D wait", "refill", "rounds(thread)"]
for it in range(2):
    nh, nc = C.c_uint64(len(hits)), C.c_uint64(len(cig))
    lib.gw_debug_pool_prof(prof, 1)
    L.check(lib.gw_align_batch(bsf._h, L.ptr(reads.reshape(-1)), L.ptr(off), n, L.ptr(hits), C.byref(nh), L.ptr(cig), C.byref(nc), None))
    lib.gw_debug_pool_prof(prof, 0)
    ms = (C.c_float * 6)(); lib.gw_aligner_last_timing(bsf._h, ms)
    tot = sum(prof[i] for i in range(7))
    print("run %d: search %.1f ms; " % (it, ms[1]) + ", ".join("%s %.1f%%" % (names[i], 100.0 * prof[i] / max(tot, 1)) for i in range(7)) + "; rounds/thread %.1f" % (prof[7] / (296 * 256.0)))
