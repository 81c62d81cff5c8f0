# experiment: two batches in flight on one GPU (two aligners, two host threads)
import sys, os, time, threading, ctypes as C
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np, torch, argparse
import bench
import genome_weaver_align_b200 as gw
from genome_weaver_align_b200 import _lib as L
args = argparse.Namespace(config=2, genome_bp=0, reads=0, ref_sample=0, cache="/tmp")
spec = bench.spec_of(args)
contigs, genome = bench.make_genome(args, spec)
fa, _, _ = bench.ensure_index(args, spec, contigs, True)
idx = gw.FMIndexOnGenome.load(fa, 0)
lib = L.lib()
dev = torch.device("cuda", 0)
T = int(sys.argv[1]) if len(sys.argv) > 1 else 2
steps = 4
work = []
for t in range(T):
    groups = bench.make_reads(spec, genome, t)
    bases, off, m = bench.flatten(groups)
    n = len(off) - 1
    w = dict(n=n, m=m, tb=int(off[-1]), d_bases=torch.from_numpy(bases).to(dev), d_off=torch.from_numpy(off.view(np.int64)).to(dev),
             d_hits=torch.empty((2 * n + 16) * 44, dtype=torch.uint8, device=dev), d_cig=torch.empty(16 * n + 64, dtype=torch.int32, device=dev),
             bsf=gw.BidirectionalSuffixFilter(idx, gw.AlignmentConfig(k=3)))
    if len(sys.argv) > 2:
        w["bsf"].set_option("ext_pct", int(sys.argv[2]))
    work.append(w)
def run(w, k):
    nh, nc = C.c_uint64(0), C.c_uint64(0)
    for _ in range(k):
        L.check(lib.gw_align_batch_device(w["bsf"]._h, w["d_bases"].data_ptr(), w["d_off"].data_ptr(), w["n"], w["tb"], w["m"], w["d_hits"].data_ptr(), 2 * w["n"] + 16,
                                          w["d_cig"].data_ptr(), 16 * w["n"] + 64, C.byref(nh), C.byref(nc), None))
for w in work: run(w, 2)
torch.cuda.synchronize()
t0 = time.perf_counter()
th = [threading.Thread(target=run, args=(w, steps)) for w in work]
for x in th: x.start()
for x in th: x.join()
torch.cuda.synchronize()
dt = time.perf_counter() - t0
print("workers %d: %d batches of %d reads in %.1f ms -> %.2f M reads/s (%.1f ms per batch)" % (T, T * steps, work[0]["n"], dt * 1e3, T * steps * work[0]["n"] / dt / 1e6, dt * 1e3 / (T * steps)))
