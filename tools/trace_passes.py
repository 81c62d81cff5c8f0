"""Per-pass times of the flow kernel (first pass and the passes with larger work areas) for one batch of a bench config (aligner option trace=1).
usage (GPU box): python tools/trace_passes.py --config 4 --reads 250000 [opt=value ...]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import genome_weaver_align_b200 as gw  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--config", type=int, default=4)
ap.add_argument("--reads", type=int, default=0)
ap.add_argument("opts", nargs="*")
a = ap.parse_args()
args = argparse.Namespace(config=a.config, genome_bp=0, reads=a.reads, ref_sample=0, cache="/tmp")
spec = bench.spec_of(args)
contigs, genome = bench.make_genome(args, spec)
fa, _, _ = bench.ensure_index(args, spec, contigs, use_gpu=True)
idx = gw.FMIndexOnGenome.load(fa, 0)
reads = [r.tobytes().decode() for g in bench.make_reads(spec, genome, 0) for r in g]
cfg = gw.AlignmentConfig(k=spec["k"], reportType=spec["report"], extended=spec["extended"])
opts = {kv.split("=")[0]: int(kv.split("=")[1]) for kv in a.opts}
opts["trace"] = 1
bsf = gw.BidirectionalSuffixFilter(idx, cfg, **opts)
bsf.align(reads[:20000])
sys.stderr.write("---- timed batch, options %r\n" % opts)
b = bsf.align(reads)
sys.stderr.write("whole %.1f ms, counters %r\n" % (b.timing[5], b.counters))
bsf.close()
idx.close()
