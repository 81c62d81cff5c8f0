"""Sweep of aligner options on one bench config: the index and the reads are set up once, every option set gets 2 warm-up and 3 timed
batches (CUDA-event times of the whole device-side path and of the flow kernel, host buffers).
usage (GPU box): python tools/sweep_opts.py [--config 2] [--reads N] "flow_batch=8,flow_age=24" "flow_batch=16" ..."""
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
ap.add_argument("--reads", type=int, default=0)
ap.add_argument("sets", nargs="*")
a = ap.parse_args()
args = argparse.Namespace(config=a.config, genome_bp=0, reads=a.reads, ref_sample=0, cache="/tmp")
spec = bench.spec_of(args)
contigs, genome = bench.make_genome(args, spec)
fa, _, _ = bench.ensure_index(args, spec, contigs, use_gpu=True)
idx = gw.FMIndexOnGenome.load(fa, 0)
reads = [r.tobytes().decode() for g in bench.make_reads(spec, genome, 0) for r in g]
cfg = gw.AlignmentConfig(k=spec["k"], reportType=spec["report"], extended=spec["extended"])
for st in [""] + list(a.sets) + [""]:
    opts = {kv.split("=")[0]: int(kv.split("=")[1]) for kv in st.split(",") if kv}
    bsf = gw.BidirectionalSuffixFilter(idx, cfg, **opts)
    t = []
    for i in range(5):
        b = bsf.align(reads)
        if i >= 2:
            t.append((b.timing[5], b.timing[1]))
    bsf.close()
    t = np.array(t)
    print("%-44s whole %8.2f ms   flow %8.2f ms   (min whole %.2f)" % (st or "(defaults)", t[:, 0].mean(), t[:, 1].mean(), t[:, 0].min()), flush=True)
idx.close()
