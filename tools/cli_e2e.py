"""End-to-end rate of the native command line (the product's multi-GPU path): ONE FASTQ file in, ONE SAM file out in input order, the reads
sharded over --gpus N index replicas inside `genome-weaver align` (host parsing, H2D, kernels, D2H, SAM formatting and writing all
inside the measured wall time; the index opening is reported separately by --stats). The SAM files of all runs must be identical.
usage: python tools/cli_e2e.py [--config 2] [--gpus 1,2] [--reads 4000000] [--batch 1048576] [--out result.json]"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--config", type=int, default=2)
ap.add_argument("--gpus", default="1")
ap.add_argument("--reads", type=int, default=4_000_000)
ap.add_argument("--batch", type=int, default=1 << 20)
ap.add_argument("--dir", default="/dev/shm")
ap.add_argument("--out", default=None)
a = ap.parse_args()

args = argparse.Namespace(config=a.config, genome_bp=0, reads=a.reads, ref_sample=0, cache="/tmp")
spec = bench.spec_of(args)
contigs, genome = bench.make_genome(args, spec)
fa, _, how = bench.ensure_index(args, spec, contigs, use_gpu=True)
groups = bench.make_reads(spec, genome, 0)
del genome, contigs
fq = os.path.join(a.dir, "gw_cli_reads.fastq")
t = time.time()
with open(fq, "wb") as f:
    i = 0
    for g in groups:
        qual = b"I" * g.shape[1]
        for lo in range(0, len(g), 100_000):
            f.write(b"".join(b"@r%d\n%s\n+\n%s\n" % (i + j, r.tobytes(), qual) for j, r in enumerate(g[lo:lo + 100_000])))
            i += len(g[lo:lo + 100_000])
print("fastq written: %.1f MB in %.1f s" % (os.path.getsize(fq) / 1e6, time.time() - t), file=sys.stderr)
cli = os.path.join(ROOT, "genome_weaver_align_b200", "genome-weaver")
sam = os.path.join(a.dir, "gw_cli_out.sam")
results, digests = [], {}
for gpus in [int(x) for x in a.gpus.split(",")]:
    for mode in ("file", "silent"):
        cmd = [cli, "align", "-r", fa, "-k", str(spec["k"]), fq, "--gpus", str(gpus), "--batch", str(a.batch), "--stats", "-R", spec["report"]]
        cmd += (["--extended"] if spec["extended"] else []) + (["--silent"] if mode == "silent" else [])
        t = time.time()
        with open(sam, "wb") as out:
            p = subprocess.run(cmd, stdout=out, stderr=subprocess.PIPE, text=True)
        wall = time.time() - t
        st = json.loads(p.stderr.strip().splitlines()[-1]) if p.returncode == 0 else {"error": p.stderr[-300:]}
        line = {"tool": "genome-weaver align", "workload": spec["name"], "mode": mode, "gpus": gpus, "reads": a.reads, "batch": a.batch,
                "wall_s": round(wall, 3), "reads_per_s_wall": round(a.reads / wall), "sam_bytes": os.path.getsize(sam), "stats": st}
        if mode == "file":
            h = hashlib.md5()
            with open(sam, "rb") as fsam:
                for blk in iter(lambda: fsam.read(1 << 24), b""):
                    h.update(blk)
            line["sam_md5"] = digests[gpus] = h.hexdigest()
        results.append(line)
        print(json.dumps(line))
        sys.stdout.flush()
ok = len(set(digests.values())) <= 1
print(json.dumps({"identical_sam_across_gpu_counts": ok, "digests": digests}))
if a.out:
    with open(a.out, "w") as f:
        json.dump({"runs": results, "identical_sam_across_gpu_counts": ok}, f, indent=1)
os.remove(fq)
os.remove(sam)
sys.exit(0 if ok else 1)
