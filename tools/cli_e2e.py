"""End-to-end rate of the native command line on the bench workload: FASTQ file in, SAM file out (host parsing, H2D, kernels, D2H, SAM
formatting and writing all inside the measured wall time; index opening included, reported separately from the stats line).
usage: python tools/cli_e2e.py [--gpus N] [--reads R] [--batch B]"""
import argparse, json, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
ap = argparse.ArgumentParser()
ap.add_argument("--gpus", type=int, default=1); ap.add_argument("--reads", type=int, default=2_000_000); ap.add_argument("--batch", type=int, default=1 << 20)
ap.add_argument("--out", default="/dev/shm/gw_cli_out.sam")
a = ap.parse_args()
args = argparse.Namespace(genome_bp=100_000_000, reads=a.reads, read_len=100, k=3, cache="/tmp")
genome = bench.make_genome(args)
fa, _, how = bench.ensure_index(args, genome, use_gpu=True)
reads = bench.make_reads(args, genome, 0)
fq = "/dev/shm/gw_cli_reads.fastq"
t = time.time()
with open(fq, "wb") as f:
    qual = b"I" * args.read_len
    f.write(b"".join(b"@r%d\n%s\n+\n%s\n" % (i, reads[i].tobytes(), qual) for i in range(len(reads))))
print("fastq written: %.1f MB in %.1f s" % (os.path.getsize(fq) / 1e6, time.time() - t), file=sys.stderr)
cli = os.path.join(ROOT, "genome_weaver_align_b200", "genome-weaver")
for mode in ("file", "silent"):
    cmd = [cli, "align", "-r", fa, "-k", "3", fq, "--gpus", str(a.gpus), "--batch", str(a.batch), "--stats"] + (["--silent"] if mode == "silent" else [])
    t = time.time()
    with open(a.out, "wb") as out:
        p = subprocess.run(cmd, stdout=out, stderr=subprocess.PIPE, text=True)
    wall = time.time() - t
    st = json.loads(p.stderr.strip().splitlines()[-1]) if p.returncode == 0 else {"error": p.stderr[-300:]}
    print(json.dumps({"tool": "genome-weaver align", "mode": mode, "gpus": a.gpus, "reads": a.reads, "batch": a.batch, "wall_s": round(wall, 3),
                      "reads_per_s_wall": round(a.reads / wall), "sam_bytes": os.path.getsize(a.out), "stats": st}))
os.remove(fq); os.remove(a.out)
