"""Scale check for the cfg 3 / cfg 4 shape: a multi-contig genome with injected repeats and N runs, indexed by
`genome-weaver bwt` on the device, aligned on the GPU through the C-ABI, and a SAMPLE of the reads compared line by line with the
CPU oracle reading the SAME index files (the oracle's own suffix sorter is too slow at this size).

usage: python tools/scale_check.py --genome-bp 1000000000 --reads 1000000 --read-len 150 --k 3 --sample 2000 [--out result.json]
Prints one JSON line: sizes, index build / open time, reads/s of the GPU path (device timing of gw_align_batch), counters,
roofline inputs (occ lookups), and the parity verdict of the sample. Test infrastructure: not part of the alignment path."""
import argparse, json, os, sys, time
import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import genome_weaver_align_b200 as gw
from genome_weaver_align_b200 import synth
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import oracle_lib


def stage(msg, _t=[time.time()]):
    sys.stderr.write("[scale_check %7.1fs] %s\n" % (time.time() - _t[0], msg))
    sys.stderr.flush()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--genome-bp", type=int, default=100_000_000)
    ap.add_argument("--contigs", type=int, default=24)
    ap.add_argument("--reads", type=int, default=1_000_000)
    ap.add_argument("--read-len", type=int, default=150)
    ap.add_argument("--k", type=float, default=3)
    ap.add_argument("--edits", type=int, default=3)
    ap.add_argument("--sample", type=int, default=2000)
    ap.add_argument("--threads", type=int, default=os.cpu_count() or 1)
    ap.add_argument("--dir", default="/tmp/gw_scale")
    ap.add_argument("--out", default=None)
    ap.add_argument("--opt", action="append", default=[])
    a = ap.parse_args()
    os.makedirs(a.dir, exist_ok=True)
    res = dict(tool="scale_check", genome_bp=a.genome_bp, contigs=a.contigs, reads=a.reads, read_len=a.read_len, k=a.k)

    t0 = time.time()
    g = synth.random_genome(a.genome_bp, 3, n_frac=0.005)
    g = synth.inject_repeats(g, 3, families=20, copies=(100, max(200, min(10000, a.genome_bp // 300000))))
    g[-1000:] = synth.random_genome(1000, 33)   # last 1 kb unique
    # contig sizes proportional to 24, 23, ... (hg19-like spread)
    w = np.arange(a.contigs, 0, -1, dtype=np.float64)
    cuts = np.concatenate([[0], np.cumsum((w / w.sum() * a.genome_bp).astype(np.int64))])
    cuts[-1] = a.genome_bp
    contigs = [("chr%d" % (i + 1), g[cuts[i]:cuts[i + 1]]) for i in range(a.contigs)]
    fasta = os.path.join(a.dir, "ref_%d.fa" % a.genome_bp)
    synth.write_fasta(fasta, contigs)
    res["synth_s"] = round(time.time() - t0, 1)
    stage("genome written")

    t0 = time.time()
    gw.bwt(fasta)
    res["bwt_s"] = round(time.time() - t0, 1)
    stage("index built")
    t0 = time.time()
    idx = gw.FMIndexOnGenome.load(fasta)
    res["open_s"] = round(time.time() - t0, 1)
    res["text_size"] = int(idx.textSize())
    stage("index opened")

    t0 = time.time()
    reads = synth.simulate_reads_fast(g, a.reads, a.read_len, a.edits, 13)
    seqs = [bytes(r) for r in reads]
    res["reads_s"] = round(time.time() - t0, 1)
    stage("reads simulated")
    del g

    cfg = gw.AlignmentConfig(k=a.k)
    al = gw.BidirectionalSuffixFilter(idx, cfg)
    for o in a.opt:
        n, v = o.split("=")
        al.set_option(n, int(v))
    al.align(seqs[: min(len(seqs), 20000)])   # warm-up (allocations)
    stage("warm-up batch aligned")
    t0 = time.time()
    batch = al.align(seqs)
    wall = time.time() - t0
    stage("batch aligned")
    ms = batch.timing
    res["gpu"] = dict(wall_s=round(wall, 3), device_ms=round(ms[5], 2), reads_per_s=round(a.reads / (ms[5] / 1e3)), family_ms=[round(x, 2) for x in ms[:5]],
                      counters=batch.counters)
    st = batch.status()
    res["status_counts"] = {int(k): int(v) for k, v in zip(*np.unique(np.asarray(st), return_counts=True))}

    # parity of a sample against the oracle on the same index files (--sample 0: GPU side only; the oracle needs about 0.5 s per
    # Mbp to open an index)
    ns = min(a.sample, a.reads)
    if ns == 0:
        res["parity"] = "not checked"
        line = json.dumps(res)
        print(line)
        if a.out:
            with open(a.out, "w") as f:
                f.write(line + "\n")
        return 0
    pick = np.linspace(0, a.reads - 1, ns).astype(np.int64)
    names = ["r%d" % i for i in range(a.reads)]
    got = batch.sam_per_read(names)
    stage("SAM formatted")
    t0 = time.time()
    oidx = oracle_lib.OracleIndex.load(fasta)
    res["oracle_open_s"] = round(time.time() - t0, 1)
    stage("oracle index opened")
    t0 = time.time()
    ref = oidx.align([names[i] for i in pick], [seqs[i].decode() for i in pick], nthreads=a.threads, k=a.k)
    res["oracle_sample_s"] = round(time.time() - t0, 2)
    stage("oracle sample aligned")
    res["oracle_reads_per_s"] = round(ns / max(time.time() - t0, 1e-9))
    bad = []
    for j, i in enumerate(pick):
        if list(got[i]) != list(ref["sam"][j]):
            bad.append(int(i))
    res["sample"] = ns
    res["sample_mismatches"] = len(bad)
    res["parity"] = "ok" if not bad else "MISMATCH"
    if bad:
        i = bad[0]
        j = int(np.nonzero(pick == i)[0][0])
        res["first_bad"] = dict(read=i, gpu=list(got[i])[:2], oracle=list(ref["sam"][j])[:2])
    line = json.dumps(res)
    print(line)
    if a.out:
        with open(a.out, "w") as f:
            f.write(line + "\n")
    return 0 if not bad else 1


if __name__ == "__main__":
    sys.exit(main())
