"""The native command line (`genome-weaver bwt`, `genome-weaver align -r ref.fa reads.fastq > out.sam`, csrc/gw_cli.cpp) against the CPU
oracle: whole SAM files, header included (reference: align/Align.java:76-104,188-195, align/SAMOutput.java:51-82)."""
import os
import subprocess

import pytest

import oracle_lib as O
from genome_weaver_align_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "genome_weaver_align_b200", "genome-weaver")


def test_cli_is_built_and_prints_usage():
    assert os.path.exists(CLI), "run __graft_entry__.build()"
    p = subprocess.run([CLI], capture_output=True, text=True)
    assert p.returncode == 1 and "genome-weaver align -r" in p.stderr
    p = subprocess.run([CLI, "align", "-r", "x.fa", "reads.bam"], capture_output=True, text=True)
    assert p.returncode == 1 and "Unsupported file type" in p.stderr      # ReadReaderFactory.java:152
    p = subprocess.run([CLI, "align", "-r", "x.fa"], capture_output=True, text=True)
    assert p.returncode == 1 and "no query is given" in p.stderr          # Align.java:79-80


def _workload(tmp_path, n_reads, k, seed):
    g = synth.random_genome(150_000, seed)
    fa = str(tmp_path / "g.fa")
    synth.write_fasta(fa, [("chr1 first contig", g[:90_000]), ("chr2", g[90_000:])])
    reads = synth.simulate_reads(g, n_reads, 100, k, seed + 1)
    reads[3] = "ACGT" * 25              # unmapped with high probability
    reads[5] = reads[5][:40] + "N" + reads[5][41:]
    names = ["read %d/1" % i for i in range(n_reads)]   # the whole header line is the name
    quals = ["".join(chr(33 + (i + j) % 40) for j in range(len(r))) for i, r in enumerate(reads)]
    return fa, names, reads, quals


@pytest.mark.gpu
def test_cli_fastq_whole_sam(tmp_path):
    fa, names, reads, quals = _workload(tmp_path, 3000, 3, 31)
    fq = str(tmp_path / "reads.fastq")
    with open(fq, "w") as f:
        for n, r, q in zip(names, reads, quals):
            f.write("@%s\n%s\n+\n%s\n" % (n, r, q))
    subprocess.check_call([CLI, "bwt", fa])
    oi = O.OracleIndex.load(fa)
    want = oi.align(names, reads, quals, k=3)
    expect = oi.sam_header() + "".join(l + "\n" for per in want["sam"] for l in per)
    for extra in (["--batch", "700"], ["--batch", "1000000"]):   # several batches / one batch: same text, input order
        p = subprocess.run([CLI, "align", "-r", fa, "-k", "3", fq, "--stats"] + extra, capture_output=True, text=True)
        assert p.returncode == 0, p.stderr
        assert p.stdout == expect
        assert '"reads": 3000' in p.stderr
    p = subprocess.run([CLI, "align", "-r", fa, "-k", "3", fq, "--silent"], capture_output=True, text=True)
    assert p.returncode == 0 and p.stdout == ""      # AlignmentConfig.java:49-50
    oi.close()


@pytest.mark.gpu
def test_cli_fasta_reads_and_single_query(tmp_path):
    fa, names, reads, _ = _workload(tmp_path, 400, 2, 41)
    rf = str(tmp_path / "reads.fa")
    with open(rf, "w") as f:
        for i, r in enumerate(reads):
            f.write(">q%d some description\n%s\n%s\n" % (i, r[:60], r[60:]))
    subprocess.check_call([CLI, "bwt", fa])
    oi = O.OracleIndex.load(fa)
    want = oi.align(["q%d" % i for i in range(len(reads))], reads, None, k=2, report_type=1)
    expect = oi.sam_header() + "".join(l + "\n" for per in want["sam"] for l in per)
    p = subprocess.run([CLI, "align", "-r", fa, "-k", "2", "-R", "allhits", rf], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    assert p.stdout == expect
    want = oi.align(["read"], [reads[7]], None)      # -q: ReadReaderFactory.singleQueryReader, default -k 0.1
    p = subprocess.run([CLI, "align", "-r", fa, "-q", reads[7]], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    assert p.stdout == oi.sam_header() + "".join(l + "\n" for l in want["sam"][0])
    oi.close()
