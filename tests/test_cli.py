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
    # compressed read files (ReadReaderFactory.java:128-140): gzip, and the snappy-java stream of ".snap"
    import gzip
    raw = open(fq, "rb").read()
    with gzip.open(fq + ".gz", "wb") as f:
        f.write(raw)
    with open(fq + ".snap", "wb") as f:
        f.write(_snappy_java_stream(raw))
    for path in (fq + ".gz", fq + ".snap"):
        p = subprocess.run([CLI, "align", "-r", fa, "-k", "3", path, "--batch", "900"], capture_output=True, text=True)
        assert p.returncode == 0, p.stderr
        assert p.stdout == expect, path
    oi.close()


def _snappy_java_stream(data, chunk=32 * 1024):
    """SnappyOutputStream framing (magic, version 1, compatible version 1, then [BE int32 length][raw Snappy block] per 32 KiB chunk) with
    literal-only Snappy blocks: valid input for any Snappy decoder"""
    out = bytearray(b"\x82SNAPPY\x00" + (1).to_bytes(4, "big") + (1).to_bytes(4, "big"))
    for p in range(0, len(data), chunk):
        piece = data[p:p + chunk]
        blk = bytearray()
        n = len(piece)
        while True:   # uncompressed length as a varint
            blk.append((n & 0x7F) | (0x80 if n > 0x7F else 0))
            n >>= 7
            if not n:
                break
        for q in range(0, len(piece), 60):   # literals of at most 60 bytes: one-byte tags
            lit = piece[q:q + 60]
            blk.append((len(lit) - 1) << 2)
            blk += lit
        out += len(blk).to_bytes(4, "big") + blk
    return bytes(out)


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


@pytest.mark.gpu
def test_cli_reads_outside_the_supported_domain_and_broken_index_files(tmp_path):
    fa, names, reads, quals = _workload(tmp_path, 50, 3, 51)
    subprocess.check_call([CLI, "bwt", fa])
    reads[10] = "ACGT" * 75                          # 300 bases: longer than the 256 the device path takes
    quals[10] = "I" * 300
    fq = str(tmp_path / "reads.fastq")
    with open(fq, "w") as f:
        for n, r, q in zip(names, reads, quals):
            f.write("@%s\n%s\n+\n%s\n" % (n, r, q))
    p = subprocess.run([CLI, "align", "-r", fa, "-k", "3", fq], capture_output=True, text=True)
    body = [l for l in p.stdout.split("\n") if l and not l.startswith("@")]
    assert p.returncode == 2 and "outside the supported domain" in p.stderr       # reported, not silently dropped
    assert len([l for l in body if l.startswith(names[10] + "\t")]) == 1           # the file keeps one record for the read ...
    assert [l for l in body if l.startswith(names[10] + "\t")][0].split("\t")[1:4] == ["68", "*", "0"]   # ... an unmapped one
    p = subprocess.run([CLI, "align", "-r", fa, "-k", "3", fq, "--keep-going"], capture_output=True, text=True)
    assert p.returncode == 0
    # index files that disagree with their headers are refused when the index is opened (no out-of-bounds reads on the device)
    import genome_weaver_align_b200 as gw
    from genome_weaver_align_b200 import _lib as L
    ssa = fa + ".f.ssa"
    good = open(ssa, "rb").read()
    try:
        for bad in (good[:-8], good[:12] + (2 ** 61).to_bytes(8, "big") + good[20:]):   # two samples short; a count whose size wraps
            open(ssa, "wb").write(bad)
            with pytest.raises(L.GwError) as e:
                gw.FMIndexOnGenome.load(fa, 0)
            assert e.value.code == L.GW_E_IO
    finally:
        open(ssa, "wb").write(good)
    silk = fa + ".i.silk"
    text = open(silk).read()
    try:
        open(silk, "w").write(text.replace("-length:90000", "-length:95000"))
        with pytest.raises(L.GwError) as e:
            gw.FMIndexOnGenome.load(fa, 0)
        assert e.value.code == L.GW_E_IO
    finally:
        open(silk, "w").write(text)
    gw.FMIndexOnGenome.load(fa, 0).close()
