"""Pins the CPU oracle against the golden vectors held by the reference's own unit tests.
Paths cited are relative to /root/reference/align/src/test/java/org/utgenome/weaver/align/."""
import ctypes as C
import random

import pytest

import oracle_lib as O

REF16 = "AAGCCTAGTTTCCTTG"  # strategy/BidirectionalSuffixFilterTest.java:47


@pytest.fixture(scope="module")
def idx16():
    i = O.OracleIndex.from_sequence("seq", REF16)
    yield i
    i.close()


def align1(idx, q, **cfg):
    cfg.setdefault("k", 2)
    r = idx.align(["read"], [q], **cfg)
    return r


def first(idx, q):
    r = align1(idx, q)
    assert r["status"][0] != 2
    return O.parse_sam(r["sam"][0][0]), r


# ---- strategy/BidirectionalSuffixFilterTest.java
def test_one_mismatch_at_tail(idx16):  # :69-75
    a, _ = first(idx16, "GCCTAC")
    assert (a["cigar"], a["start"], a["strand"], a["nm"]) == ("5M1S", 3, "+", 0)


def test_one_mismatch(idx16):  # :78-87
    a, _ = first(idx16, "GCCAAGTT")
    assert (a["cigar"], a["start"], a["strand"], a["nm"]) == ("8M", 3, "+", 1)


def test_one_mismatch_reverse(idx16):  # :90-96
    a, _ = first(idx16, O.revcomp("GCGTAGTT"))
    assert (a["cigar"], a["start"], a["strand"], a["nm"]) == ("8M", 3, "-", 1)


def test_bidirectional_search(idx16):  # :99-105
    a, _ = first(idx16, "AACCCTAGTTTCGTT")
    assert (a["cigar"], a["start"], a["strand"], a["nm"]) == ("15M", 1, "+", 2)


def test_bidirectional_search_reverse(idx16):  # :108-114
    a, _ = first(idx16, O.revcomp("AACCCTAGTTTCGTT"))
    assert (a["cigar"], a["start"], a["strand"], a["nm"]) == ("15M", 1, "-", 2)


def test_two_mismatch_at_head(idx16):  # :117-123
    a, _ = first(idx16, "TTGCCTAGTTT")
    assert (a["cigar"], a["start"], a["strand"], a["nm"]) == ("2S9M", 3, "+", 0)


def test_forward_exact(idx16):  # :126-132
    a, _ = first(idx16, "GCCTAGT")
    assert (a["cigar"], a["start"], a["strand"], a["nm"]) == ("7M", 3, "+", 0)


def test_reverse_exact(idx16):  # :135-141
    a, _ = first(idx16, O.revcomp("GCCTAGT"))
    assert (a["cigar"], a["start"], a["strand"], a["nm"]) == ("7M", 3, "-", 0)


def test_split_exact(idx16):  # :144-158
    r = align1(idx16, "AAGCCTATCCTTG")
    lines = r["sam"][0]
    assert len(lines) == 2
    a, s = O.parse_sam(lines[0]), O.parse_sam(lines[1])
    assert (a["cigar"], a["start"], a["strand"], a["nm"]) == ("7M", 1, "+", 0)
    assert (s["cigar"], s["start"], s["strand"], s["nm"]) == ("6M", 11, "+", 0)


def test_split_exact_reverse(idx16):  # :161-173
    r = align1(idx16, O.revcomp("AAGCCTATCCTTG"))
    lines = r["sam"][0]
    assert len(lines) == 2
    a, s = O.parse_sam(lines[0]), O.parse_sam(lines[1])
    assert (a["cigar"], a["start"], a["strand"], a["nm"]) == ("7M", 1, "-", 0)
    assert (s["cigar"], s["start"], s["strand"], s["nm"]) == ("6M", 11, "-", 0)


def test_split2_runs(idx16):  # :176-178 (no assertion in the reference)
    assert align1(idx16, "AGCCTATTCCTT")["status"][0] != 2


def test_long_read(idx16):  # :181-187
    a, _ = first(idx16, "AAGCCTAGATTCCGTG")
    assert (a["cigar"], a["start"], a["strand"], a["nm"]) == ("16M", 1, "+", 2)


def test_clip(idx16):  # :190-198 (end = start + unclipped size = 11)
    a, _ = first(idx16, "AAGCCTAGTTAAAAAA")
    assert (a["cigar"], a["start"]) == ("10M6S", 1)


def test_one_deletion(idx16):  # :212-220
    a, _ = first(idx16, "AAGCCTGTTT")
    assert (a["cigar"], a["start"], a["strand"], a["nm"]) == ("6M1D4M", 1, "+", 1)


def test_one_insertion(idx16):  # :223-231
    a, _ = first(idx16, "AAGCCTCAGTT")
    assert (a["cigar"], a["start"], a["strand"], a["nm"]) == ("6M1I4M", 1, "+", 1)


# :201-209 clip2 and :234-242 twoInsertion do not reproduce under a literal reading of the reference
# (very likely red on the reference HEAD; cannot be confirmed without a JVM). Recorded as expected-red.
@pytest.mark.xfail(reason="expected-red on reference HEAD (SURVEY.md 4.3)", strict=False)
def test_clip2_expected_red(idx16):
    a, _ = first(idx16, "TTTTTTGAGTTTCCTTG")
    assert (a["cigar"], a["start"]) == ("7S10M", 7)


@pytest.mark.xfail(reason="expected-red on reference HEAD (SURVEY.md 4.3)", strict=False)
def test_two_insertion_expected_red(idx16):
    a, _ = first(idx16, "AAGCCAATAGTTT")
    assert (a["cigar"], a["start"], a["nm"]) == ("5M2I6M", 1, 2)


# ---- BitParallelSmithWatermanTest.java
@pytest.mark.parametrize("ref,query,k,pos,cigar,nm", [
    ("TATACCAAGATCTAGAGATCTGG", "TACCAAGATAGAGATCTGG", 31, 2, "8M2D11M", 2),  # :184-191
    ("GATCTA", "GCTATA", 31, 0, "6M", 2),  # :194-200
    ("TATACCAAGATCTAGAGATCTGG", "TACCAAGATCTCTAGAGATCTGG", 31, 2, "8M2I13M", 2),  # :203-210
    ("TATACCAAGATCTAGAGATCTGG", "ACCAAGATCTAGAG", 31, 3, "14M", 0),  # :213-220
    ("TATACCAAGATCTAGAGATCTGG", "GGCGCACCAAGATCTAGAG", 31, 3, "5S14M", 0),  # :223-230
    ("TATACCAAGATCTAGAGTCTGG", "ACCAAGATCTAGAGAAAA", 31, 3, "14M4S", 0),  # :233-240
    ("GCTTCAGTTTCCTGACACTTAAAAAAAAAAGAGTTGCTTATTATTTTAATGAGACTAATGCTTACACTCTGAGTTACTTGTAAGGTGATTGGTTACTTTAATGTTATTATAAGTAATTT",
     "CTGACACTTAAAAAAAAAAGAGTTGCTTATTATTTTAATGAGACTAATGCTTACACTCTGAGTTACTTGTAAGGTGATTGGTTACTTTAATGTTATT", 11, 11, "97M", 0),  # :289-301
])
def test_bpsw_goldens(ref, query, k, pos, cigar, nm):
    a = O.align_block_detailed(ref, query, k)
    assert a is not None
    assert (a["pos"], a["cigar"], a["nm"]) == (pos, cigar, nm)


def test_bpsw_exact_match_sweep():  # :304-319
    r = ("GCTTCAGTTTCCTGACACTTAAAAAAAAAAGAGTTGCTTATTATTTTAATGAGACTAATGCTTACACTCTGAGTTACTTGTAAGGTGATTGGTTACTTTAATGTTATTATAAGTAATTT"
         "ATTGGTTACTTTAATGTTATTATAAGTAAT")
    for k in range(30, len(r)):
        for i in range(0, len(r) - k):
            a = O.align_block_detailed(r, r[i:i + k], 2)
            assert a is not None and (a["pos"], a["cigar"], a["nm"]) == (i, "%dM" % k, 0), (k, i, a)


# ---- FMIndexOnOccTableTest.java:41-68 (test.fa = TATAATAATATAATA)
def test_fm_backward_search_goldens():
    idx = O.OracleIndex.from_sequence("seq", "TATAATAATATAATA")
    n = idx.text_size()
    si = (0, n - 1)
    expect = [(3, (9, 15)), (0, (3, 9)), (3, (12, 15)), (0, (6, 9)), (0, (1, 3))]
    for ch, e in expect:
        si = idx.backward_search(0, ch, si[0], si[1])
        assert si == e
    # ACGTSequenceTest.java:210-213: the BWT of test.fa is TTTTATTAAAAAAAA, A-count 9
    assert idx.bwt(0) == "TTTTATTAAAAAAAA"
    idx.close()


# ---- sais/CyclicSAISTest.java:69-80
def test_cyclic_sais_int():
    t = (C.c_int * 6)(3, 2, 2, 3, 1, 0)
    sa = (C.c_int * 6)()
    O.lib().gwo_cyclic_sais_int(t, 6, sa)
    assert list(sa) == [5, 4, 1, 2, 3, 0]


# ---- BWAlignTest.java:78-95 (test2.fa, two contigs, default k=0.1 -> exact only)
def test_bwalign_align3(tmp_path):
    fa = tmp_path / "test2.fa"
    fa.write_text(">seq1\nAACCTATCTATACCCCGGGGAATTATA\n>seq2\nACGTCCTTATACTTTATCTCATGGGATA\n")
    O.OracleIndex.build_files(str(fa))
    idx = O.OracleIndex.load(str(fa))
    r = idx.align(["read"], ["TAAAGTAT"])
    a = O.parse_sam(r["sam"][0][0])
    assert (a["chr"], a["strand"], a["start"]) == ("seq2", "-", 9)
    assert idx.sam_header() == "@SQ\tSN:seq1\tLN:27\n@SQ\tSN:seq2\tLN:28\n"
    idx.close()


# ---- ACGTSequenceTest.java:156-208: fastCount / fastCountACGTN == naive count over all (x, y)
def test_fast_count_equals_naive():
    rng = random.Random(0)
    s = "".join(rng.choice("ACGTN") for _ in range(69))
    out = (C.c_longlong * 5)()
    for x in range(0, len(s), 3):
        for y in range(x, len(s) + 1, 2):
            O.lib().gwo_fast_count_acgtn(s.encode(), x, y, out)
            assert list(out) == [s[x:y].count(c) for c in "ACGTN"]
            for ci, c in enumerate("ACGTN"):
                assert O.lib().gwo_fast_count(s.encode(), ci, x, y) == s[x:y].count(c)


# ---- QueryMaskTest.java:36-67: check(query, answer, d, ch, offset, margin, boundary) ==
#      getBidirectionalPatternMask64(d, offset, boundary, offset, ch, margin) read as an m-bit binary string (MSB first)
@pytest.mark.parametrize("answer,d,offset,margin,boundary", [
    ("00001011", 0, 0, 0, 0), ("00000101", 0, 1, 0, 0), ("00000001", 0, 3, 0, 3), ("00011000", 1, 3, 2, 3),
    ("00000110", 1, 3, 0, 3), ("00000011", 1, 2, 0, 3), ("00000001", 1, 1, 0, 3), ("00000000", 1, 0, 0, 3)])
def test_query_mask_vectors(answer, d, offset, margin, boundary):
    q = b"AAGATTGC"
    # nextACGTIndex argument is `offset` in the test (same value passed for cursor)
    v = O.lib().gwo_query_mask64(q, d, offset, boundary, offset, 0, margin)
    m = len(q)
    assert "".join("1" if (v >> i) & 1 else "0" for i in range(m - 1, -1, -1)) == answer


def test_snappy_roundtrip():
    rng = random.Random(1)
    for n in (0, 1, 59, 60, 61, 255, 256, 257, 65535, 65536, 70000):
        data = bytes(rng.randrange(256) for _ in range(n))
        assert O.lib().gwo_compress_roundtrip(data, n) == 0


def test_unmapped_sam_line(idx16):
    r = align1(idx16, "GGGGGGGGGGGG")
    assert r["status"][0] == 1
    assert r["sam"][0] == ["read\t68\t*\t0\t1\t\t*\t0\t0\tGGGGGGGGGGGG\t*"]


def test_sam_text_exact_forward_and_reverse(idx16):
    r = idx16.align(["r1", "r2"], ["GCCTAGT", O.revcomp("GCCTAGT")], quals=["IIIIIIH", "ABCDEFG"], k=2)
    assert r["sam"][0] == ["r1\t66\tseq\t3\t1\t7M\t*\t0\t0\tGCCTAGT\tIIIIIIH\tNM:i:0\tXP:Z:U\tX0:i:1"]
    assert r["sam"][1] == ["r2\t82\tseq\t3\t1\t7M\t*\t0\t0\tGCCTAGT\tGFEDCBA\tNM:i:0\tXP:Z:U\tX0:i:1"]
