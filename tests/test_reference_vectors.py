"""Unit-test vectors of the reference re-hosted on the oracle (CPU), plus host-side pieces of the product that need no GPU:
the staircase tables of libgwalign against the oracle's for every (m, kk), and the exported symbols against include/gw_align.h.

Reference tests (align/src/test/java/org/utgenome/weaver/align/): BitVectorTest.java:64-84 (lshift, lshift7 — the members on the
alignment path), CIGARTest.java:35-80, SparseSuffixArrayTest.java:44-92, BWTransformTest.java:73-113."""
import ctypes as C
import os
import re
import subprocess

import numpy as np

import oracle_lib as O
from genome_weaver_align_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _lshift(bits, n, not_first=0):
    out = C.create_string_buffer(len(bits) + 1)
    assert O.lib().gwo_bitvector_lshift(bits.encode(), n, not_first, out, len(bits) + 1) == 0
    return out.value.decode()


def test_bitvector_lshift():
    # BitVectorTest.lshift (:64-77): shifting towards the high end drops the leading bits and appends zeros
    orig = "001011010001000001111110010101010011101111110001010101001111110000000001101010100010101010011100001010100010100111"
    for i in range(len(orig)):
        assert _lshift(orig, i) == orig[i:] + "0" * i, i
    # BitVectorTest.lshift7 (:79-84): new BitVector(28).not().lshift(7)
    assert _lshift("0" * 28, 7, not_first=1) == "1111111111111111111110000000"
    # BitVector._not with a length that is a multiple of 64 zeroes the last block (Java shift masking)
    assert _lshift("0" * 128, 0, not_first=1) == "0" * 64 + "1" * 64


def test_cigar_merge_parse_construct():
    out = C.create_string_buffer(256)
    # CIGARTest.merge (:37-42)
    assert O.lib().gwo_cigar_merge(b"12S", b"2S62M", out, 256) == 2 and out.value == b"14S62M"
    # CIGARTest.parse (:44-65): 4 elements 5M 24I 9M 62S
    assert O.lib().gwo_cigar_merge(b"5M24I9M62S", b"", out, 256) == 4 and out.value == b"5M24I9M62S"
    # CIGARTest.construct (:67-80)
    lens = (C.c_int * 9)(10, 5, 10, 2, 3, 43, 100, 1, 134)
    assert O.lib().gwo_cigar_build(lens, b"SMIMDNMPH", 9, out, 256) == 9 and out.value == b"10S5M10I2M3D43N100M1P134H"


def test_sparse_suffix_array_roundtrip(tmp_path):
    # SparseSuffixArrayTest.shortSeq (:44-62) with L = 32, longSeq (:64-92) with L = 16 on a longer sequence
    short = "AACCTATCTATACCCCGGGGAATTATATACGTCCTTATACTTTATCTCATGGGATA"
    long_ = synth.random_genome(3000, 17).tobytes().decode()
    for seq, L in ((short, 32), (long_, 16), (long_, 32)):
        sa = (C.c_int * len(seq))()
        assert O.lib().gwo_ssa_roundtrip(seq.encode(), L, str(tmp_path / "x.ssa").encode(), sa) == 0
        want = sorted(range(len(seq)), key=lambda i: ((seq + seq)[i:i + len(seq)], i))
        assert list(sa) == want


def test_bwt_files_hold_text_and_reversed_text(tmp_path):
    # BWTransformTest.transform (:73-113): after `bwt`, .f.3bit decodes to the sequence and .r.3bit to its reverse
    g = synth.random_genome(5000, 23, n_frac=0.02)
    fa = str(tmp_path / "sample.fa")
    synth.write_fasta(fa, [("s1", g[:3000]), ("s2", g[3000:])])
    O.OracleIndex.build_files(fa)
    text = g.tobytes().decode()
    for ext, want in ((".f.3bit", text), (".r.3bit", text[::-1])):
        p = O.lib().gwo_load_acgt_string((fa + ext).encode())
        assert p
        assert O._take_str(p) == want, ext


def test_staircase_tables_match_the_oracle():
    # the product's host-side StaircaseFilter tables (csrc/gw_staircase.cpp) against the oracle's, including where the reference's
    # constructor throws, for every read length and every kk the aligner can ask for
    from genome_weaver_align_b200 import aligner
    for m in range(1, 257):
        nb = (m + 63) // 64
        for kk in range(0, 27):
            got, thrown = aligner.staircase_table(m, kk)
            want = (C.c_ulonglong * ((kk + 1) * nb))()
            rc = O.lib().gwo_staircase_table(m, kk, 0, want, (kk + 1) * nb)
            assert thrown == (rc == 2), (m, kk)
            if not thrown:
                assert [int(x) for x in got.reshape(-1)] == list(want), (m, kk)
            else:   # the `extended` domain: out-of-range blocks read as 0 on both sides
                got, thrown = aligner.staircase_table(m, kk, extended=True)
                assert not thrown and O.lib().gwo_staircase_table(m, kk, 1, want, (kk + 1) * nb) == 0
                assert [int(x) for x in got.reshape(-1)] == list(want), (m, kk)


def test_exported_symbols_match_the_header():
    # include/gw_align.h, the ctypes table and the shared object agree on the entry points
    from genome_weaver_align_b200 import _lib
    header = open(os.path.join(ROOT, "include", "gw_align.h")).read()
    header = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    declared = set(re.findall(r"\b(gw_[a-z0-9_]+)\s*\(", header))
    assert declared == set(_lib.SYMBOLS), (declared ^ set(_lib.SYMBOLS))
    nm = subprocess.run(["nm", "-D", "--defined-only", _lib.SO_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r"\bT (gw_[a-z0-9_]+)\b", nm))
    assert declared <= exported, declared - exported
    _lib.lib()   # loads, and every declared symbol resolves with its signature


def test_no_alloc_verifier_checks_its_limit_before_touching_the_device():
    # gw_verify_pairs_noalloc (no counterpart in the reference): a batch whose DP history does not fit the static device array is refused
    # with GW_E_ARG and a message, before any CUDA call (so this holds on a machine without a GPU too)
    from genome_weaver_align_b200 import _lib
    lib = _lib.lib()
    n = (64 << 20) // ((106 + 2) * 2 * 16) + 1
    rc = lib.gw_verify_pairs_noalloc(0, None, 1, 1, 1, 1, n, 106, 100, 3, 1, 1, 40)   # (the pointers only need to be non-null: nothing is read)
    assert rc == _lib.GW_E_ARG
    msg = lib.gw_last_error().decode()
    assert "static scratch" in msg and str(n) in msg
    assert lib.gw_verify_pairs_noalloc(0, None, 1, 1, 1, 1, 10, 5000, 100, 3, 1, 1, 40) == _lib.GW_E_ARG   # window longer than supported
