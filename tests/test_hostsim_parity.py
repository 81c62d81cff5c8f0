"""CPU-side check of the device automaton logic: the per-thread kernel bodies of the product, compiled for the host by the
test-only simulator (tests/hostsim), against the oracle. The real parity tests (CUDA path through the C-ABI) are in
test_gpu_parity.py; this file exists so that logic errors are caught on a machine without a GPU."""
import os

import pytest

import hostsim_lib as H
import oracle_lib as O
from genome_weaver_align_b200 import synth


def _mk(tmp, name, contigs):
    fa = os.path.join(str(tmp), name)
    synth.write_fasta(fa, contigs)
    O.OracleIndex.build_files(fa)
    return fa, O.OracleIndex.load(fa), H.SimIndex(fa)


def _compare(oi, si, reads, slab_words=4096, **cfg):
    names = ["r%d" % i for i in range(len(reads))]
    quals = ["I" * len(r) for r in reads]
    ro = oi.align(names, reads, quals, **cfg)
    rs = si.align(names, reads, quals, slab_words=slab_words, **cfg)
    bad = [i for i in range(len(reads)) if ro["sam"][i] != rs["sam"][i] or ro["status"][i] != rs["status"][i] or ro["counters"][i] != rs["counters"][i]]
    assert not bad, (bad[:5], ro["sam"][bad[0]], rs["sam"][bad[0]], ro["counters"][bad[0]], rs["counters"][bad[0]])
    return ro, rs


@pytest.fixture(scope="module")
def rnd(tmp_path_factory):
    g = synth.random_genome(50_000, 2, n_frac=0.01)
    fa, oi, si = _mk(tmp_path_factory.mktemp("rnd"), "g.fa", [("chr1", g[:30_000]), ("chr2", g[30_000:])])
    yield g, oi, si
    oi.close()
    si.close()


def test_rank_locate(rnd):
    g, oi, si = rnd
    for w in (0, 1):
        for p in [0, 1, 127, 128, 129, 5000, 49_999, 50_000, 50_007]:
            assert oi.rank_acgtn(w, p) == si.rank_acgtn(w, p)
        for r in [0, 1, 31, 32, 33, 777, 49_999]:
            assert oi.locate(w, r) == si.locate(w, r)


@pytest.mark.parametrize("k,length,n", [(2, 100, 400), (3, 100, 400), (3, 36, 150), (3, 155, 120), (6, 150, 80), (0.1, 100, 100), (0, 50, 100)])
def test_sam_parity(rnd, k, length, n):
    g, oi, si = rnd
    _compare(oi, si, synth.simulate_reads(g, n, length, int(k) if k >= 1 else 6, length + int(10 * k)), k=k)


def test_abort_domain_and_small_slabs(rnd):
    g, oi, si = rnd
    ro, rs = _compare(oi, si, synth.simulate_reads(g, 60, 250, 6, 9), k=6)
    assert 2 in ro["status"]
    ro, rs = _compare(oi, si, synth.simulate_reads(g, 200, 100, 3, 12), slab_words=1024, k=3)
    assert rs["overflow_reads"] > 0   # the larger work-area tiers were exercised


@pytest.mark.parametrize("env", [dict(GWSIM_NOCACHE="1"), dict(GWSIM_SPILL="1"), dict(GWSIM_SPILL="5", GWSIM_NOCACHE="1")])
def test_lane_register_copy_is_transparent(rnd, monkeypatch, env):
    # a search lane keeps the read's counters, queue root and current state in registers; re-reading the state from the work area
    # before every step, or writing the lane back and re-reading it between slices, must not change any result
    g, oi, si = rnd
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    _compare(oi, si, synth.simulate_reads(g, 300, 100, 3, 13), k=3)
    _compare(oi, si, synth.simulate_reads(g, 100, 150, 6, 14), k=6)


def test_extended_domain(rnd):
    # config 4 (250 bp, k = 6): in the `extended` domain the reads on which the reference's StaircaseFilter throws are aligned
    g, oi, si = rnd
    ro, rs = _compare(oi, si, synth.simulate_reads(g, 80, 250, 6, 9), k=6, extended=1)
    assert 2 not in ro["status"] and ro["status"].count(0) > 60
    _compare(oi, si, synth.simulate_reads(g, 60, 200, 3, 10), k=3, extended=1, report_type=1)


def test_edge_reads(rnd):
    g, oi, si = rnd
    s = g.tobytes().decode()
    reads = [s[1000:1100], s[5:41], "N" * 50, s[2000:2150], "ACGT", "A", s[0:100], s[-100:], s[29_950:30_050], s[3000:3050] + "NN" + s[3052:3100],
             O.revcomp(s[7000:7250]), s[-60:] + s[:40], "ACGTNNNNACGT"]
    _compare(oi, si, reads, k=3)
    _compare(oi, si, reads, k=0.1)


def test_repeats_and_low_complexity(tmp_path):
    g2 = synth.inject_repeats(synth.random_genome(60_000, 4), 5, families=5, min_len=200, max_len=800, copies=(5, 20))
    fa, oi, si = _mk(tmp_path, "rep.fa", [("c", g2)])
    _compare(oi, si, synth.simulate_reads(g2, 300, 100, 3, 6), k=3, report_type=1)
    _compare(oi, si, synth.simulate_reads(g2, 150, 100, 3, 7), k=3, report_type=2, top_l=3)
    _compare(oi, si, synth.simulate_reads(g2, 150, 100, 3, 8), k=3, num_split_allowed=0)
    g3 = synth.low_complexity_genome(40_000, 5)
    fa, oi3, si3 = _mk(tmp_path, "lc.fa", [("lc", g3)])
    for L in (36, 75, 125):
        _compare(oi3, si3, synth.simulate_reads(g3, 120, L, 3, L), k=3, report_type=1)


def test_verifier_body_against_oracle():
    import random
    rnd_ = random.Random(5)
    for _ in range(300):
        m = rnd_.randint(1, 250)
        q = "".join(rnd_.choice("ACGT") for _ in range(m))
        ref = list("".join(rnd_.choice("ACGT") for _ in range(rnd_.randint(0, 8))) + q + "".join(rnd_.choice("ACGT") for _ in range(rnd_.randint(0, 8))))
        for _e in range(rnd_.randint(0, 6)):
            p = rnd_.randrange(len(ref))
            r = rnd_.random()
            if r < 0.5:
                ref[p] = rnd_.choice("ACGTN")
            elif r < 0.75:
                ref.insert(p, rnd_.choice("ACGT"))
            else:
                del ref[p]
        ref = "".join(ref)
        a, b = O.align_block_detailed(ref, q, 31), H.verify_pair(ref, q, 31)
        assert (a is None) == (b is None)
        if a:
            assert (a["cigar"], a["pos"], a["nm"]) == (b["cigar"], b["pos"], b["nm"]), (ref, q)


def test_verifier_many_differences_and_unrelated_windows():
    # alignments with dozens of differences, unrelated windows, windows shorter than the query: the device verifier body must still give
    # what alignBlockDetailed gives
    import random
    rnd_ = random.Random(11)
    for t in range(400):
        m = rnd_.choice([5, 30, 63, 64, 65, 100, 127, 128, 129, 150, 200, 250, 256])
        q = "".join(rnd_.choice("ACGT") for _ in range(m))
        kind = t % 4
        if kind == 0:   # unrelated window
            ref = "".join(rnd_.choice("ACGT") for _ in range(rnd_.randint(1, m + 20)))
        else:
            ref = list("".join(rnd_.choice("ACGT") for _ in range(rnd_.randint(0, 12))) + q + "".join(rnd_.choice("ACGT") for _ in range(rnd_.randint(0, 12))))
            for _e in range(rnd_.randint(0, 45 if kind == 1 else 12)):
                p_ = rnd_.randrange(len(ref))
                r = rnd_.random()
                if r < 0.4:
                    ref[p_] = rnd_.choice("ACGTN")
                elif r < 0.7:
                    ref.insert(p_, rnd_.choice("ACGT"))
                elif len(ref) > 1:
                    del ref[p_]
            ref = "".join(ref)
        a, b = O.align_block_detailed(ref, q, 31), H.verify_pair(ref, q, 31)
        assert (a is None) == (b is None), (ref, q)
        if a:
            assert (a["cigar"], a["pos"], a["nm"]) == (b["cigar"], b["pos"], b["nm"]), (ref, q)


@pytest.mark.parametrize("blocks", ["0,0", "1000,1000", "40,3", "0,50"])
def test_moving_into_larger_work_areas_while_running(tmp_path, monkeypatch, blocks):
    # gw_promote_*: a read whose work area is nearly full moves into a block 8 x as large and from there into one 64 x as large; when no
    # block is left it runs until its work area is full and is run again from the start with a larger one. Small first work areas,
    # repeat-rich and low-complexity genomes, ALLHITS: every read goes through some of this, and no result may depend on which part.
    monkeypatch.setenv("GWSIM_PROMOTE", blocks)
    g = synth.inject_repeats(synth.random_genome(40_000, 21), 22, families=4, min_len=200, max_len=600, copies=(5, 20))
    lc = synth.low_complexity_genome(40_000, 23)
    fa, oi, si = _mk(tmp_path, "m.fa", [("rep", g), ("low", lc)])
    import numpy as np
    both = np.concatenate([g, lc])
    for cfg, reads in ((dict(k=3, report_type=1), synth.simulate_reads(both, 250, 100, 3, 24)),
                       (dict(k=3), synth.simulate_reads(both, 150, 150, 3, 25)),
                       (dict(k=5, report_type=1, extended=1), synth.simulate_reads(both, 60, 200, 5, 26))):
        ro, rs = _compare(oi, si, reads, slab_words=1024, **cfg)
        if blocks == "0,0":
            assert rs["overflow_reads"] > 0   # the workload does leave the first work area
    oi.close()
    si.close()


def test_hit_list_grows_in_the_arena(tmp_path, monkeypatch):
    # gw_add_result: a full hit list moves into the arena with twice the capacity. The simulator starts every read with room for ONE hit
    # (GWSIM_HITCAP), so every read that reports a split pair or several equally good alignments goes through it, some of them repeatedly.
    monkeypatch.setenv("GWSIM_HITCAP", "1")
    g = synth.inject_repeats(synth.random_genome(40_000, 41), 42, families=4, min_len=150, max_len=500, copies=(6, 20))
    fa, oi, si = _mk(tmp_path, "h.fa", [("c", g)])
    reads = synth.simulate_reads(g, 300, 100, 3, 43)
    cut = 45
    reads += [g[1000:1000 + cut].tobytes().decode() + g[9000:9000 + 100 - cut].tobytes().decode() for _ in range(3)]   # split alignments
    for cfg in (dict(k=3, report_type=1), dict(k=3, report_type=2, top_l=4), dict(k=3)):
        ro, rs = _compare(oi, si, reads, **cfg)
        assert max(len(x) for x in ro["sam"]) >= 2   # some read does report more than one record
    oi.close()
    si.close()
