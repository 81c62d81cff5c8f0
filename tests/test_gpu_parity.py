"""GPU parity tests: the CUDA path, called through the C-ABI (libgwalign.so), against the CPU oracle on the same seeded inputs,
against the reference's golden vectors, and through size-independent properties. Integer work: the bar is bit-exact."""
import os
import random

import numpy as np
import pytest

import oracle_lib as O

pytestmark = pytest.mark.gpu

gw = None


@pytest.fixture(scope="module", autouse=True)
def _load():
    global gw
    import genome_weaver_align_b200 as g
    gw = g
    yield


from genome_weaver_align_b200 import synth  # noqa: E402  (pure numpy helper; does not load the CUDA library)

REF16 = "AAGCCTAGTTTCCTTG"


def names_quals(reads):
    return ["r%d" % i for i in range(len(reads))], ["I" * len(r) for r in reads]


def compare(idx, oidx, reads, config, ocfg, **opts):
    names, quals = names_quals(reads)
    bsf = gw.BidirectionalSuffixFilter(idx, config, **opts)
    batch = bsf.align(reads)
    got = batch.sam_per_read(names, quals)
    want = oidx.align(names, reads, quals, **ocfg)
    st = batch.status()
    bad = [i for i in range(len(reads)) if got[i] != want["sam"][i] or int(st[i]) != want["status"][i]]
    assert not bad, "%d/%d reads differ; first %d: %r vs %r" % (len(bad), len(reads), bad[0], got[bad[0]], want["sam"][bad[0]])
    # the four per-read debug counters of the reference (numFMIndexSearches, numSW, numCutOff, numFiltered), read by read; reads on
    # which the reference throws stop at an implementation-defined point, so only their status is compared
    okc = [i for i in range(len(reads)) if want["status"][i] != 2]
    badc = [i for i in okc if list(map(int, batch.read_counters[i])) != list(want["counters"][i])]
    assert not badc, "%d/%d reads differ in counters; first %d: %r vs %r" % (len(badc), len(okc), badc[0], list(batch.read_counters[badc[0]]), want["counters"][badc[0]])
    if len(okc) == len(reads):
        for j, name in enumerate(("fm_searches", "sw_calls", "cutoffs", "filtered")):
            assert batch.counters[name] == sum(want["counters"][i][j] for i in okc), name
    bsf.close()
    return batch, want


# ---- golden vectors of the reference (strategy/BidirectionalSuffixFilterTest.java:47-242), through the device
GOLDEN = [("GCCTAC", "5M1S", 3, "+", 0), ("GCCAAGTT", "8M", 3, "+", 1), (O.revcomp("GCGTAGTT"), "8M", 3, "-", 1),
          ("AACCCTAGTTTCGTT", "15M", 1, "+", 2), (O.revcomp("AACCCTAGTTTCGTT"), "15M", 1, "-", 2), ("TTGCCTAGTTT", "2S9M", 3, "+", 0),
          ("GCCTAGT", "7M", 3, "+", 0), (O.revcomp("GCCTAGT"), "7M", 3, "-", 0), ("AAGCCTAGATTCCGTG", "16M", 1, "+", 2),
          ("AAGCCTAGTTAAAAAA", "10M6S", 1, "+", None), ("AAGCCTGTTT", "6M1D4M", 1, "+", 1), ("AAGCCTCAGTT", "6M1I4M", 1, "+", 1)]


def test_reference_golden_vectors():
    idx = gw.FMIndexOnGenome.buildFromSequence("seq", REF16)
    bsf = gw.BidirectionalSuffixFilter(idx, gw.AlignmentConfig(k=2))
    reads = [g[0] for g in GOLDEN] + ["AAGCCTATCCTTG"]
    names, quals = names_quals(reads)
    sam = bsf.align(reads).sam_per_read(names, quals)
    for (q, cigar, start, strand, nm), lines in zip(GOLDEN, sam):
        a = O.parse_sam(lines[0])
        assert (a["cigar"], a["start"], a["strand"]) == (cigar, start, strand), q
        if nm is not None:
            assert a["nm"] == nm
    split = sam[-1]   # :144-158 split alignment -> two records
    assert len(split) == 2
    a, s = O.parse_sam(split[0]), O.parse_sam(split[1])
    assert (a["cigar"], a["start"], s["cigar"], s["start"]) == ("7M", 1, "6M", 11)
    bsf.close()
    idx.close()


def test_verifier_golden_vectors():
    # BitParallelSmithWatermanTest.java:184-240
    cases = [("GATCTA", "GCTATA", 2), ("TATAATAATA", "TAATA", 0)]
    got = gw.alignBlockDetailed(cases_pairs := [(r, q) for r, q, _ in cases], 31)
    for (r, q, _), g in zip(cases, got):
        assert g == {k: v for k, v in O.align_block_detailed(r, q, 31).items() if k in g}
    rnd = random.Random(5)
    pairs = []
    for t in range(600):
        m = rnd.randint(1, 250)
        q = "".join(rnd.choice("ACGT") for _ in range(m))
        ref = list("".join(rnd.choice("ACGT") for _ in range(rnd.randint(0, 8))) + q + "".join(rnd.choice("ACGT") for _ in range(rnd.randint(0, 8))))
        for _e in range(rnd.randint(0, 6 if t % 3 else 18)):   # every third pair: many differences (the cigar still fits the 40 elements asked for)
            p = rnd.randrange(len(ref))
            r = rnd.random()
            if r < 0.5:
                ref[p] = rnd.choice("ACGTN")
            elif r < 0.75:
                ref.insert(p, rnd.choice("ACGT"))
            elif len(ref) > 1:
                del ref[p]
        pairs.append(("".join(ref), q))
    pairs.append(("", "ACGT"))
    pairs.append(("ACGTACGT", "A"))
    got = gw.alignBlockDetailed(pairs, 31)
    for (r, q), g in zip(pairs, got):
        w = O.align_block_detailed(r, q, 31)
        assert (g is None) == (w is None), (r, q)
        if g:
            assert (g["cigar"], g["pos"], g["nm"]) == (w["cigar"], w["pos"], w["nm"]), (r, q)


@pytest.fixture(scope="module")
def small(tmp_path_factory):
    d = tmp_path_factory.mktemp("small")
    g = synth.random_genome(300_000, 31, n_frac=0.003)
    fa = str(d / "g.fa")
    synth.write_fasta(fa, [("chr1", g[:100_000]), ("chr2", g[100_000:250_000]), ("chr3", g[250_000:])])
    gw.bwt(fa, 0)
    idx = gw.FMIndexOnGenome.load(fa, 0)
    oidx = O.OracleIndex.load(fa)   # same files: the oracle reads what the device builder wrote
    yield dict(g=g, fa=fa, idx=idx, oidx=oidx, dir=d)
    idx.close()
    oidx.close()


def test_index_files_match_oracle_builder(small, tmp_path):
    # `genome-weaver bwt` on the device writes the same seven files as the oracle's builder (same SA of cyclic rotations)
    import shutil
    fa2 = str(tmp_path / "g.fa")
    shutil.copy(small["fa"], fa2)
    O.OracleIndex.build_files(fa2)
    for ext in (".i.silk", ".f.3bit", ".r.3bit", ".f.bwt", ".r.bwt", ".f.ssa", ".r.ssa"):
        assert open(small["fa"] + ext, "rb").read() == open(fa2 + ext, "rb").read(), ext


def test_cyclic_suffix_array_golden_and_random():
    assert list(gw.cyclic_suffix_array("TATAATAATATAATA")) == sorted(range(15), key=lambda i: ("TATAATAATATAATA" * 2)[i:i + 15])
    rnd = random.Random(3)
    for n in (1, 2, 5, 26, 27, 28, 100, 1000):
        s = "".join(rnd.choice("ACGTN") for _ in range(n))
        want = sorted(range(n), key=lambda i: ([("ACGTN".index(c)) for c in (s + s)[i:i + n]], i))
        assert list(gw.cyclic_suffix_array(s)) == want, n
    per = "ACGTTGCA" * 16   # non-primitive text: ties broken by position
    n = len(per)
    assert list(gw.cyclic_suffix_array(per)) == sorted(range(n), key=lambda i: ((per + per)[i:i + n], i))


def test_rank_and_locate(small):
    idx, oidx = small["idx"], small["oidx"]
    N = idx.textSize()
    rnd = np.random.default_rng(1)
    pos = np.concatenate([rnd.integers(0, N + 1, size=2000), [0, 1, 127, 128, 129, N - 1, N, N + 5]]).astype(np.uint32)
    for which in (0, 1):
        got = idx.rankACGTN(which, pos)
        for p, g in zip(pos[::17], got[::17]):
            assert list(map(int, g)) == oidx.rank_acgtn(which, int(p))
        rows = rnd.integers(0, N, size=500).astype(np.uint32)
        loc = idx.locate(which, rows)
        for r, l in zip(rows[::5], loc[::5]):
            assert int(l) == oidx.locate(which, int(r))


def test_locate_table_steps(small):
    # the locate table built on the device (suffix-array value of every step-th row) must give what SparseSuffixArray.get gives,
    # whatever its density; alignment results must not depend on it either
    idx, oidx = small["idx"], small["oidx"]
    N = idx.textSize()
    assert idx.locateStep() == 1   # what load() chose: the whole suffix array fits easily
    rnd = np.random.default_rng(2)
    rows = np.concatenate([rnd.integers(0, N, size=3000), [0, 1, 31, 32, 33, N - 1]]).astype(np.uint32)
    want = [oidx.locate(0, int(r)) for r in rows[::3]], [oidx.locate(1, int(r)) for r in rows[::3]]
    reads = synth.simulate_reads(small["g"], 800, 100, 3, 41)
    try:
        for step in (1, 2, 16, 32, 4):
            idx.locateTable(step)
            assert idx.locateStep() == step
            for which in (0, 1):
                assert [int(x) for x in idx.locate(which, rows)[::3]] == want[which], (step, which)
            batch, _ = compare(idx, oidx, reads, gw.AlignmentConfig(k=3), dict(k=3))
            assert (batch.counters["occ_lookups_locate"] == 0) == (step == 1)
    finally:
        idx.locateTable(0)


@pytest.mark.parametrize("coop", [0, 1])
def test_quick_scan_both_forms(small, coop):
    reads = synth.simulate_reads(small["g"], 300, 100, 3, 40) + ["ACGT", "A", "ACGTNACGT" * 3]
    import ctypes as C
    for strand in (0, 1):
        got = small["idx"].quickScan(reads, strand, coop)
        for i in range(0, len(reads), 7):
            q = reads[i].replace("N", "A")
            if strand == 1:
                q = q.translate(str.maketrans("ACGT", "TGCA"))
                q = "".join("A" if reads[i][j] == "N" else q[j] for j in range(len(q)))
            # oracle: FMQuickScan == m backward-search steps with restart on the whole range
            lb, ub, mm, mark, ls, le, has = 0, small["oidx"].text_size(), 0, 0, 0, 0, False
            N = ub
            for j, ch in enumerate(q):
                lb, ub = small["oidx"].backward_search(1 if strand == 0 else 0, "ACGT".index(ch), lb, ub)
                if lb >= ub:
                    mm += 1
                    if not has or (le - ls) < (j - mark):
                        has, ls, le = True, mark, j
                    lb, ub, mark = 0, N, j + 1
            if not has or (le - ls) < (len(q) - mark):
                ls, le = mark, len(q)
            assert list(got[i]) == [mm, ls, le, lb, ub], (i, strand)


@pytest.mark.parametrize("k,length,n", [(2, 100, 3000), (3, 100, 3000), (3, 36, 1000), (3, 150, 1000), (6, 150, 400), (0.1, 75, 600)])
def test_sam_parity_random(small, k, length, n):
    reads = synth.simulate_reads(small["g"], n, length, int(k) if k >= 1 else 6, int(k * 10) + length)
    compare(small["idx"], small["oidx"], reads, gw.AlignmentConfig(k=k), dict(k=k))


def test_sam_parity_scalar_scan_and_small_slab(small):
    reads = synth.simulate_reads(small["g"], 1500, 100, 3, 77)
    compare(small["idx"], small["oidx"], reads, gw.AlignmentConfig(k=3), dict(k=3), scan_coop=0, slab_words=1024, wave_reads=400)


def test_sam_parity_large_work_areas_in_chunks(small):
    # reads that need a large work area share a buffer of tier_budget bytes: with a tiny budget they run in many chunks and their
    # outcomes are written back to the small work areas between chunks
    reads = synth.simulate_reads(small["g"], 1500, 100, 3, 79)
    compare(small["idx"], small["oidx"], reads, gw.AlignmentConfig(k=3), dict(k=3), slab_words=1024, tier_budget_mb=1)
    compare(small["idx"], small["oidx"], reads, gw.AlignmentConfig(k=3), dict(k=3), slab_words=768, tier_budget_mb=1, ext_pct=0)


@pytest.mark.parametrize("opts", [dict(slab_words=1024, ext_mult=2), dict(slab_words=1024, ext_pct=1, ext_mult=7), dict(slab_words=768, ext_mult=1),
                                  dict(slab_words=1024, ext_pct=0), dict(slab_words=1024, ext_mult=3),
                                  dict(slab_words=1024, ext_mult=2, big_slab_words=8192, ext2_pm=500),
                                  dict(slab_words=1024, ext_pct=2, ext_mult=2, big_slab_words=16384, ext2_pm=3),
                                  dict(slab_words=768, ext_mult=1, big_slab_words=4096, ext2_pm=1000), dict(slab_words=1024, ext2_pm=0),
                                  dict(mem_budget_mb=64), dict(mem_budget_mb=8, wave_reads=500)])
def test_sam_parity_moves_to_larger_work_areas(small, opts):
    # a read whose work area is nearly full is moved into a larger block taken from a pool and goes on there; when the pool is empty,
    # the block fills up too or the queue overflows, the read is run again from the start with a larger work area: the outcome must
    # not depend on which of these happened
    reads = synth.simulate_reads(small["g"], 1500, 100, 3, 80)
    compare(small["idx"], small["oidx"], reads, gw.AlignmentConfig(k=3), dict(k=3), **opts)


@pytest.mark.parametrize("opts", [dict(flow_batch=1, flow_age=1), dict(flow_batch=32, flow_age=1000, flow_flush=32),
                                  dict(flow_fetch=1, flow_flush=1, flow_blocks=3), dict(flow_blocks=1), dict(flow_blocks=1000),
                                  dict(flow_order=0), dict(flow_order=1, flow_blocks=7)])
def test_sam_parity_flow_kernel_tunables(small, opts):
    # how the warps of the flow kernel batch their work (deferred continuations, verification batches, grid size) must not change
    # any result
    reads = synth.simulate_reads(small["g"], 1500, 100, 3, 78)
    compare(small["idx"], small["oidx"], reads, gw.AlignmentConfig(k=3), dict(k=3), **opts)


def test_ragged_empty_and_edge_inputs(small):
    bsf = gw.BidirectionalSuffixFilter(small["idx"], gw.AlignmentConfig(k=3))
    assert len(bsf.align([]).hits) == 0                           # empty batch
    bsf.close()
    g = small["g"].tobytes().decode()
    reads = [g[1000:1100], g[5:41], "N" * 50, g[2000:2150], "ACGT", g[0:100], g[-100:], g[99_950:100_050], "A" * 300, ""]
    reads += [g[3000:3100][:50] + "NN" + g[3000:3100][52:], O.revcomp(g[7000:7250])]
    batch, want = compare(small["idx"], small["oidx"], [r for r in reads if 0 < len(r) <= 256], gw.AlignmentConfig(k=3), dict(k=3))
    bsf = gw.BidirectionalSuffixFilter(small["idx"], gw.AlignmentConfig(k=3))
    st = bsf.align(reads).status()
    assert st[8] == 3 and st[9] == 3                              # longer than 256 / empty: outside the supported domain, reported as such
    bsf.close()


def test_reference_abort_domain(small):
    # 250 bp reads with k=6: the reference throws on part of this domain; status 2 must agree read by read
    reads = synth.simulate_reads(small["g"], 200, 250, 6, 91)
    batch, want = compare(small["idx"], small["oidx"], reads, gw.AlignmentConfig(k=6), dict(k=6))
    assert 2 in want["status"]


def test_extended_domain_config4_reads(small):
    # BASELINE.json configs[3] reads (250 bp, k = 6) in the `extended` domain: no read comes back GW_READ_REF_ABORT, all are aligned
    # exactly as the oracle's extended mode aligns them
    reads = synth.simulate_reads(small["g"], 400, 250, 6, 92)
    batch, want = compare(small["idx"], small["oidx"], reads, gw.AlignmentConfig(k=6, extended=True), dict(k=6, extended=1))
    assert 2 not in want["status"] and want["status"].count(0) > 300


def test_repeats_all_hits_and_topl(tmp_path):
    g = synth.inject_repeats(synth.random_genome(150_000, 4), 5, families=6, min_len=200, max_len=900, copies=(5, 30))
    fa = str(tmp_path / "rep.fa")
    synth.write_fasta(fa, [("c1", g[:70_000]), ("c2", g[70_000:])])
    gw.bwt(fa, 0)
    idx, oidx = gw.FMIndexOnGenome.load(fa, 0), O.OracleIndex.load(fa)
    reads = synth.simulate_reads(g, 1500, 100, 3, 6)
    compare(idx, oidx, reads, gw.AlignmentConfig(k=3, reportType="allhits"), dict(k=3, report_type=1))
    compare(idx, oidx, reads[:500], gw.AlignmentConfig(k=3, reportType="topl", topL=3), dict(k=3, report_type=2, top_l=3))
    lc = synth.low_complexity_genome(80_000, 5)
    fa2 = str(tmp_path / "lc.fa")
    synth.write_fasta(fa2, [("lc", lc)])
    gw.bwt(fa2, 0)
    idx2, oidx2 = gw.FMIndexOnGenome.load(fa2, 0), O.OracleIndex.load(fa2)
    for L in (36, 50, 75, 100, 125, 150, 200, 250):   # the read-length sweep of cfg 5 (the reference throws for part of m >= 156: status compared too)
        compare(idx2, oidx2, synth.simulate_reads(lc, 160, L, 3, L), gw.AlignmentConfig(k=3, reportType="allhits"), dict(k=3, report_type=1))
    for x in (idx, idx2):
        x.close()


def test_properties_at_scale(small):
    # size-independent properties on a batch too large for the oracle to be the checker in seconds:
    # exact reads map to their origin with NM 0; results do not depend on batch composition/order; device-resident == host call
    g = small["g"]
    arr = synth.simulate_reads_fast(g, 60_000, 100, 3, 5)
    reads = [r.tobytes().decode() for r in arr]
    names, quals = names_quals(reads)
    bsf = gw.BidirectionalSuffixFilter(small["idx"], gw.AlignmentConfig(k=3))
    full = bsf.align(reads).sam_per_read(names, quals)
    perm = np.random.default_rng(0).permutation(len(reads))[:5000]
    sub = bsf.align([reads[i] for i in perm]).sam_per_read([names[i] for i in perm], [quals[i] for i in perm])
    for j, i in enumerate(perm):
        assert sub[j] == full[i]
    sample = list(range(0, len(reads), 60))
    want = small["oidx"].align([names[i] for i in sample], [reads[i] for i in sample], [quals[i] for i in sample], k=3)
    for j, i in enumerate(sample):
        assert full[i] == want["sam"][j]
    bsf.close()


def test_config1_full_size(tmp_path):
    # BASELINE.json configs[0] at its stated size (SURVEY.md 8d cfg 1): 4.6 Mbp i.i.d. genome (seed 1), 10 000 x 100 bp reads with
    # e ~ U{0,1,2} edits (seed 11), -k 2; parity = the whole SAM file, header included, plus the per-read counters
    g = synth.random_genome(4_600_000, 1)
    fa = str(tmp_path / "cfg1.fa")
    synth.write_fasta(fa, [("chr1", g)])
    gw.bwt(fa, 0)
    idx, oidx = gw.FMIndexOnGenome.load(fa, 0), O.OracleIndex.load(fa)
    assert idx.toSAMHeader() == oidx.sam_header()
    reads = synth.simulate_reads(g, 10_000, 100, 2, 11)
    compare(idx, oidx, reads, gw.AlignmentConfig(k=2), dict(k=2))
    idx.close()
    oidx.close()


def test_config5_read_length_sweep_5mbp(tmp_path):
    # a 5 Mbp instance of BASELINE.json configs[4]: low-complexity genome (seed 5), ALLHITS, all eight read lengths x 2 000 reads, -k 3
    # (for m >= 156 the reference throws on part of the reads: their status must agree read by read)
    lc = synth.low_complexity_genome(5_000_000, 5)
    fa = str(tmp_path / "cfg5.fa")
    synth.write_fasta(fa, [("lc", lc)])
    gw.bwt(fa, 0)
    idx, oidx = gw.FMIndexOnGenome.load(fa, 0), O.OracleIndex.load(fa)
    for L in (36, 50, 75, 100, 125, 150, 200, 250):
        compare(idx, oidx, synth.simulate_reads(lc, 2000, L, 3, 15 + L), gw.AlignmentConfig(k=3, reportType="allhits"), dict(k=3, report_type=1))
    idx.close()
    oidx.close()


def test_extended_k5_low_complexity_small_work_areas(tmp_path):
    # long reads with k = 5 on a low-complexity genome, ALLHITS, first work areas of 8 KB: most reads move into a block of the first size,
    # many from there into one of the second (their queue moves to the end of the block each time), the rest is run again in the passes
    # with larger work areas; the extended domain is on because part of these (m, k) make the reference's StaircaseFilter throw
    lc = synth.low_complexity_genome(400_000, 9)
    fa = str(tmp_path / "lc.fa")
    synth.write_fasta(fa, [("lc", lc)])
    gw.bwt(fa, 0)
    idx, oidx = gw.FMIndexOnGenome.load(fa, 0), O.OracleIndex.load(fa)
    reads = synth.simulate_reads(lc, 1200, 200, 5, 31) + synth.simulate_reads(lc, 800, 120, 4, 32)
    for opts in (dict(slab_words=1024), dict(slab_words=1024, ext_pct=3, ext2_pm=5), dict()):
        compare(idx, oidx, reads, gw.AlignmentConfig(k=5, reportType="allhits", extended=True), dict(k=5, report_type=1, extended=1), **opts)
    idx.close()
    oidx.close()


def test_fetch_after_capacity_error(small):
    # gw_align_batch with output buffers that are too small returns GW_E_CAPACITY and the sizes needed; gw_align_batch_fetch then hands
    # out the retained records of that batch (the batch is not aligned again) and they are what a call with large buffers returns
    import ctypes as C
    from genome_weaver_align_b200 import _lib as L
    reads = synth.simulate_reads(small["g"], 400, 100, 3, 83)
    bsf = gw.BidirectionalSuffixFilter(small["idx"], gw.AlignmentConfig(k=3))
    want = bsf.align(reads)
    b, off = L.concat(reads)
    hits, cig = np.zeros(4, dtype=L.HIT_DTYPE), np.zeros(4, dtype=np.uint32)
    nh, nc = C.c_uint64(4), C.c_uint64(4)
    cnt = L.Counters()
    rc = L.lib().gw_align_batch(bsf._h, L.ptr(b), L.ptr(off), len(reads), L.ptr(hits), C.byref(nh), L.ptr(cig), C.byref(nc), C.byref(cnt))
    assert rc == L.GW_E_CAPACITY and nh.value == len(want.hits) and nc.value == len(want.cigars)
    searches = cnt.fm_searches
    hits, cig = np.zeros(nh.value, dtype=L.HIT_DTYPE), np.zeros(nc.value, dtype=np.uint32)
    rc = L.lib().gw_align_batch_fetch(bsf._h, L.ptr(hits), C.byref(nh), L.ptr(cig), C.byref(nc))
    assert rc == 0 and hits.tobytes() == want.hits.tobytes() and cig.tobytes() == want.cigars.tobytes()
    assert cnt.fm_searches == searches == want.counters["fm_searches"]
    nh2, nc2 = C.c_uint64(1), C.c_uint64(1)
    assert L.lib().gw_align_batch_fetch(bsf._h, L.ptr(hits), C.byref(nh2), L.ptr(cig), C.byref(nc2)) == L.GW_E_CAPACITY and nh2.value == nh.value
    bsf.close()
