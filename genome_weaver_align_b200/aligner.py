"""Host-side mirror of the reference's aligner interface over the C-ABI (include/gw_align.h).

Reference types mirrored (paths relative to align/src/main/java/org/utgenome/weaver/):
  FMIndexOnGenome            align/FMIndexOnGenome.java:60-119      load(fasta) / build_from_sequence(name, seq)
  AlignmentConfig            align/AlignmentConfig.java:40-72, align/AlignmentScoreConfig.java:37-77
  BidirectionalSuffixFilter  align/strategy/BidirectionalSuffixFilter.java:97-125   new BSF(fmIndex, reference, config); align(reads)
  Aligner / Reporter         align/Aligner.java:30-33, parallel/Reporter.java:27-30   align(read, out) / out.emit(record)
  SAMOutput                  align/SAMOutput.java:51-82
  BWTransform ("bwt")        align/BWTransform.java:72-90
The JVM reference aligns one read per call; here a call takes a batch (the device runs all reads of the batch concurrently) and
reports results in input order, which is the order the reference's sequential loop produces.
"""
import ctypes as C

import numpy as np

from . import _lib as L

REPORT_TYPES = {"besthit": 0, "allhits": 1, "topl": 2}


class AlignmentConfig:
    """-k -s -M -N -S -P -W -R -L of `genome-weaver align` (same names and defaults as the reference)"""

    def __init__(self, k=0.1, numSplitAlowed=1, matchScore=1, mismatchPenalty=3, splitOpenPenalty=11, indelEndSkip=5, bandWidth=31,
                 reportType="besthit", topL=5, extended=False):
        self.k, self.numSplitAlowed, self.matchScore, self.mismatchPenalty = k, numSplitAlowed, matchScore, mismatchPenalty
        self.splitOpenPenalty, self.indelEndSkip, self.bandWidth, self.topL = splitOpenPenalty, indelEndSkip, bandWidth, topL
        self.reportType = reportType
        # not in the reference: reads on which its StaircaseFilter constructor throws (m >= 156) are aligned with the oracle's `extended`
        # semantics (out-of-range bit-vector blocks read as 0) instead of coming back READ_REF_ABORT
        self.extended = extended

    def getMaximumEditDistance(self, readLength):
        return L.lib().gw_max_edit_distance_c(C.c_float(self.k), readLength)

    def _c(self):
        c = L.Config()
        L.lib().gw_config_default(C.byref(c))
        c.k = self.k
        c.num_split_allowed, c.match_score, c.mismatch_penalty = self.numSplitAlowed, self.matchScore, self.mismatchPenalty
        c.split_open_penalty, c.indel_end_skip, c.band_width, c.top_l = self.splitOpenPenalty, self.indelEndSkip, self.bandWidth, self.topL
        rt = self.reportType
        c.report_type = REPORT_TYPES[rt.lower()] if isinstance(rt, str) else int(rt)
        c.extended = 1 if self.extended else 0
        return c


def bwt(fasta_path, device=0):
    """`genome-weaver bwt <fasta>`: writes <prefix>.i.silk .f.3bit .r.3bit .f.ssa .r.ssa .f.bwt .r.bwt (suffix sorting on the device)"""
    L.check(L.lib().gw_index_build(fasta_path.encode(), device))


class FMIndexOnGenome:
    """the opened index: both FM-indexes, sampled suffix arrays and the forward reference, resident in HBM of one device"""

    def __init__(self, handle):
        self._h = handle

    @classmethod
    def load(cls, fasta_path, device=0):
        h = C.c_void_p()
        L.check(L.lib().gw_index_open(fasta_path.encode(), device, C.byref(h)))
        return cls(h)

    @classmethod
    def buildFromSequence(cls, name, seq, device=0):
        h = C.c_void_p()
        L.check(L.lib().gw_index_from_sequence(name.encode(), seq.encode(), device, C.byref(h)))
        return cls(h)

    def locateTable(self, step):
        """suffix-array value of every step-th row in HBM (1, 2, 4, 8, 16; 32: the .ssa samples only; 0: automatic, as load() does)"""
        L.check(L.lib().gw_index_locate_table(self._h, step))

    def locateStep(self):
        return L.lib().gw_index_locate_step(self._h)

    def textSize(self):
        return L.lib().gw_index_text_size(self._h)

    def contigs(self):
        lib = L.lib()
        return [(lib.gw_index_contig_name(self._h, i).decode(), lib.gw_index_contig_length(self._h, i), lib.gw_index_contig_offset(self._h, i))
                for i in range(lib.gw_index_num_contigs(self._h))]

    def device(self):
        return L.lib().gw_index_device(self._h)

    def toSAMHeader(self):
        n = C.c_uint64(0)
        L.lib().gw_sam_header(self._h, None, C.byref(n))
        buf = C.create_string_buffer(max(int(n.value), 1))
        L.check(L.lib().gw_sam_header(self._h, C.cast(buf, C.c_void_p), C.byref(n)))
        return buf.raw[:n.value].decode()

    # ---- single reference methods (parity tests)
    def rankACGTN(self, which, positions):
        pos = np.ascontiguousarray(positions, dtype=np.uint32)
        out = np.zeros((len(pos), 5), dtype=np.uint32)
        L.check(L.lib().gw_rank_acgtn(self._h, which, L.ptr(pos), len(pos), L.ptr(out)))
        return out

    def locate(self, which, sa_rows):
        rows = np.ascontiguousarray(sa_rows, dtype=np.uint32)
        out = np.zeros(len(rows), dtype=np.uint32)
        L.check(L.lib().gw_locate(self._h, which, L.ptr(rows), len(rows), L.ptr(out)))
        return out

    def quickScan(self, seqs, strand, cooperative=1):
        b, off = L.concat(seqs)
        out = np.zeros((len(seqs), 5), dtype=np.uint32)
        L.check(L.lib().gw_quick_scan(self._h, L.ptr(b), L.ptr(off), len(seqs), strand, cooperative, L.ptr(out)))
        return out.astype(np.int64)

    def close(self):
        if self._h:
            L.lib().gw_index_close(self._h)
            self._h = None


class AlignmentBatch:
    """hit records of one batch (gw_hit_t array + BAM-encoded cigar elements), convertible to SAM text"""

    def __init__(self, index, seqs_bytes, seq_off, hits, cigars, counters, timing, read_counters=None):
        self.index, self.bases, self.base_off, self.hits, self.cigars, self.counters, self.timing = index, seqs_bytes, seq_off, hits, cigars, counters, timing
        # (n, 4) uint32: numFMIndexSearches, numSW, numCutOff, numFiltered of every read (BidirectionalSuffixFilter.java:588-592)
        self.read_counters = read_counters

    def sam(self, names, quals=None):
        """SAMOutput.emit of every record, in input order; returns (text, lines_per_read)"""
        n = len(self.base_off) - 1
        nb, noff = L.concat(names)
        qb, qoff = (None, None) if quals is None else L.concat(quals)
        lines = np.zeros(max(n, 1), dtype=np.uint32)
        cap = C.c_uint64(max(64, 400 * len(self.hits) + int(self.base_off[-1]) * 2 + len(nb) * 2))
        while True:
            buf = C.create_string_buffer(int(cap.value))
            want = C.c_uint64(cap.value)
            rc = L.lib().gw_sam_format(self.index._h, L.ptr(nb), L.ptr(noff), L.ptr(self.bases), L.ptr(self.base_off), L.ptr(qb), L.ptr(qoff), n,
                                       L.ptr(self.hits), len(self.hits), L.ptr(self.cigars), C.cast(buf, C.c_void_p), C.byref(want), L.ptr(lines))
            if rc == L.GW_E_CAPACITY:
                cap = C.c_uint64(int(want.value) + 64)
                continue
            L.check(rc)
            return buf.raw[:want.value].decode(), lines[:n]

    def sam_per_read(self, names, quals=None):
        text, lines = self.sam(names, quals)
        all_lines = text.split("\n")[:-1] if text else []
        out, p = [], 0
        for c in lines:
            out.append(all_lines[p:p + int(c)])
            p += int(c)
        return out

    def status(self):
        """GW_READ_* per read (from the head records)"""
        n = len(self.base_off) - 1
        st = np.full(n, -1, dtype=np.int32)
        heads = self.hits[self.hits["is_head"] == 1]
        st[heads["read_index"][::-1]] = heads["status"][::-1]
        return st


class BidirectionalSuffixFilter:
    """the BSF aligner: new BidirectionalSuffixFilter(fmIndex, reference, config). The forward reference is part of the opened index."""

    def __init__(self, fmIndex, config=None, **options):
        self.fmIndex = fmIndex
        self.config = config or AlignmentConfig()
        h = C.c_void_p()
        c = self.config._c()
        L.check(L.lib().gw_aligner_create(fmIndex._h, C.byref(c), C.byref(h)))
        self._h = h
        for k, v in options.items():
            self.set_option(k, v)

    def set_option(self, name, value):
        L.check(L.lib().gw_aligner_set_option(self._h, name.encode(), int(value)))

    def align(self, seqs):
        """Aligner.align for a batch of read sequences (str/bytes); returns an AlignmentBatch"""
        b, off = L.concat(seqs)
        n = len(seqs)
        cnt = L.Counters()
        cap_h, cap_c = max(2 * n, 16), max(8 * n, 64)
        while True:
            hits = np.zeros(cap_h, dtype=L.HIT_DTYPE)
            cig = np.zeros(cap_c, dtype=np.uint32)
            nh, nc = C.c_uint64(cap_h), C.c_uint64(cap_c)
            rc = L.lib().gw_align_batch(self._h, L.ptr(b), L.ptr(off), n, L.ptr(hits), C.byref(nh), L.ptr(cig), C.byref(nc), C.byref(cnt))
            if rc == L.GW_E_CAPACITY:   # the batch is aligned and its records are retained on the device: fetch them into larger buffers
                cap_h, cap_c = int(nh.value) + 16, int(nc.value) + 64
                hits = np.zeros(cap_h, dtype=L.HIT_DTYPE)
                cig = np.zeros(cap_c, dtype=np.uint32)
                nh, nc = C.c_uint64(cap_h), C.c_uint64(cap_c)
                rc = L.lib().gw_align_batch_fetch(self._h, L.ptr(hits), C.byref(nh), L.ptr(cig), C.byref(nc))
            L.check(rc)
            break
        ms = (C.c_float * 6)()
        L.lib().gw_aligner_last_timing(self._h, ms)
        rc4 = np.zeros((max(n, 1), 4), dtype=np.uint32)
        L.check(L.lib().gw_aligner_read_counters(self._h, 0, n, L.ptr(rc4)))
        return AlignmentBatch(self.fmIndex, b, off, hits[:nh.value].copy(), cig[:max(int(nc.value), 1)].copy(), cnt.as_dict(), list(ms), rc4[:n])

    def close(self):
        if self._h:
            L.lib().gw_aligner_destroy(self._h)
            self._h = None


class SAMOutput:
    """Reporter that prints AlignmentRecords as SAM (align/SAMOutput.java): init() writes the @SQ header, emit() the lines"""

    def __init__(self, fmIndex, out):
        self.fmIndex, self.out = fmIndex, out

    def init(self):
        self.out.write(self.fmIndex.toSAMHeader())

    def emit(self, batch, names, quals=None):
        text, _ = batch.sam(names, quals)
        self.out.write(text)


_OPS = "MIDNSHPX"


def cigar_string(ops):
    return "".join("%d%s" % (int(v) >> 4, _OPS[int(v) & 7]) for v in ops)


def alignBlockDetailed(pairs, k, device=0):
    """BitParallelSmithWaterman.alignBlockDetailed(ref, query, k) for a list of (ref, query) strings; None where the reference returns null"""
    refs, ro = L.concat([r.replace(" ", "") for r, _ in pairs])
    qs, qo = L.concat([q.replace(" ", "") for _, q in pairs])
    n = len(pairs)
    res = np.zeros((max(n, 1), 4), dtype=np.int32)
    stride = 40
    cig = np.zeros((max(n, 1), stride), dtype=np.uint32)
    L.check(L.lib().gw_verify_pairs(device, L.ptr(refs), L.ptr(ro), L.ptr(qs), L.ptr(qo), n, k, L.ptr(res), L.ptr(cig), stride))
    out = []
    for i in range(n):
        if not res[i, 0]:
            out.append(None)
        else:
            m = int(qo[i + 1] - qo[i])
            out.append(dict(pos=int(res[i, 1]), nm=int(res[i, 2]), score=m - int(res[i, 2]), cigar=cigar_string(cig[i, :res[i, 3]])))
    return out


def staircase_table(m, kk, extended=False):
    """StaircaseFilter(m, kk) masks as a (kk+1, ceil(m/64)) uint64 array, and whether the reference constructor throws"""
    nb = (m + 63) // 64
    out = np.zeros((kk + 1) * nb, dtype=np.uint64)
    rc = L.lib().gw_staircase_table(m, kk, 1 if extended else 0, L.ptr(out), len(out))
    if rc < 0:
        L.check(rc)
    return out.reshape(kk + 1, nb), rc == 2


def cyclic_suffix_array(seq, device=0):
    b = seq.encode() if isinstance(seq, str) else bytes(seq)
    out = np.zeros(max(len(b), 1), dtype=np.uint32)
    L.check(L.lib().gw_cyclic_suffix_array(L.ptr(b), len(b), device, L.ptr(out)))
    return out[:len(b)]
