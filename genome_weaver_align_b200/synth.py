"""Synthetic genomes and reads for the configs of BASELINE.json / SURVEY.md 8(d).

Deterministic (numpy PCG64 with the stated seeds). Used by tests and bench.py to create inputs; not part of
the alignment path itself."""
import numpy as np

_ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
_COMP = np.zeros(256, dtype=np.uint8)
for a, b in zip(b"ACGTN", b"TGCAN"):
    _COMP[a] = b


def random_genome(n, seed, n_frac=0.0):
    """i.i.d. uniform ACGT genome as a uint8 ASCII array; optional runs of N covering about n_frac of it."""
    rng = np.random.default_rng(seed)
    g = _ACGT[rng.integers(0, 4, size=n, dtype=np.uint8)]
    if n_frac > 0:
        run = 50
        for s in rng.integers(0, max(1, n - run), size=max(1, int(n * n_frac / run))):
            g[s:s + run] = ord("N")
    return g


def inject_repeats(g, seed, families=20, min_len=300, max_len=6000, copies=(100, 1000), max_div=0.10):
    """copy `families` repeat units over the genome with 0..max_div divergence (cfg 3 style)."""
    rng = np.random.default_rng(seed)
    n = len(g)
    for _ in range(families):
        L = int(rng.integers(min_len, max_len + 1))
        unit = _ACGT[rng.integers(0, 4, size=L, dtype=np.uint8)]
        for _c in range(int(rng.integers(copies[0], copies[1] + 1))):
            if n <= L + 1000:
                break
            pos = int(rng.integers(0, n - L - 1000))
            cp = unit.copy()
            div = rng.random() * max_div
            nm = int(L * div)
            if nm:
                idx = rng.integers(0, L, size=nm)
                cp[idx] = _ACGT[rng.integers(0, 4, size=nm, dtype=np.uint8)]
            g[pos:pos + L] = cp
    return g


def low_complexity_genome(n, seed):
    """tandem repeats (unit 2..50 bp) and 2-/3-letter biased segments (cfg 5 style)."""
    rng = np.random.default_rng(seed)
    out = np.empty(n, dtype=np.uint8)
    p = 0
    while p < n:
        seg = int(rng.integers(200, 5000))
        seg = min(seg, n - p)
        kind = rng.integers(0, 3)
        if kind == 0:
            u = _ACGT[rng.integers(0, 4, size=int(rng.integers(2, 51)), dtype=np.uint8)]
            out[p:p + seg] = np.resize(u, seg)
        elif kind == 1:
            letters = rng.permutation(4)[: int(rng.integers(2, 4))]
            out[p:p + seg] = _ACGT[letters[rng.integers(0, len(letters), size=seg)]]
        else:
            out[p:p + seg] = _ACGT[rng.integers(0, 4, size=seg, dtype=np.uint8)]
        p += seg
    return out


def write_fasta(path, contigs, width=70):
    """contigs: list of (name, uint8 array)"""
    with open(path, "wb") as f:
        for name, seq in contigs:
            f.write(b">" + name.encode() + b"\n")
            n = len(seq)
            full = n // width * width
            if full:
                body = np.empty((full // width, width + 1), dtype=np.uint8)
                body[:, :width] = seq[:full].reshape(-1, width)
                body[:, width] = 10
                f.write(body.tobytes())
            if n > full:
                f.write(seq[full:].tobytes() + b"\n")


def revcomp(a):
    return _COMP[a[::-1]]


def simulate_reads(genome, n_reads, length, max_edits, seed, end_skip=5, exact_edits=None):
    """returns list of ASCII strings. Each read: uniform start, strand 50/50, e ~ U{0..max_edits} edits
    (70% substitution / 15% insertion / 15% deletion, none within end_skip bp of the read ends)."""
    rng = np.random.default_rng(seed)
    n = len(genome)
    reads = []
    starts = rng.integers(0, n - length - max_edits - 1, size=n_reads)
    strands = rng.integers(0, 2, size=n_reads)
    nedits = rng.integers(0, max_edits + 1, size=n_reads) if exact_edits is None else np.full(n_reads, exact_edits)
    for i in range(n_reads):
        s = int(starts[i])
        frag = list(genome[s:s + length + max_edits + 1].tobytes().decode())
        for _ in range(int(nedits[i])):
            pos = int(rng.integers(end_skip + 1, length - end_skip - 1))
            r = rng.random()
            if r < 0.70:
                c = frag[pos]
                frag[pos] = "ACGT"[("ACGT".find(c) + 1 + int(rng.integers(0, 3))) % 4] if c in "ACGT" else "A"
            elif r < 0.85:
                frag.insert(pos, "ACGT"[int(rng.integers(0, 4))])
            else:
                del frag[pos]
        rd = "".join(frag[:length])
        if strands[i]:
            rd = rd[::-1].translate(str.maketrans("ACGTN", "TGCAN"))
        reads.append(rd)
    return reads


def simulate_reads_fast(genome, n_reads, length, max_edits, seed, end_skip=5, exact_edits=None):
    """vectorised variant for millions of reads: substitutions, single-base insertions and deletions applied with
    numpy gathers (same error model as simulate_reads, different random stream). Returns a (n_reads, length) uint8 array."""
    rng = np.random.default_rng(seed)
    n = len(genome)
    span = length + max_edits + 1
    starts = rng.integers(0, n - span - 1, size=n_reads)
    idx = starts[:, None] + np.arange(span)[None, :]
    # source offset map: off[r, j] = genome offset feeding read position j
    off = np.tile(np.arange(length, dtype=np.int64), (n_reads, 1))
    sub_mask = np.zeros((n_reads, length), dtype=bool)
    ins_mask = np.zeros((n_reads, length), dtype=bool)
    nedits = rng.integers(0, max_edits + 1, size=n_reads)
    if exact_edits is not None:
        nedits = np.full(n_reads, exact_edits)
    for e in range(max_edits):
        active = nedits > e
        pos = rng.integers(end_skip + 1, length - end_skip - 1, size=n_reads)
        r = rng.random(n_reads)
        is_sub = active & (r < 0.70)
        is_ins = active & (r >= 0.70) & (r < 0.85)
        is_del = active & (r >= 0.85)
        rows = np.nonzero(is_sub)[0]
        sub_mask[rows, pos[rows]] = True
        cols = np.arange(length)[None, :]
        # insertion at pos: positions > pos read one genome base earlier; position pos is random
        rows = np.nonzero(is_ins)[0]
        if len(rows):
            ins_mask[rows, pos[rows]] = True
            off[rows] -= (cols > pos[rows, None]).astype(np.int64)
        rows = np.nonzero(is_del)[0]
        if len(rows):
            off[rows] += (cols >= pos[rows, None]).astype(np.int64)
    off = np.clip(off, 0, span - 1)
    reads = genome[np.take_along_axis(idx, off, axis=1)]
    rnd = _ACGT[rng.integers(0, 4, size=(n_reads, length), dtype=np.uint8)]
    # substitution: force a different base by rotating within ACGT
    code = np.zeros(256, dtype=np.uint8)
    for i, c in enumerate(b"ACGT"):
        code[c] = i
    rot = _ACGT[(code[reads] + 1 + rng.integers(0, 3, size=reads.shape, dtype=np.uint8)) % 4]
    reads = np.where(sub_mask, rot, reads)
    reads = np.where(ins_mask, rnd, reads)
    strands = rng.integers(0, 2, size=n_reads).astype(bool)
    rc = _COMP[reads[:, ::-1]]
    reads = np.where(strands[:, None], rc, reads)
    return np.ascontiguousarray(reads)
