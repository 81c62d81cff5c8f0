"""CPU fuzz of the device automaton (the kernel bodies compiled by tests/hostsim) against the oracle: SAM lines, status and the
four per-read counters, over random / repeat-rich / low-complexity genomes, read lengths 20..256, k 0..6 and fractional k,
BESTHIT / ALLHITS / TOPL, small work areas.  usage: python tools/fuzz_hostsim.py [n_rounds] [seed]"""
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

import hostsim_lib as H  # noqa: E402
import oracle_lib as O  # noqa: E402
from genome_weaver_align_b200 import synth  # noqa: E402


def chimeric(g, n, length, seed):
    """reads made of two genome pieces (split alignments) with a few edits"""
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n):
        cut = int(rng.integers(8, length - 8))
        a, b = int(rng.integers(0, len(g) - length)), int(rng.integers(0, len(g) - length))
        r = bytearray(g[a:a + cut].tobytes() + g[b:b + length - cut].tobytes())
        for _e in range(int(rng.integers(0, 3))):
            r[int(rng.integers(0, length))] = b"ACGT"[int(rng.integers(0, 4))]
        s = r.decode()
        out.append(O.revcomp(s) if rng.random() < 0.5 else s)
    return out


def main():
    rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 6
    seed0 = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    d = tempfile.mkdtemp(prefix="gw_fuzz_")
    total, t0 = 0, time.time()
    for rd in range(rounds):
        seed = seed0 * 1000 + rd
        rng = np.random.default_rng(seed)
        kind = rd % 3
        n_g = int(rng.integers(30_000, 120_000))
        if kind == 0:
            g = synth.random_genome(n_g, seed, n_frac=0.004)
        elif kind == 1:
            g = synth.inject_repeats(synth.random_genome(n_g, seed), seed + 1, families=5, min_len=150, max_len=700, copies=(4, 25))
        else:
            g = synth.low_complexity_genome(n_g, seed)
        fa = os.path.join(d, "g%d.fa" % rd)
        cut = n_g // 3
        synth.write_fasta(fa, [("c1", g[:cut]), ("c2", g[cut:])])
        O.OracleIndex.build_files(fa)
        oi, si = O.OracleIndex.load(fa), H.SimIndex(fa)
        cases = []
        for (k, L, n) in [(3, 100, 600), (2, 100, 300), (3, 36, 300), (3, 150, 250), (6, 150, 120), (0.1, 75, 200), (1, 50, 200), (0, 40, 100),
                          (4, 128, 150), (3, 64, 150), (3, 65, 150), (6, 250, 60), (3, 200, 80)]:
            e = int(k) if k >= 1 else 6
            cases.append((dict(k=k), synth.simulate_reads(g, n, L, e, seed + L + int(10 * k))))
        cases.append((dict(k=3), chimeric(g, 400, 100, seed + 7)))
        cases.append((dict(k=3, report_type=1), synth.simulate_reads(g, 300, 100, 3, seed + 8) + chimeric(g, 100, 80, seed + 9)))
        cases.append((dict(k=3, report_type=2, top_l=3), synth.simulate_reads(g, 200, 100, 3, seed + 10)))
        cases.append((dict(k=3, num_split_allowed=0), synth.simulate_reads(g, 200, 100, 4, seed + 11)))
        ragged = [g[i * 97:i * 97 + 1 + (i * 7) % 256].tobytes().decode() for i in range(120)]
        cases.append((dict(k=0.1), ragged))
        for ci, (cfg, reads) in enumerate(cases):
            names = ["r%d" % i for i in range(len(reads))]
            quals = ["I" * len(r) for r in reads]
            ro = oi.align(names, reads, quals, **cfg)
            for slab in ((4096, 1024) if ci % 4 == 0 else (4096,)):
                rs = si.align(names, reads, quals, slab_words=slab, **cfg)
                bad = [i for i in range(len(reads)) if ro["sam"][i] != rs["sam"][i] or ro["status"][i] != rs["status"][i] or
                       (ro["status"][i] != 2 and ro["counters"][i] != rs["counters"][i])]
                if bad:
                    i = bad[0]
                    print("MISMATCH round %d case %d cfg %r slab %d: %d reads; first %d\n  read %s\n  oracle %r %r %r\n  sim    %r %r %r" % (
                        rd, ci, cfg, slab, len(bad), i, reads[i], ro["sam"][i], ro["status"][i], ro["counters"][i], rs["sam"][i], rs["status"][i],
                        rs["counters"][i]))
                    sys.exit(1)
            total += len(reads)
        oi.close()
        si.close()
        print("round %d (genome kind %d, %d bp): ok, %d reads so far, %.0f s" % (rd, kind, n_g, total, time.time() - t0), flush=True)
    print("fuzz ok: %d reads, 0 differences" % total)


if __name__ == "__main__":
    main()
