"""Writes tests/golden/*.json: SAM lines, per-read status and debug counters of the CPU oracle on small seeded workloads.
The oracle is pinned to the reference's own unit-test vectors (tests/test_oracle_golden.py); these files freeze its output on
random workloads, so that a later change of the oracle OR of the CUDA path shows up as a diff against a committed artefact.
usage: python tools/make_golden.py   (needs only the oracle: runs here, without a GPU)"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import oracle_lib as O
from genome_weaver_align_b200 import synth

CASES = {   # name: (genome_bp, genome_seed, n_frac, reads, length, max_edits, reads_seed, config)
    "random_100bp_k3": (200_000, 101, 0.0, 400, 100, 3, 201, dict(k=3)),
    "random_36bp_k2": (200_000, 101, 0.0, 300, 36, 2, 202, dict(k=2)),
    "ncontaining_150bp_k3_allhits": (150_000, 103, 0.01, 200, 150, 3, 203, dict(k=3, report_type=1)),
}


def genome_of(case):
    bp, seed, n_frac = CASES[case][:3]
    return synth.random_genome(bp, seed, n_frac=n_frac)


def reads_of(case, g):
    _, _, _, n, length, edits, seed, _ = CASES[case]
    return synth.simulate_reads(g, n, length, edits, seed)


def main():
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    for case, spec in CASES.items():
        g = genome_of(case)
        reads = reads_of(case, g)
        oidx = O.OracleIndex.from_sequence("chr1", g.tobytes().decode())
        names = ["r%d" % i for i in range(len(reads))]
        quals = ["I" * len(r) for r in reads]
        res = oidx.align(names, reads, quals, **spec[7])
        json.dump(dict(case=case, spec=[x if not isinstance(x, dict) else x for x in spec], sam=res["sam"], status=res["status"], counters=res["counters"]),
                  open(os.path.join(out_dir, case + ".json"), "w"), separators=(",", ":"))
        print(case, len(reads), "reads,", sum(len(x) for x in res["sam"]), "SAM lines")


if __name__ == "__main__":
    main()
