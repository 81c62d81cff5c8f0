"""Committed golden fixtures (tests/golden/*.json, written by tools/make_golden.py from the CPU oracle on seeded workloads):
the oracle must still reproduce them (CPU test) and the CUDA path, through the C-ABI, must produce the same SAM lines, statuses
and debug counters (GPU test)."""
import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import make_golden as MG  # noqa: E402
import oracle_lib as O  # noqa: E402

CASES = sorted(MG.CASES)


def load(case):
    return json.load(open(os.path.join(ROOT, "tests", "golden", case + ".json")))


@pytest.mark.parametrize("case", CASES)
def test_oracle_reproduces_golden_fixture(case):
    gold = load(case)
    g = MG.genome_of(case)
    reads = MG.reads_of(case, g)
    oidx = O.OracleIndex.from_sequence("chr1", g.tobytes().decode())
    names = ["r%d" % i for i in range(len(reads))]
    res = oidx.align(names, reads, ["I" * len(r) for r in reads], **MG.CASES[case][7])
    assert res["sam"] == gold["sam"]
    assert res["status"] == gold["status"]
    assert res["counters"] == gold["counters"]


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_gpu_reproduces_golden_fixture(case):
    import genome_weaver_align_b200 as gw
    gold = load(case)
    g = MG.genome_of(case)
    reads = MG.reads_of(case, g)
    cfg = MG.CASES[case][7]
    idx = gw.FMIndexOnGenome.buildFromSequence("chr1", g.tobytes().decode())
    config = gw.AlignmentConfig(k=cfg["k"], reportType=cfg.get("report_type", 0))
    bsf = gw.BidirectionalSuffixFilter(idx, config)
    batch = bsf.align(reads)
    names = ["r%d" % i for i in range(len(reads))]
    got = batch.sam_per_read(names, ["I" * len(r) for r in reads])
    assert [list(x) for x in got] == gold["sam"]
    assert [int(x) for x in batch.status()] == gold["status"]
    ok = [i for i in range(len(reads)) if gold["status"][i] != 2]
    if len(ok) == len(reads):
        assert batch.counters["fm_searches"] == sum(c[0] for c in gold["counters"])
        assert batch.counters["sw_calls"] == sum(c[1] for c in gold["counters"])
    bsf.close()
    idx.close()
