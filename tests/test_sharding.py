"""N>1 host logic on CPU: world_size-2 gloo. Reads shard over ranks in contiguous blocks, nothing is exchanged on the data path, rank 0
gathers the SAM text in input order. The per-rank aligner here is the CPU oracle (this is a test of the sharding and
gathering, not of the kernels); the single-process oracle output is the expectation."""
import os
import socket
import sys

import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))

from genome_weaver_align_b200 import sharding, synth  # noqa: E402


def test_shard_blocks_cover_every_read_once():
    for n, block, world in [(0, 4, 2), (1, 4, 2), (10, 3, 2), (1000, 64, 8), (7, 1, 3), (8, 100, 4)]:
        seen = []
        for r in range(world):
            seen += sharding.shard_blocks(n, block, world, r)
        seen.sort()
        assert [b for b, _, _ in seen] == list(range((n + block - 1) // block))
        assert sum(e - s for _, s, e in seen) == n
        assert all(seen[i][2] == seen[i + 1][1] for i in range(len(seen) - 1))
    with pytest.raises(ValueError):
        sharding.shard_blocks(10, 0, 2, 0)
    with pytest.raises(ValueError):
        sharding.shard_blocks(10, 4, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_main(rank, world, port, fa, reads, block, out_path):
    import oracle_lib as O
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    oi = O.OracleIndex.load(fa)       # every rank holds its own replica of the index
    names = ["r%d" % i for i in range(len(reads))]

    def align_block(s, e):
        res = oi.align(names[s:e], reads[s:e], None, k=2)
        return "".join(l + "\n" for per in res["sam"] for l in per)

    text = sharding.align_sharded(align_block, len(reads), block)
    slow = sharding.max_over_ranks(1.0 + rank)
    assert slow == float(world)
    if rank == 0:
        open(out_path, "w").write(text)
    else:
        assert text is None
    oi.close()
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_gather_sam_in_input_order(tmp_path):
    import oracle_lib as O
    g = synth.random_genome(60_000, 7)
    fa = str(tmp_path / "g.fa")
    synth.write_fasta(fa, [("chr1", g)])
    O.OracleIndex.build_files(fa)
    reads = synth.simulate_reads(g, 230, 100, 2, 8)
    oi = O.OracleIndex.load(fa)
    want = oi.align(["r%d" % i for i in range(len(reads))], reads, None, k=2)
    expect = "".join(l + "\n" for per in want["sam"] for l in per)
    oi.close()
    out = str(tmp_path / "out.sam")
    mp.spawn(_rank_main, args=(2, _free_port(), fa, reads, 32, out), nprocs=2, join=True)   # 8 blocks, ragged last one
    assert open(out).read() == expect
