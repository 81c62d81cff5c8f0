"""Read sharding over the GPUs of one box: reads are independent units (align/Align.java:188-195) and the output order is
the input order, so the read list is cut into contiguous blocks of `block` reads, block b goes to rank b % world, every rank holds a
full replica of the index, and nothing is exchanged on the data path. The only communication is the host-side gather of SAM text
(rank 0 writes the blocks back in input order) and the max-over-ranks reduction of the timed region.

Works with any torch.distributed backend (nccl on the GPU box, gloo in the CPU tests); `align_block` is the per-rank aligner call
(BidirectionalSuffixFilter.align(...).sam(...) in production)."""
import torch
import torch.distributed as dist


def shard_blocks(n_reads, block, world, rank):
    """[(block_index, start, end)] of the blocks this rank aligns"""
    if block <= 0 or world <= 0 or not (0 <= rank < world):
        raise ValueError("bad sharding arguments")
    n_blocks = (n_reads + block - 1) // block
    return [(b, b * block, min(n_reads, (b + 1) * block)) for b in range(rank, n_blocks, world)]


def align_sharded(align_block, n_reads, block, group=None):
    """runs align_block(start, end) -> SAM text on this rank's blocks; rank 0 returns the whole SAM body in input order, others None"""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    mine = [(b, align_block(s, e)) for b, s, e in shard_blocks(n_reads, block, world, rank)]
    if world == 1:
        return "".join(t for _, t in sorted(mine))
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(mine, gathered, dst=0, group=group)
    if rank != 0:
        return None
    parts = sorted(p for per_rank in gathered for p in per_rank)
    assert [b for b, _ in parts] == list(range(len(parts))), "a block is missing or duplicated"
    return "".join(t for _, t in parts)


def max_over_ranks(seconds, device=None, group=None):
    """the job's time for a timed region = the slowest rank's (bench.py contract)"""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return float(seconds)
    t = torch.tensor([seconds], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())
