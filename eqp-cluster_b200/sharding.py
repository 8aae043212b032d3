"""Multi-GPU plumbing: chunk -> rank assignment and the sample-level sum of count vectors.

Chunks are independent (fragments never span chunks, ChangeFastqIdsAndSplit.java:229-249) and chunk results combine by
addition (combineCounts.py:199-210), so the only exchange on the path is ONE reduce(sum) of the per-rank vectors.  The vectors
are 2^-38 fixed-point integers: the sum is exact, so neither the rank count nor the reduction tree changes a bit.
torch.distributed (NCCL on GPUs, gloo in the CPU tests) is the transport.
"""
import torch
import torch.distributed as dist


def assign_chunks(work, world_size):
    """work: list of (sample, chunk, cost).  Longest-first greedy assignment; returns one list of (sample, chunk) per rank."""
    ranks = [[] for _ in range(world_size)]
    load = [0.0] * world_size
    for sample, chunk, cost in sorted(work, key=lambda w: (-w[2], w[0], w[1])):
        r = min(range(world_size), key=lambda k: (load[k], k))
        ranks[r].append((sample, chunk))
        load[r] += cost
    return ranks


def sample_owner(samples, world_size):
    """Rank that writes a sample's combined vector (round robin over the sorted sample names)."""
    return {s: i % world_size for i, s in enumerate(sorted(samples))}


def reduce_vectors(per_sample, owners, group=None):
    """per_sample: {sample: int64 tensor (this rank's sum over its chunks of that sample, zeros if it holds none)}.
    One reduce(sum) per sample to its owner; returns {sample: tensor} for the samples this rank owns."""
    rank = dist.get_rank(group)
    out = {}
    for s in sorted(per_sample):
        t = per_sample[s]
        assert t.dtype == torch.int64
        dist.reduce(t, dst=owners[s], op=dist.ReduceOp.SUM, group=group)
        if owners[s] == rank:
            out[s] = t
    return out


def fixed_to_float(t, shift=38):
    return t.to(torch.float64) / float(1 << shift)
