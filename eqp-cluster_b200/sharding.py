"""Multi-GPU plumbing: chunk -> rank assignment and the sample-level sum of count vectors.

Chunks are independent (fragments never span chunks, ChangeFastqIdsAndSplit.java:229-249) and chunk results combine by
addition (combineCounts.py:199-210), so the only exchange on the path is ONE reduce(sum) of the per-rank vectors.  The vectors
are WIDE fixed-point accumulators (include/eqp_count.h "sample level"): 2n integers, fraction limbs in 2^-38 reads followed by
integer limbs.  Integer sums are exact, so neither the rank count nor the reduction tree changes a bit, and because a limb grows
by less than 2^38 per chunk neither limb can overflow a (signed or unsigned) 64-bit word before 2^25 chunks have been added: a
highly expressed feature of a 100-200 M read sample is safe, which a single 2^-38 word is not (it wraps at 2^26 reads).
Transport: eqp_reduce_sample (NCCL inside libeqpcount.so; CountContext.reduce_sample) on the GPUs, or torch.distributed
(gloo in the CPU tests, NCCL) through reduce_vectors below.
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


def wide_zeros(n, device=None):
    """Zeroed wide accumulator for a vector of n features."""
    return torch.zeros(2 * n, dtype=torch.int64, device=device)


def wide_add_chunk(acc, chunk_fixed, shift=38):
    """Host/torch restatement of eqp_accumulate for the CPU tests: chunk_fixed = a chunk's n fixed-point values as int64 bit patterns."""
    n = acc.numel() // 2
    mask = (1 << shift) - 1
    acc[:n] += chunk_fixed & mask
    acc[n:] += (chunk_fixed >> shift) & ((1 << (64 - shift)) - 1)          # logical shift of the unsigned pattern
    return acc


def wide_to_float(t, shift=38):
    n = t.numel() // 2
    return t[n:].to(torch.float64) + t[:n].to(torch.float64) / float(1 << shift)
