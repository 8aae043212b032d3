"""World-size-2 gloo tests (CPU) of the multi-GPU host logic: chunk assignment and the exact sample-level reduce."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import eqp_cluster_b200 as eqp
from conftest import ROOT


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import eqp_cluster_b200 as eqp_
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    work = [("s%d" % s, c, 1.0 + 0.1 * c) for s in range(3) for c in range(4)]            # 3 samples x 4 chunks
    mine = eqp_.sharding.assign_chunks(work, world)[rank]
    owners = eqp_.sharding.sample_owner({w[0] for w in work}, world)
    g = torch.Generator().manual_seed(1234)
    chunk_vec = {(s, c): torch.randint(0, 1 << 40, (1000,), generator=g, dtype=torch.int64) for s, c, _ in work}   # same on every rank
    per_sample = {s: torch.zeros(1000, dtype=torch.int64) for s in owners}
    for s, c in mine:
        per_sample[s] += chunk_vec[(s, c)]
    got = eqp_.sharding.reduce_vectors(per_sample, owners)
    ok = True
    for s, t in got.items():
        want = sum(chunk_vec[(s, c)] for c in range(4))
        ok &= bool(torch.equal(t, want))
    q.put((rank, sorted(got), ok, len(mine)))
    dist.destroy_process_group()


def test_two_rank_reduce_is_exact():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(r[2] for r in res)
    assert sorted(res[0][1] + res[1][1]) == ["s0", "s1", "s2"]           # every sample has exactly one owner
    assert res[0][3] + res[1][3] == 12 and abs(res[0][3] - res[1][3]) <= 1


def test_assignment_is_balanced_and_complete():
    work = [("s%02d" % s, c, 25e6 if c < 3 else 7e6) for s in range(32) for c in range(4)]    # config 5: 32 samples x 4 chunks
    ranks = eqp.sharding.assign_chunks(work, 8)
    flat = [x for r in ranks for x in r]
    assert sorted(flat) == sorted((s, c) for s, c, _ in work)
    cost = {(s, c): k for s, c, k in work}
    loads = [sum(cost[x] for x in r) for r in ranks]
    assert max(loads) - min(loads) <= 25e6
