#!/usr/bin/env python
"""cohort.py — configs 4 and 5 of BASELINE.json through the shipped code path, one process per GPU (torchrun).

  --workload cohort   (config 5) S samples x C chunks sharded over the ranks (sharding.assign_chunks), every chunk counted by
                      eqp_count_chunk from host buffers, eqp_accumulate into the sample's WIDE accumulator, one eqp_reduce_sample
                      (ncclReduce inside libeqpcount.so) per sample onto its owner rank, combine.write_sample_files there, then the
                      `Id x sample` project matrices with combine.combine_counts (combineCounts.py semantics, process-samples.sh:1726-1823).
                      The last --verify samples are small ones (--verify-reads per chunk) whose sample files are checked against
                      `combineCounts.py -A` over the ORACLE's per-chunk count files.
  --workload junction (config 4) one sample of C chunks with 40 % spliced reads (seeds 1004...), sharded by chunk; the head of every chunk
                      (--verify-reads fragments) is also counted on its own and its junction rows compared with the oracle.

Chunks are generated (tools/worldgen), counted and deleted one at a time per rank; every chunk is its own generated text (no tiling).
Rank 0 prints one JSON line.  torchrun --nproc-per-node N tools/cohort.py ... ; with one GPU it runs as a single rank.
"""
import argparse
import hashlib
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
KINDS = ("transcript", "junction", "genome")


def sh(cmd):
    subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)


def generate(b, anno, d, seed, first, n, layout, spliced, procs):
    """One chunk as `procs` worldgen processes over consecutive fragment-id ranges; returns three bytes objects."""
    os.makedirs(d, exist_ok=True)
    per = (n + procs - 1) // procs
    jobs, dirs = [], []
    for w in range(procs):
        lo = first + w * per
        k = min(per, first + n - lo)
        if k <= 0:
            break
        dd = os.path.join(d, "p%02d" % w)
        os.makedirs(dd, exist_ok=True)
        dirs.append(dd)
        cmd = [b.WORLDGEN_BIN, "reads", anno, dd, str(seed), str(lo), str(k), layout] + ([str(spliced)] if spliced else [])
        jobs.append(subprocess.Popen(cmd, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL))
    for j in jobs:
        if j.wait():
            raise RuntimeError("worldgen failed")
    sfx = "pe" if layout == "pe" else "sr"
    out = []
    for kind in KINDS:
        parts = []
        for i, dd in enumerate(dirs):
            data = open(os.path.join(dd, "%s-%s.sam" % (kind, sfx)), "rb").read()
            if i and data.startswith(b"@"):                               # only the chunk's first block may carry the header (bowtie2-align.sh:462-467)
                data = b"".join(l for l in data.splitlines(True) if not l.startswith(b"@"))
            parts.append(data)
        out.append(b"".join(parts))
    shutil.rmtree(d, ignore_errors=True)
    return out


def oracle_chunk(b, anno, d, seed, first, n, layout, spliced, wait=True):
    """Generates the chunk as plain files and runs the oracle on it; returns the oracle's output dir (wait=False: (dir, process))."""
    o = os.path.join(d, "oracle")
    os.makedirs(o, exist_ok=True)
    gen = " ".join([b.WORLDGEN_BIN, "reads", anno, d, str(seed), str(first), str(n), layout] + ([str(spliced)] if spliced else []))
    run = " ".join([b.ORACLE_BIN, "run", anno, d, layout, o, "--quiet"])
    p = subprocess.Popen(["bash", "-c", gen + " && " + run], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    if not wait:
        return o, p
    if p.wait():
        raise RuntimeError("oracle run failed")
    return o


def rows(path):
    return [(l.split("\t")[0], float(l.rstrip("\n").split("\t")[1])) for l in open(path)]


def diff_rows(got, want):
    if [g[0] for g in got] != [w[0] for w in want]:
        return {"ids_equal": False, "rows": len(want), "mismatch_gt_1e-9": abs(len(got) - len(want)) + 1, "worst_rel": None}
    worst, bad = 0.0, 0
    for (_, a), (_, w) in zip(got, want):
        rel = abs(a - w) / max(abs(a), abs(w)) if a != w else 0.0
        worst, bad = max(worst, rel), bad + (rel > 1e-9)
    return {"ids_equal": True, "rows": len(want), "mismatch_gt_1e-9": int(bad), "worst_rel": worst}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cohort", choices=["cohort", "junction"])
    ap.add_argument("--samples", type=int, default=32)
    ap.add_argument("--chunks", type=int, default=4)
    ap.add_argument("--reads", type=int, default=5_000_000, help="reads per chunk")
    ap.add_argument("--full-sample-reads", type=int, default=0, help="cohort: reads per chunk of sample 1 (0 = like the others; 25000000 = one full-size sample)")
    ap.add_argument("--verify", type=int, default=2, help="cohort: number of (small) samples verified against the oracle")
    ap.add_argument("--verify-reads", type=int, default=150_000, help="reads per chunk of a verified sample / of a verified chunk head")
    ap.add_argument("--genes", type=int, default=60000)
    ap.add_argument("--layout", default="se", choices=["se", "pe"])
    ap.add_argument("--workdir", default=None)
    ap.add_argument("--out", default=None, help="directory for the sample / project count files (default: <workdir>/counts)")
    args = ap.parse_args()

    import numpy as np
    import torch
    import torch.distributed as dist
    import eqp_cluster_b200 as eqp
    b = eqp.build
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    workdir = args.workdir or os.path.join(tempfile.gettempdir(), "eqp_cohort_%d" % os.getuid())
    outdir = args.out or os.path.join(workdir, "counts")
    anno = os.path.join(workdir, "anno")
    if rank == 0:
        b.build_all()
        os.makedirs(outdir, exist_ok=True)
        if not os.path.exists(os.path.join(anno, "world.seed")):
            os.makedirs(anno, exist_ok=True)
            sh([b.WORLDGEN_BIN, "anno", anno, "76", str(args.genes)])
    if world > 1:
        dist.barrier()
    t_start = time.time()
    junction = args.workload == "junction"
    spliced = 0.4 if junction else 0
    n_samples = 1 if junction else args.samples
    names = ["s%02d" % (i + 1) for i in range(n_samples)]
    verified = set() if junction else set(names[n_samples - args.verify:])

    def chunk_reads(s):
        if s in verified:
            return args.verify_reads
        if not junction and args.full_sample_reads and s == names[0]:
            return args.full_sample_reads
        return args.reads
    work = [(s, c, float(chunk_reads(s))) for s in names for c in range(args.chunks)]
    mine = eqp.sharding.assign_chunks(work, world)[rank]
    owners = eqp.sharding.sample_owner(names, world)
    ctx = eqp.CountContext(local).load_annotation(anno)
    if world > 1:
        ids = [eqp.comm_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.comm_init(ids[0], world, rank)
    n = ctx.vector_length
    acc = {}
    procs = max(1, (os.cpu_count() or 1) // world)
    params = eqp.default_params()
    params.paired_end = 1 if args.layout == "pe" else 0
    t_gen = t_count = dev_ms = 0.0
    mapped = text_bytes = 0
    head_checks = []
    for s, c in sorted(mine):
        si = names.index(s)
        seed = (1004 + c) if junction else (2000 + (si + 1) * 10 + c)      # config 4 seeds 1004..1007, config 5 seeds 2000 + s*10 + c
        nreads = chunk_reads(s)
        first = 1 + c * nreads
        t0 = time.time()
        sams = generate(b, anno, os.path.join(workdir, "gen_r%d" % rank), seed, first, nreads, args.layout, spliced, procs)
        t_gen += time.time() - t0
        t0 = time.time()
        res = ctx.count_chunk(*sams, params=params, fetch=False)
        if s not in acc:
            acc[s] = eqp.sharding.wide_zeros(n, device=dev)
            torch.cuda.current_stream().synchronize()
        ctx.accumulate(acc[s].data_ptr())
        t_count += time.time() - t0
        dev_ms += res.device_ms
        mapped += res.number_mapped_reads
        text_bytes += sum(len(x) for x in sams)
        if junction:                                                      # the head of the chunk on its own, against the oracle
            k = args.verify_reads
            key = b"\nF%09d\t" % (first + k)
            heads = [x[:x.find(key) + 1] if x.find(key) >= 0 else x for x in sams]
            got = ctx.count_chunk(*heads, params=params)
            o = oracle_chunk(b, anno, os.path.join(workdir, "head_r%d_c%d" % (rank, c)), seed, first, k, args.layout, spliced)
            d = diff_rows(got.rows("junction"), rows(os.path.join(o, "junction.cnt")))
            d.update({"chunk": c, "reads": k, "nonzero_junction_rows": len(got.rows("junction"))})
            head_checks.append(d)
        del sams
    # ---- the one exchange: per sample, sum of the ranks' wide accumulators onto the owner (combineCounts.py -A)
    if world > 1:
        dist.barrier()
    t0 = time.time()
    zero = eqp.sharding.wide_zeros(n, device=dev) if world > 1 else None
    for s in names:                                                       # an owner that counted no chunk of its sample still receives the sum
        if owners[s] == rank and s not in acc:
            acc[s] = eqp.sharding.wide_zeros(n, device=dev)
    torch.cuda.synchronize()
    for s in names:
        if world > 1:
            buf = acc.get(s)
            if buf is None:
                zero.zero_()
                torch.cuda.current_stream().synchronize()
                buf = zero
            ctx.reduce_sample(buf.data_ptr(), root=owners[s])
    t_reduce = time.time() - t0
    # ---- writers
    t0 = time.time()
    written = {}
    for s in names:
        if owners[s] == rank:
            written[s] = eqp.combine.write_sample_files(ctx.feature_ids, acc[s].cpu().numpy(), os.path.join(outdir, s + "-bowtie2-" + ("pe" if args.layout == "pe" else "sr")))
    t_write = time.time() - t0
    # ---- verification of the small samples: combineCounts.py -A over the oracle's chunk files
    checks = []
    for s in sorted(verified):
        if owners[s] != rank:
            continue
        si = names.index(s)
        files = {m: [] for m in eqp.MODES}
        jobs = [oracle_chunk(b, anno, os.path.join(workdir, "verify_%s_c%d" % (s, c)), 2000 + (si + 1) * 10 + c, 1 + c * args.verify_reads, args.verify_reads, args.layout, 0, wait=False)
                for c in range(args.chunks)]                               # the chunks' oracle runs side by side
        for o, p in jobs:
            if p.wait():
                raise RuntimeError("oracle run failed")
            for m in eqp.MODES:
                files[m].append(os.path.join(o, m + ".cnt"))
        for m in eqp.MODES:
            want = os.path.join(workdir, "want-%s-%s.cnt" % (s, m))
            eqp.combine.combine_counts(files[m], want, add_up=True)
            d = diff_rows(rows(written[s][m]), rows(want))
            d.update({"sample": s, "mode": m})
            checks.append(d)
    gathered = [None] * world
    stats = {"rank": rank, "chunks": len(mine), "mapped": mapped, "text_bytes": text_bytes, "t_gen": t_gen, "t_count": t_count, "dev_ms": dev_ms,
             "t_reduce": t_reduce, "t_write": t_write, "checks": checks, "head_checks": head_checks, "written": written}
    if world > 1:
        dist.all_gather_object(gathered, stats)
    else:
        gathered = [stats]
    if rank == 0:
        t0 = time.time()
        sha = {}
        all_written = {}
        for g in gathered:
            all_written.update(g["written"])
        if not junction:
            sfx = "-bowtie2-" + ("pe" if args.layout == "pe" else "sr")
            for m in eqp.MODES:                                           # process-samples.sh:1796-1822: Id x sample matrix, columns in sample order
                out = os.path.join(outdir, "project%s-%s.cnt" % (sfx, m))
                eqp.combine.combine_counts([all_written[s][m] for s in names], out, file_suffix="%s-%s.cnt" % (sfx, m))
                sha[m] = hashlib.sha256(open(out, "rb").read()).hexdigest()
        else:
            for m in eqp.MODES:
                sha[m] = hashlib.sha256(open(all_written[names[0]][m], "rb").read()).hexdigest()
        t_matrix = time.time() - t0
        total_mapped = sum(g["mapped"] for g in gathered)
        count_wall = max(g["t_count"] for g in gathered)
        line = {"workload": "config 5: %d samples x %d chunks" % (n_samples, args.chunks) if not junction else "config 4: one %d-chunk sample, 40%% spliced reads, seeds 1004.." % args.chunks,
                "n_gpus": world, "reads_per_chunk": args.reads, "full_sample_reads_per_chunk": args.full_sample_reads, "verify_reads_per_chunk": args.verify_reads, "layout": args.layout, "genes": args.genes,
                "chunks": len(work), "mapped_reads": total_mapped, "sam_text_bytes": sum(g["text_bytes"] for g in gathered),
                "counting_seconds_max_rank": count_wall, "device_seconds_max_rank": max(g["dev_ms"] for g in gathered) / 1e3,
                "reads_per_s_counting": total_mapped / count_wall if count_wall else None,
                "samples_per_s_counting": n_samples / count_wall if count_wall else None,
                "generation_seconds_max_rank": max(g["t_gen"] for g in gathered), "reduce_seconds": max(g["t_reduce"] for g in gathered),
                "write_seconds_max_rank": max(g["t_write"] for g in gathered), "matrix_seconds": t_matrix, "wall_seconds": time.time() - t_start,
                "reduce": "eqp_reduce_sample (ncclReduce(sum, uint64) of the %d-element wide accumulator inside libeqpcount.so), one per sample" % (2 * n),
                "chunks_per_rank": [g["chunks"] for g in gathered],
                "sha256": sha, "verified_against_oracle": [c for g in gathered for c in g["checks"]], "chunk_heads_against_oracle": [c for g in gathered for c in g["head_checks"]],
                "timing_note": "counting_seconds = wall around eqp_count_chunk (host buffers: H2D + counting + D2H) + eqp_accumulate per chunk; chunk generation is outside it"}
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
