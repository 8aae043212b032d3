#!/usr/bin/env python
"""Parity at the bench's annotation scale (60,000 genes, 1.17 M junction references): the CUDA path through the C ABI against
the CPU oracle on larger seeded chunks than the pytest cases use (the oracle needs ~70 s per million reads).
Usage: python tools/parity_large.py [--reads 1000000] [--layouts se,pe] [--genes 60000]   (needs a GPU; test infrastructure)
Prints one JSON line per layout: integer outputs must be equal, fractional counts within 1e-9 relative."""
import argparse
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (helpers only: annotation / sample generation)
import eqp_cluster_b200 as eqp  # noqa: E402


def read_cnt(path):
    out = {}
    for line in open(path):
        k, v = line.rstrip("\n").split("\t")
        out[k] = float(v)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=1_000_000)
    ap.add_argument("--layouts", default="se,pe")
    ap.add_argument("--genes", type=int, default=60000)
    args = ap.parse_args()
    b = eqp.build
    b.build_tools(); b.build_oracle(); b.build_library()
    workdir = os.path.join(tempfile.gettempdir(), "eqp_parity_%d" % os.getuid())
    os.makedirs(workdir, exist_ok=True)
    anno = bench.make_annotation(b, workdir, args.genes)
    ctx = eqp.CountContext(0).load_annotation(anno)
    ok = True
    for layout in args.layouts.split(","):
        seed = 1001 if layout == "se" else 1003
        d, o = bench.oracle_sample(b, anno, workdir, "parity_%s_%d" % (layout, args.reads), seed, 1, args.reads, layout)
        t0 = time.time()
        bench.run_oracle_parallel(b, anno, [(d, o)], layout)
        t_oracle = time.time() - t0
        sams = [open(bench.sam_path(d, k, layout), "rb").read() for k in bench.KINDS]
        p = eqp.default_params(); p.paired_end = 1 if layout == "pe" else 0
        res = ctx.count_chunk(*sams, params=p)
        again = ctx.count_chunk(*sams, params=p)
        stats = {}
        for line in open(os.path.join(o, "stats.txt")):
            k, _, v = line.strip().partition("=")
            stats[k] = v
        rep = {"layout": layout, "reads": args.reads, "genes": args.genes, "oracle_s": round(t_oracle, 1), "device_ms": round(res.device_ms, 2),
               "mapped": res.number_mapped_reads, "mapped_equal": res.number_mapped_reads == int(stats["NUMBER_MAPPED_READS"]),
               "expressed_equal": res.number_expressed_reads == int(stats["NUMBER_EXPRESSED_READS"]),
               "weight_objects_equal": res.num_weight_objects[0] == int(stats["NUM_WEIGHT_OBJECTS"]),
               "bit_reproducible": all((res.fixed[m] == again.fixed[m]).all() for m in eqp.MODES)}
        for m in eqp.MODES:
            want = read_cnt(os.path.join(o, m + ".cnt"))
            got = res.as_dict(m, nonzero_only=(m == "junction"))
            worst, bad = 0.0, 0
            for k in set(got) | set(want):
                a, w = got.get(k, 0.0), want.get(k, 0.0)
                if a != w:
                    r = abs(a - w) / max(abs(a), abs(w))
                    worst = max(worst, r)
                    bad += r > 1e-9
            rep[m] = {"features": len(want), "nonzero": sum(1 for v in want.values() if v), "mismatch_gt_1e-9": bad, "worst_rel": worst,
                      "sum_ours": sum(got.values()), "sum_oracle": sum(want.values())}
            ok &= bad == 0
        ok &= rep["mapped_equal"] and rep["expressed_equal"] and rep["weight_objects_equal"] and rep["bit_reproducible"]
        print(json.dumps(rep), flush=True)
    ctx.close()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
