#!/usr/bin/env python
"""inflate_bench.py — development tool: one generated chunk counted from its text in host memory (eqp_count_chunk) and from its three
.sam.gz members (eqp_count_chunk_gz, device inflate), for several segment sizes; checks that both give the same vectors.

  python tools/inflate_bench.py [--reads N] [--level 6] [--segs 16384,32768,65536] [--layout se]
"""
import argparse
import json
import os
import shutil
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=4_000_000)
    ap.add_argument("--level", type=int, default=6)
    ap.add_argument("--segs", default="16384,32768,65536")
    ap.add_argument("--layout", default="se")
    ap.add_argument("--genes", type=int, default=60000)
    ap.add_argument("--repeat", type=int, default=2)
    ap.add_argument("--libs", default="", help="comma-separated variant names under eqp-cluster_b200/lib/variants (tools/variants.py build); default: the product library")
    args = ap.parse_args()
    import numpy as np
    import torch
    import eqp_cluster_b200 as eqp
    import bench
    b = eqp.build
    b.build_all()
    work = os.path.join(tempfile.gettempdir(), "eqp_inflate_bench")
    os.makedirs(work, exist_ok=True)
    anno = bench.make_annotation(b, work, args.genes)
    dirs = bench.generate_blocks(b, anno, os.path.join(work, "chunk"), 1001, 1, args.reads, args.layout, os.cpu_count() or 1)
    t0 = time.time()
    gz = []
    for kind in bench.KINDS:
        out = os.path.join(work, kind + ".sam.gz")
        subprocess.run([b.PGZIP_BIN, out, str(args.level), str(os.cpu_count() or 1)] + [bench.sam_path(d, kind, args.layout) for d in dirs], check=True)
        n = os.path.getsize(out)
        buf = torch.empty(n, dtype=torch.uint8, pin_memory=True)
        with open(out, "rb") as f:
            assert f.readinto(memoryview(buf.numpy())) == n
        gz.append(buf)
        os.remove(out)
    t_gz = time.time() - t0
    host = bench.load_blocks_pinned(torch, dirs, args.layout)
    shutil.rmtree(os.path.join(work, "chunk"), ignore_errors=True)
    libs = [("product", None)] + [(n, os.path.join(ROOT, "eqp-cluster_b200", "lib", "variants", "libeqpcount_%s.so" % n)) for n in args.libs.split(",") if n]
    ctx = eqp.CountContext(0).load_annotation(anno)
    targs = [(h.data_ptr(), h.numel()) for h in host]
    gargs = [(g.data_ptr(), g.numel()) for g in gz]
    ref = None
    ms = []
    for _ in range(args.repeat + 1):
        t0 = time.perf_counter()
        ref = ctx.count_chunk(*targs)
        ms.append((time.perf_counter() - t0) * 1e3)
    print(json.dumps({"leg": "text in pinned host memory (eqp_count_chunk, no prefetch)", "reads": args.reads, "text_bytes": sum(h.numel() for h in host),
                      "gz_bytes": sum(g.numel() for g in gz), "gzip_level": args.level, "pgzip_seconds": round(t_gz, 1), "wall_ms": [round(x, 1) for x in ms], "device_ms": ref.device_ms}), flush=True)
    for lib_name, lib_path in libs:
      if lib_path:
        ctx.close()
        ctx = eqp.CountContext(0, lib_path).load_annotation(anno)
      for seg in [s for s in args.segs.split(",") if s]:
        os.environ["EQP_INFLATE_SEG"] = seg
        ms, infos, same = [], None, None
        for _ in range(args.repeat + 1):
            t0 = time.perf_counter()
            try:
                c = ctx.count_chunk_gz(*gargs)
            except eqp.EqpError as e:
                print(json.dumps({"leg": "gz", "lib": lib_name, "seg": int(seg), "error": str(e)}), flush=True)
                break
            ms.append((time.perf_counter() - t0) * 1e3)
            infos = [ctx.last_inflate(f) for f in range(3)]
            same = all(np.array_equal(c.fixed[m], ref.fixed[m]) for m in eqp.MODES)
        else:
            tot = {k: round(sum(i[k] for i in infos), 1) for k in ("search_ms", "decode_ms", "resolve_ms", "emit_ms")}
            print(json.dumps({"leg": "three .sam.gz members in pinned host memory (eqp_count_chunk_gz)", "lib": lib_name, "seg": int(seg), "wall_ms": [round(x, 1) for x in ms], "count_device_ms": c.device_ms,
                              "same_vectors_as_text": same, "inflate_ms": tot, "segments": [i["segments"] for i in infos], "chain": [i["chain"] for i in infos],
                              "bogus_starts": sum(i["bogus_starts"] for i in infos), "waves": [i["waves"] for i in infos]}), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
