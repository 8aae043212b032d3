#!/usr/bin/env python
"""variants.py — A/B timing of several builds of libeqpcount.so on ONE generated chunk (development tool).

  python tools/variants.py build  name=-DFLAG1,-DFLAG2 ...     compile eqp-cluster_b200/lib/variants/libeqpcount_<name>.so (here, no GPU needed)
  python tools/variants.py run [--reads N] [--layout se|pe]     on the GPU box: generate one chunk, then for every variant (and the product
                                                                library) count it 1 + 3 times from device-resident text and print the stage times
"""
import glob
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
VDIR = os.path.join(ROOT, "eqp-cluster_b200", "lib", "variants")


def main():
    import eqp_cluster_b200 as eqp
    b = eqp.build
    if sys.argv[1] == "build":
        os.makedirs(VDIR, exist_ok=True)
        for spec in sys.argv[2:]:
            name, _, flags = spec.partition("=")
            out = os.path.join(VDIR, "libeqpcount_%s.so" % name)
            b._run([b.nvcc_path()] + b.NVCC_FLAGS + [f for f in flags.split(",") if f] + ["-o", out, os.path.join(b.CSRC, "eqp_capi.cu"), "-lz", "-lpthread"])
            print("built", out)
        return
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("cmd")
    ap.add_argument("--reads", type=int, default=25_000_000)
    ap.add_argument("--layout", default="se")
    ap.add_argument("--genes", type=int, default=60000)
    ap.add_argument("--knobs", default="", help="comma-separated worldgen annotation knobs, e.g. ld=0,tr=0,inv=0,ver=0,ver2=0,scaf=0 (the round-1 world)")
    ap.add_argument("--only", default="", help="comma-separated variant names to run (default: all)")
    args = ap.parse_args()
    import torch
    sys.path.insert(0, ROOT)
    import bench
    work = os.path.join(tempfile.gettempdir(), "eqp_variants")
    os.makedirs(work, exist_ok=True)
    b.build_tools()
    if args.knobs:
        work = os.path.join(work, "knobs_" + args.knobs.replace(",", "_").replace("=", ""))
        os.makedirs(os.path.join(work, "anno"), exist_ok=True)
        if not os.path.exists(os.path.join(work, "anno", "world.seed")):
            subprocess.run([b.WORLDGEN_BIN, "anno", os.path.join(work, "anno"), "76", str(args.genes)] + args.knobs.split(","), check=True, stdout=subprocess.DEVNULL)
    anno = bench.make_annotation(b, work, args.genes)
    dirs = bench.generate_blocks(b, anno, os.path.join(work, "chunk"), 1001, 1, args.reads, args.layout, os.cpu_count() or 1)
    host = bench.load_blocks_pinned(torch, dirs, args.layout)
    import shutil
    shutil.rmtree(os.path.join(work, "chunk"), ignore_errors=True)       # 45 GB of text per chunk: two worlds on one disk do not fit
    dev = [h.cuda() for h in host]
    torch.cuda.synchronize()
    dargs = [(d.data_ptr(), d.numel()) for d in dev]
    libs = [("product", b.LIB_PATH)] + [(os.path.basename(p)[12:-3], p) for p in sorted(glob.glob(os.path.join(VDIR, "libeqpcount_*.so")))]
    runs = [(n, p) for n, p in libs if not args.only or n in args.only.split(",")]
    for name, path in runs:
        r, w = os.pipe()
        saved = os.dup(2)
        ctx = eqp.CountContext(0, path).load_annotation(anno)
        try:
            ctx.count_chunk(*dargs, on_device=True, fetch=False)
        except eqp.EqpError as e:
            print(json.dumps({"variant": name, "error": str(e)}), flush=True)
            os.close(r); os.close(w); os.close(saved)
            ctx.close()
            continue
        os.environ["EQP_STAGE_TIMING"] = "1"
        ms = []
        os.dup2(w, 2)
        try:
            for _ in range(3):
                ms.append(ctx.count_chunk(*dargs, on_device=True, fetch=False).device_ms)
        finally:
            os.dup2(saved, 2)
            os.close(w)
            os.environ.pop("EQP_STAGE_TIMING", None)
        text = os.read(r, 1 << 20).decode(errors="replace")
        os.close(r)
        import re
        runs, cur = [], {}
        for line in text.splitlines():
            m = re.match(r"\[eqp stage\] (.+?)\s+([\d.]+) ms", line)
            if not m:
                continue
            if m.group(1) == "total":
                runs.append(cur)
                cur = {}
            else:
                cur[m.group(1)] = cur.get(m.group(1), 0.0) + float(m.group(2))
        best = {k: round(min(r[k] for r in runs if k in r), 2) for k in (runs[0] if runs else {})}
        print(json.dumps({"variant": name, "device_ms": [round(x, 2) for x in ms], "stages_min_ms": best}), flush=True)
        ctx.close()


if __name__ == "__main__":
    main()
