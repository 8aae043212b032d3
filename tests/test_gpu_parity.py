"""GPU parity tests (-m gpu): the CUDA path, called through the C ABI, against the oracle on the same seeded inputs."""
import os

import numpy as np
import pytest

import eqp_cluster_b200 as eqp
from conftest import GOLDEN, SHAPES, assert_counts_match, assert_rows_match, read_cnt, read_rows, read_sams, read_stats, read_weights

pytestmark = pytest.mark.gpu

FRAGSTAT = np.dtype([("nPe", "<u2"), ("nSr1", "<u2"), ("nSr2", "<u2"), ("flags", "<u2"), ("edPe", "<i4"), ("edSr1", "<i4"), ("edSr2", "<i4")])


@pytest.fixture(scope="module")
def ctx(world):
    c = eqp.CountContext(0).load_annotation(world.anno)
    yield c
    c.close()


def check_chunk(ctx, chunk, oracle_dir, params=None, exact=False):
    res = ctx.count_chunk(*read_sams(chunk["dir"], chunk["layout"]), params=params)
    st = read_stats(os.path.join(oracle_dir, "stats.txt"))
    assert res.number_mapped_reads == int(st["NUMBER_MAPPED_READS"])
    assert res.number_expressed_reads == int(st["NUMBER_EXPRESSED_READS"])
    assert res.num_weight_objects[0] == int(st["NUM_WEIGHT_OBJECTS"])
    assert res.kernel_launches > 0 and res.device_ms > 0
    for i, m in enumerate(eqp.MODES):
        # rows in print order: every id of the map for gene / exon (ComputeCounts.java:877-897), the non-zero rows of the filtered junction map
        assert_rows_match(res.rows(m), read_rows(os.path.join(oracle_dir, m + ".cnt")), exact=exact, what=m)
        assert abs(res.sum_weights[i] - float(st["SUM_WEIGHTS_" + m.upper()])) <= 0.01, m    # ComputeCounts.java:989
    return res


def check_weights(ctx, oracle_dir):
    """a20 of the weight pass and the final.wgt entry of every fragment in the intersection, against bed.wgt / final.wgt."""
    base = int(ctx.debug_fetch("meta", np.uint32)[0])
    a20, nfinal = ctx.debug_fetch("a20", np.uint32), ctx.debug_fetch("weights:final", np.uint32)
    got = {"F%09d" % (base + s): (int(a20[s]), int(nfinal[s])) for s in np.nonzero(a20)[0]}
    assert got == read_weights(oracle_dir)


@pytest.mark.parametrize("name,seed,layout,spliced,n", [
    ("se_small", 1001, "se", None, 3000), ("pe_small", 1003, "pe", None, 3000), ("se_spliced", 1004, "se", 0.4, 3000),
    ("se_mid", 1011, "se", None, 40000), ("pe_mid", 1013, "pe", None, 40000)])
def test_counts_match_oracle(world, ctx, name, seed, layout, spliced, n):
    c = world.chunk(name, seed, 1, n, layout, spliced)
    check_chunk(ctx, c, c["oracle"])


@pytest.mark.parametrize("layout,seed", [("se", 1001), ("pe", 1003)])
@pytest.mark.parametrize("shape", sorted(SHAPES))
def test_annotation_shapes(shape_worlds, shape, layout, seed):
    """One dense world per shape of SURVEY 8d (length differences / a17, tr- exons, long tr- fillers, .inv-strand, two versions,
    3-column junction map) and their mix: counts, statistics and per-fragment weights against the oracle."""
    w = shape_worlds(shape)
    c = w.chunk(layout, seed, 1, 6000, layout)
    with eqp.CountContext(0).load_annotation(w.anno) as cx:
        check_chunk(cx, c, c["oracle"])
        check_weights(cx, c["oracle"])
        assert cx.map_columns["junction"] == (3 if shape == "map3" else 2)


@pytest.mark.parametrize("direction,code", [("forward", 1), ("backward", 2)])
@pytest.mark.parametrize("layout,seed", [("se", 1001), ("pe", 1003)])
def test_strand_specific_mode(shape_worlds, layout, seed, direction, code):
    w = shape_worlds("mixed")
    c = w.chunk(layout, seed, 1, 6000, layout)
    o = w.oracle_run(layout, ("--stranded=" + direction,))
    p = eqp.default_params()
    p.stranded = code
    with eqp.CountContext(0).load_annotation(w.anno) as cx:
        check_chunk(cx, c, o, params=p)
        check_weights(cx, o)


@pytest.mark.parametrize("layout,seed", [("se", 1201), ("pe", 1203)])
def test_genomic_aligner_input(shape_worlds, layout, seed):
    """the SAM of a spliced genome aligner (N CIGARs, HI:i hit indices) counted with eqp_params.genomic_input
    against the oracle's restatement of compute-genomic-counts.sh (gene and exon runs)."""
    w = shape_worlds("mixed")
    c = w.genomic_chunk("genomic_" + layout, seed, 1, 6000, layout)
    p = eqp.default_params()
    p.genomic_input, p.modes = 1, 3
    with eqp.CountContext(0).load_annotation(w.anno) as cx:
        res = cx.count_chunk(open(c["sam"], "rb").read(), b"", b"", params=p)
        st = read_stats(os.path.join(c["oracle"], "stats.txt"))
        assert res.number_mapped_reads == int(st["NUMBER_MAPPED_READS"]) and res.number_expressed_reads == int(st["NUMBER_EXPRESSED_READS"])
        assert res.num_weight_objects[0] == int(st["NUM_WEIGHT_OBJECTS"])
        for m in ("gene", "exon"):
            assert_rows_match(res.rows(m), read_rows(os.path.join(c["oracle"], m + ".cnt")), what="genomic " + m)
        p.modes = 7
        with pytest.raises(eqp.EqpError) as e:
            cx.count_chunk(open(c["sam"], "rb").read(), b"", b"", params=p)
        assert e.value.code == -5                                           # the junction run of that script is not built
        with pytest.raises(eqp.EqpError):                                   # an HI tag outside the genomic mode is still refused
            cx.count_chunk(open(c["sam"], "rb").read(), b"", b"")


def test_weight_intermediates_match_oracle(world, ctx):
    """*-sam.wgt (a4), trans-junction weights (a6 -e) and final.wgt (a6/a20) against the oracle's files."""
    c = world.chunk("pe_small", 1003, 1, 3000, "pe")
    ctx.count_chunk(*read_sams(c["dir"], "pe"))
    meta = ctx.debug_fetch("meta", np.uint32)
    base = int(meta[0])
    for tap, fname in (("weights:T", "transcript-sam.wgt"), ("weights:J", "junction-sam.wgt"), ("weights:G", "genome-sam.wgt"), ("weights:TJ", "trans-junction-sam.wgt")):
        st = ctx.debug_fetch(tap, FRAGSTAT)
        got = {}
        for s in np.nonzero(st["flags"] & 1)[0]:
            e = st[s]
            cls = "paired-end" if e["flags"] & 2 else ("single-read (both)" if ((e["flags"] >> 2) & 3) > 2 else "single-read")
            got["F%09d" % (base + s)] = (int(e["nPe"]) + max(int(e["nSr1"]), int(e["nSr2"])), int(e["edPe"]) + int(e["edSr1"]) + int(e["edSr2"]), cls)
        want = {}
        for line in open(os.path.join(c["oracle"], fname)):
            t = line.rstrip("\n").split("\t")
            want[t[0]] = (int(t[1]), int(t[2]), t[3])
        assert got == want, fname
    a20 = ctx.debug_fetch("a20", np.uint32)
    nfinal = ctx.debug_fetch("weights:final", np.uint32)
    bedw = {l.split("\t")[0]: int(l.split("\t")[1]) for l in open(os.path.join(c["oracle"], "bed.wgt"))}
    finalw = {l.split("\t")[0]: int(l.split("\t")[1]) for l in open(os.path.join(c["oracle"], "final.wgt"))}
    for s in np.nonzero(a20)[0]:
        f = "F%09d" % (base + s)
        assert bedw[f] == a20[s] and finalw[f] == nfinal[s], f
    assert len(bedw) == int(np.count_nonzero(a20))


@pytest.mark.parametrize("flags,field", [(("--unweighted",), "unweighted"), (("--unambig",), "unambiguous"), (("--containment", "--nonsplice"), None)])
def test_switches(world, ctx, flags, field):
    c = world.chunk("pe_small", 1003, 1, 3000, "pe")
    o = world.oracle_run("pe_small", flags)
    p = eqp.default_params()
    if field:
        setattr(p, field, 1)
    else:
        p.containment, p.nonsplice_conforming = 1, 1
    check_chunk(ctx, c, o, params=p, exact=("--unweighted" in flags))      # integer counts: bit-exact


@pytest.mark.parametrize("case", ["appB", "appC"])
def test_golden_hand_traced(case):
    g = os.path.join(GOLDEN, case)
    with eqp.CountContext(0).load_annotation(os.path.join(g, "anno")) as c:
        res = c.count_chunk(*read_sams(os.path.join(g, "chunk"), "se"))
        for m in eqp.MODES:
            assert res.as_dict(m, nonzero_only=(m == "junction")) == read_cnt(os.path.join(g, "expected", m + ".cnt"))


def test_edge_cases(world, ctx):
    """Empty inputs, header-only inputs, no trailing newline, CRLF, and a repeated call on the same context."""
    c = world.chunk("se_small", 1001, 1, 3000, "se")
    t, j, g = read_sams(c["dir"], "se")
    res = ctx.count_chunk(b"", b"", b"")
    assert res.number_mapped_reads == 0 and all(float(res.counts[m].sum()) == 0.0 for m in eqp.MODES)
    res = ctx.count_chunk(b"@HD\tVN:1.0\n@SQ\tSN:chr1\tLN:1000\n", b"", b"@PG\tID:bowtie2")
    assert res.number_mapped_reads == 0
    ref = ctx.count_chunk(t, j, g)
    hdr = b"@HD\tVN:1.0\tSO:unsorted\n"
    again = ctx.count_chunk(hdr + t.rstrip(b"\n"), j.replace(b"\n", b"\r\n"), g)
    for m in eqp.MODES:
        assert np.array_equal(ref.fixed[m], again.fixed[m]), m
    assert ref.number_mapped_reads == again.number_mapped_reads


def test_input_resident_on_device(world, ctx):
    import torch
    c = world.chunk("se_small", 1001, 1, 3000, "se")
    sams = read_sams(c["dir"], "se")
    ref = ctx.count_chunk(*sams)
    dev = [torch.frombuffer(bytearray(s), dtype=torch.uint8).cuda() for s in sams]
    torch.cuda.synchronize()
    res = ctx.count_chunk(*[(d.data_ptr(), d.numel()) for d in dev], on_device=True)
    for m in eqp.MODES:
        assert np.array_equal(ref.fixed[m], res.fixed[m])


def test_errors_are_reported_not_swallowed(world, ctx):
    c = world.chunk("se_small", 1001, 1, 3000, "se")
    t, j, g = read_sams(c["dir"], "se")
    lines = t.split(b"\n")
    bad = b"\n".join(lines[50:60] + lines[:50])                        # out of order (CombineSamFiles.java:191-193)
    with pytest.raises(eqp.EqpError) as e:
        ctx.count_chunk(bad + b"\n", j, g)
    assert e.value.code == -4
    with pytest.raises(eqp.EqpError):
        ctx.count_chunk(t.replace(b"F000000007", b"Q000000007"), j, g)  # SamReader.java:157
    p = eqp.default_params()
    p.stranded = 3
    with pytest.raises(eqp.EqpError) as e:
        ctx.count_chunk(t, j, g, params=p)
    assert e.value.code == -1


def test_chunk_vectors_add_up(world, ctx):
    """Size-independent property: counting two halves of a chunk and adding the fixed-point vectors equals counting the whole
    chunk (chunks combine by addition, combineCounts.py:201-210); eqp_accumulate does the add on the device."""
    import torch
    whole = world.chunk("se_mid", 1011, 1, 40000, "se")
    sams = read_sams(whole["dir"], "se")
    res = ctx.count_chunk(*sams)
    # gene and exon vectors only: the junction vector also depends on the per-chunk map filter (a27), which is not additive
    want = np.concatenate([res.fixed[m] for m in ("gene", "exon")])

    def split(s, frag):
        k = s.find(b"\nF%09d" % frag)
        return (s, b"") if k < 0 else (s[:k + 1], s[k + 1:])
    halves = [split(s, 20001) for s in sams]
    n = ctx.vector_length
    acc = eqp.sharding.wide_zeros(n, device="cuda")
    torch.cuda.synchronize()
    for h in (0, 1):
        ctx.count_chunk(*[x[h] for x in halves])
        assert ctx.device_vector()[1] == n
        ctx.accumulate(acc.data_ptr())
    wide = acc.cpu().numpy().view(np.uint64)
    got = (wide[n:] << np.uint64(38)) + wide[:n]                         # limbs back to one 2^-38 word (no overflow at this size)
    assert n == want.size + res.fixed["junction"].size and np.array_equal(got[:want.size], want)
    assert np.array_equal(eqp.wide_to_double(wide)[:want.size], want.astype(np.float64) / float(1 << 38))
    # the reduce with NCCL called inside the library: a one-rank communicator on this GPU (the multi-rank run is tools/cohort.py)
    ctx.comm_init(eqp.comm_id(), 1, 0)
    ctx.reduce_sample(acc.data_ptr(), root=0)
    assert np.array_equal(acc.cpu().numpy().view(np.uint64), wide)


def test_wide_accumulator_does_not_wrap(world, ctx):
    """A feature that collects more than 2^26 reads' worth of weight over the chunks of a sample wraps a single 2^-38 word; the wide
    accumulator (fraction limb + integer limb) holds it."""
    import torch
    c = world.chunk("se_small", 1001, 1, 3000, "se")
    res = ctx.count_chunk(*read_sams(c["dir"], "se"))
    n = ctx.vector_length
    acc = eqp.sharding.wide_zeros(n, device="cuda")
    k = int(np.argmax(res.fixed["gene"]))
    top = float(res.counts["gene"][k])
    acc[n + k] = (1 << 27)                                                # as if earlier chunks had put 134 M reads on the top gene
    torch.cuda.synchronize()
    ctx.accumulate(acc.data_ptr())
    assert eqp.wide_to_double(acc.cpu().numpy())[k] == float(1 << 27) + top


def test_sample_files_from_the_reduced_vector(world, ctx, tmp_path):
    """Writer step after the reduce (combine.write_sample_files) against `combineCounts.py -A` semantics applied to per-chunk
    `.cnt` files (combine.combine_counts): same ids in the same order, values equal to 1e-9 (fp64 sum of chunk values vs exact
    fixed-point sum)."""
    import torch
    cmb = eqp.combine
    acc, files = None, {m: [] for m in eqp.MODES}
    for c, (seed, first) in enumerate(((1021, 1), (1022, 20001))):
        ch = world.chunk("sf_%d" % c, seed, first, 20000, "se")
        res = ctx.count_chunk(*read_sams(ch["dir"], "se"))
        if acc is None:
            acc = eqp.sharding.wide_zeros(ctx.vector_length, device="cuda")
            torch.cuda.synchronize()
        ctx.accumulate(acc.data_ptr())
        for m in eqp.MODES:                                              # chunk files as compute-counts writes them ("0" rows, junctions non-zero only)
            p = os.path.join(str(tmp_path), "s-C%03d-%s.cnt" % (c + 1, m))
            with open(p, "w") as f:
                for i, v in zip(ctx.feature_ids[m], res.counts[m]):
                    if v != 0 or m != "junction":
                        f.write("%s\t%s\n" % (i, repr(float(v)) if v else "0"))
            files[m].append(p)
    paths = cmb.write_sample_files(ctx.feature_ids, acc.cpu().numpy(), os.path.join(str(tmp_path), "s"))
    for m in eqp.MODES:
        want = os.path.join(str(tmp_path), "want-%s.cnt" % m)
        cmb.combine_counts(files[m], want, add_up=True)
        w, g = open(want).read().splitlines(), open(paths[m]).read().splitlines()
        assert [r.split("\t")[0] for r in w] == [r.split("\t")[0] for r in g], m
        assert any(float(r.split("\t")[1]) > 0 for r in g), m
        for rw, rg in zip(w, g):
            a, b = float(rw.split("\t")[1]), float(rg.split("\t")[1])
            assert abs(a - b) <= 1e-9 * max(a, b), (m, rw, rg)


def test_prefetched_chunks_give_the_same_counts(world, ctx):
    """eqp_prefetch_chunk (double-buffered H2D): same bits as the blocking copy, in order; a non-matching announcement is dropped;
    a third announcement is refused."""
    a = read_sams(world.chunk("pf_a", 1031, 1, 20000, "se")["dir"], "se")
    b = read_sams(world.chunk("pf_b", 1032, 20001, 15000, "se")["dir"], "se")
    ra, rb = ctx.count_chunk(*a), ctx.count_chunk(*b)
    keep = {m: (ra.fixed[m].copy(), rb.fixed[m].copy()) for m in eqp.MODES}
    ctx.prefetch_chunk(*a)
    ctx.prefetch_chunk(*b)
    with pytest.raises(eqp.EqpError):
        ctx.prefetch_chunk(*a)
    pa = ctx.count_chunk(*a)
    fa = {m: pa.fixed[m].copy() for m in eqp.MODES}
    ctx.prefetch_chunk(*a)                                               # b is next in the queue, then a
    pb = ctx.count_chunk(*b)
    fb = {m: pb.fixed[m].copy() for m in eqp.MODES}
    pb2 = ctx.count_chunk(*b)                                            # announced a, counted b: announcement dropped
    for m in eqp.MODES:
        assert np.array_equal(fa[m], keep[m][0]) and np.array_equal(fb[m], keep[m][1]) and np.array_equal(pb2.fixed[m], keep[m][1]), m
    assert pa.number_mapped_reads == ra.number_mapped_reads and pb2.number_mapped_reads == rb.number_mapped_reads


def test_streamed_chunk_gives_the_same_counts(world, ctx):
    """eqp_chunk_begin / feed / end: the three texts fed piecewise and interleaved (as three inflate threads would) count to the same bits."""
    c = world.chunk("se_mid", 1011, 1, 40000, "se")
    sams = read_sams(c["dir"], "se")
    ref = ctx.count_chunk(*sams)

    def pieces():
        step, offs = 3_000_001, [0, 0, 0]
        while any(offs[f] < len(sams[f]) for f in range(3)):
            for f in (2, 0, 1):
                if offs[f] < len(sams[f]):
                    yield f, sams[f][offs[f]:offs[f] + step]
                    offs[f] += step
    got = ctx.count_stream(pieces())
    for m in eqp.MODES:
        assert np.array_equal(ref.fixed[m], got.fixed[m]), m
    assert got.number_mapped_reads == ref.number_mapped_reads
    empty = ctx.count_stream(iter(()))
    assert empty.number_mapped_reads == 0


BLOCK = np.dtype([("bucket", "<i4"), ("start", "<i4"), ("end", "<i4"), ("L", "<u2"), ("I", "<u2"), ("D", "<u2"), ("seg", "u1"), ("maxSeg", "u1"),
                  ("aln", "<u4"), ("readIdx", "u1"), ("paired", "u1"), ("strand", "u1"), ("pad", "u1"), ("slot", "<u4")])
HIT = np.dtype([("block", "<u4"), ("row", "<i4"), ("overlap", "<i4"), ("pad", "<u4")])


def test_selection_blocks_and_join_taps(world, ctx):
    """The intermediates behind the counts, against the oracle's files: which records survive the two CombineSamFiles runs
    (`selection` tap vs combined.sam), the BED6 blocks (`blocks` tap vs combined.bed, a8-a10), and the join (`hits` tap vs
    intersection.bed, a11) — the latter is the device's bucketed binary search against a file the oracle wrote and
    tests/test_shapes.py::test_join_against_brute_force checks against an O(|A| x |B|) restatement."""
    c = world.chunk("pe_small", 1003, 1, 3000, "pe")
    sams = read_sams(c["dir"], "pe")
    ctx.count_chunk(*sams)
    meta = ctx.debug_fetch("meta", np.uint32)
    base, n_slots = int(meta[0]), int(meta[1])
    # --- selection: per slot up to 4 parts (file, read-index filter); expand to the multiset of mapped+unmapped lines the reference prints
    sel = ctx.debug_fetch("selection", np.uint32)
    by_frag = [{}, {}, {}]
    for f in range(3):
        for line in sams[f].split(b"\n"):
            if line and not line.startswith(b"@"):
                by_frag[f].setdefault(int(line[1:10]), []).append(line)

    def read_index(line):
        flag = int(line.split(b"\t")[1])
        return 1 if not flag & 1 or flag & 64 or not flag & 128 else 2
    got = []
    for s in range(n_slots):
        v = int(sel[s])
        for k in range((v >> 16) & 7):
            f, flt = (v >> (4 * k)) & 3, (v >> (4 * k + 2)) & 3
            got += [l for l in by_frag[f].get(base + s, []) if not flt or read_index(l) == flt]
    want = [l for l in open(os.path.join(c["oracle"], "combined.sam"), "rb").read().split(b"\n") if l and not l.startswith(b"@")]
    assert got == want
    # --- blocks: (reference name, start, end, L, I, D, seg, maxSeg, P/S + read index, strand) per fragment, alignment ids relabelled
    blk = ctx.debug_fetch("blocks", BLOCK)
    names = ctx.bucket_names()
    gb = sorted((base + int(b["slot"]), names[b["bucket"]], int(b["start"]), int(b["end"]), int(b["L"]), int(b["I"]), int(b["D"]), int(b["seg"]), int(b["maxSeg"]),
                 ("P" if b["paired"] else "S") + str(int(b["readIdx"])), chr(b["strand"])) for b in blk)
    wb = []
    for l in open(os.path.join(c["oracle"], "combined.bed")):
        t = l.rstrip("\n").split("\t")
        nm = t[3]
        frag = int(nm[1:10])
        L, I, D = (int(nm.split("-" + k)[1].split("-")[0]) for k in ("L", "I", "D"))
        seg, rest = nm.split("-F")[1].split("-")
        wb.append((frag, t[0], int(t[1]), int(t[2]), L, I, D, int(seg), int(rest.split("/")[0]), rest.split("/")[1], t[5]))
    versionless = [x for x in wb if x[1] in set(names)]                   # references without rows in the B file produce no device blocks
    assert gb == sorted(versionless) and len(gb) > 3000
    # --- hits: (block coordinates, B row coordinates, overlap) against intersection.bed
    hits = ctx.debug_fetch("hits", HIT)
    rows = ctx.row_table()
    gh = sorted((names[blk[h["block"]]["bucket"]], int(blk[h["block"]]["start"]), int(blk[h["block"]]["end"]), base + int(blk[h["block"]]["slot"]),
                 int(rows[h["row"]][0]), int(rows[h["row"]][1]), int(h["overlap"])) for h in hits)
    wh = sorted((t[0], int(t[1]), int(t[2]), int(t[3][1:10]), int(t[7]), int(t[8]), int(t[12])) for t in (l.rstrip("\n").split("\t") for l in open(os.path.join(c["oracle"], "intersection.bed"))))
    assert gh == wh and len(gh) > 5000


@pytest.mark.parametrize("layout,seed", [("se", 1001), ("pe", 1003)])
def test_parity_at_bench_annotation_scale(tools, tmp_path_factory, layout, seed):
    """400,000 reads on the 60,000-gene annotation bench.py uses (SURVEY 8d shape mix at its default rates): what
    tools/parity_large.py runs at 1 M reads, sized so that the oracle needs about a minute per layout."""
    from conftest import World
    w = test_parity_at_bench_annotation_scale.__dict__.setdefault("world", None) or World(tools, tmp_path_factory.mktemp("world60k"), n_genes=60000)
    test_parity_at_bench_annotation_scale.world = w
    c = w.chunk(layout, seed, 1, 400000, layout)
    cx = test_parity_at_bench_annotation_scale.__dict__.get("ctx") or eqp.CountContext(0).load_annotation(w.anno)
    test_parity_at_bench_annotation_scale.ctx = cx
    check_chunk(cx, c, c["oracle"])
    check_weights(cx, c["oracle"])
