"""GPU parity tests (-m gpu): the CUDA path, called through the C ABI, against the oracle on the same seeded inputs."""
import os

import numpy as np
import pytest

import eqp_cluster_b200 as eqp
from conftest import GOLDEN, assert_counts_match, read_cnt, read_sams, read_stats

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
    for m in eqp.MODES:
        want = read_cnt(os.path.join(oracle_dir, m + ".cnt"))
        got = res.as_dict(m, nonzero_only=(m == "junction"))
        if m != "junction":
            assert set(got) == set(want), m                           # every id of the map is printed (ComputeCounts.java:877-897)
        assert_counts_match(got, want, exact=exact, what=m)
    return res


@pytest.mark.parametrize("name,seed,layout,spliced,n", [
    ("se_small", 1001, "se", None, 3000), ("pe_small", 1003, "pe", None, 3000), ("se_spliced", 1004, "se", 0.4, 3000),
    ("se_mid", 1011, "se", None, 40000), ("pe_mid", 1013, "pe", None, 40000)])
def test_counts_match_oracle(world, ctx, name, seed, layout, spliced, n):
    c = world.chunk(name, seed, 1, n, layout, spliced)
    check_chunk(ctx, c, c["oracle"])


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


def test_golden_appendix_b():
    g = os.path.join(GOLDEN, "appB")
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
    p.stranded = 1
    with pytest.raises(eqp.EqpError) as e:
        ctx.count_chunk(t, j, g, params=p)
    assert e.value.code == -5


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
    n = None
    acc = None
    for h in (0, 1):
        ctx.count_chunk(*[x[h] for x in halves])
        if acc is None:
            _, n = ctx.device_vector()
            acc = torch.zeros(n, dtype=torch.int64, device="cuda")
        ctx.accumulate(acc.data_ptr())
    got = acc.cpu().numpy().view(np.uint64)[:want.size]
    assert n == want.size + res.fixed["junction"].size and np.array_equal(got, want)


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
            _, n = ctx.device_vector()
            acc = torch.zeros(n, dtype=torch.int64, device="cuda")
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
