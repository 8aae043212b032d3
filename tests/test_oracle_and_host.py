"""CPU suite (-m "not gpu"): the oracle against the hand-traced golden vectors, the device functions emulated on the host
against the oracle, and the C-ABI library's exports.  No CUDA compute happens here."""
import ctypes
import os
import re

import pytest

import eqp_cluster_b200 as eqp
from conftest import GOLDEN, ROOT, assert_counts_match, read_cnt, read_stats, run


# ---------------------------------------------------------------- oracle vs golden vectors
def test_oracle_appendix_b_golden(tools, tmp_path):
    """SURVEY.md App. B worked example and its three weighting variants, traced by hand through the reference code."""
    g = os.path.join(GOLDEN, "appB")
    run([tools["oracle"], "run", os.path.join(g, "anno"), os.path.join(g, "chunk"), "se", str(tmp_path), "--quiet"])
    for name in ("gene.cnt", "exon.cnt", "junction.cnt", "final.wgt", "transcript-sam.wgt"):
        want = open(os.path.join(g, "expected", name)).read()
        assert open(os.path.join(str(tmp_path), name)).read() == want, name
    st = read_stats(os.path.join(str(tmp_path), "stats.txt"))
    assert st["NUMBER_MAPPED_READS"] == "3" and st["NUMBER_EXPRESSED_READS"] == "3"


def test_oracle_junction_filter_golden(tools, tmp_path):
    """tests/golden/appC: the per-chunk junction-map filter (a27) and a junction-reference read, traced by hand (expected/README.md)."""
    g = os.path.join(GOLDEN, "appC")
    run([tools["oracle"], "run", os.path.join(g, "anno"), os.path.join(g, "chunk"), "se", str(tmp_path), "--quiet"])
    for name in ("gene.cnt", "exon.cnt", "junction.cnt"):
        assert open(os.path.join(str(tmp_path), name)).read() == open(os.path.join(g, "expected", name)).read(), name
    assert "JCT_24" not in open(os.path.join(str(tmp_path), "exon_junction_filtered.map")).read()


def test_oracle_self_checks(world):
    """Invariants the reference asserts at run time hold on the oracle's intermediates."""
    c = world.chunk("se_small", 1001, 1, 3000, "se")
    prev = None
    for line in open(os.path.join(c["oracle"], "final.wgt")):          # CombineReadWeightFiles.java:434-451
        f = line.split("\t")[0]
        assert re.fullmatch(r"F\d{9}", f) and (prev is None or f > prev)
        prev = f
    for line in open(os.path.join(c["oracle"], "intersection.bed")):    # BedRecord.java:306-309, :335-337, :683-687
        t = line.rstrip("\n").split("\t")
        assert len(t) == 13 and t[0] == t[6]
        assert 1 <= int(t[12]) <= min(int(t[2]), int(t[8])) - max(int(t[1]), int(t[7]))
    for line in open(os.path.join(c["oracle"], "combined.bed")):        # SamRecord.java:463-466: block read lengths are positive
        name = line.split("\t")[3]
        assert int(re.search(r"-L(\d+)-", name).group(1)) > 0


# ---------------------------------------------------------------- device functions (host emulation) vs oracle
@pytest.mark.parametrize("name,seed,layout,spliced", [("se_small", 1001, "se", None), ("pe_small", 1003, "pe", None), ("se_spliced", 1004, "se", 0.4)])
def test_device_functions_match_oracle(world, name, seed, layout, spliced):
    c = world.chunk(name, seed, 1, 3000, layout, spliced)
    emu = world.emu_run(name)
    st = read_stats(os.path.join(c["oracle"], "stats.txt"))
    for k in ("NUMBER_MAPPED_READS", "NUMBER_EXPRESSED_READS", "NUM_WEIGHT_OBJECTS"):
        assert emu["stats"][k] == st[k], k
    for m in ("gene", "exon", "junction"):
        want = read_cnt(os.path.join(c["oracle"], m + ".cnt"))
        got = emu[m] if m == "junction" else {k: emu[m].get(k, 0.0) for k in want}
        assert_counts_match(got, want, what="%s %s" % (name, m))


def test_duplicate_chain_walk_matches_oracle(world):
    """a22-a25 read a fragment's hits flat up to EQP_FLAT_SCAN hits and walk the duplicate chains beyond: the chain path alone."""
    c = world.chunk("pe_small", 1003, 1, 3000, "pe")
    emu = world.emu_run("pe_small", variant="_chain")
    for m in ("gene", "exon", "junction"):
        want = read_cnt(os.path.join(c["oracle"], m + ".cnt"))
        got = emu[m] if m == "junction" else {k: emu[m].get(k, 0.0) for k in want}
        assert_counts_match(got, want, what="chain %s" % m)


@pytest.mark.parametrize("flags", [("--unweighted",), ("--unambig",), ("--containment", "--nonsplice")])
def test_device_functions_switches(world, flags):
    world.chunk("pe_small", 1003, 1, 3000, "pe")
    o = world.oracle_run("pe_small", flags)
    emu = world.emu_run("pe_small", flags)
    for m in ("gene", "exon", "junction"):
        want = read_cnt(os.path.join(o, m + ".cnt"))
        got = emu[m] if m == "junction" else {k: emu[m].get(k, 0.0) for k in want}
        assert_counts_match(got, want, exact=("--unweighted" in flags), what="%s %s" % (flags, m))


@pytest.mark.parametrize("case", ["appB", "appC"])
def test_device_functions_golden(tools, tmp_path, case):
    g = os.path.join(GOLDEN, case)
    run([tools["emu"], os.path.join(g, "anno"), os.path.join(g, "chunk"), "se", str(tmp_path)])
    for m in ("gene", "exon", "junction"):
        got = {l.split("\t")[0]: float(l.split("\t")[2]) for l in open(os.path.join(str(tmp_path), m + ".tsv"))}
        assert got == read_cnt(os.path.join(g, "expected", m + ".cnt"))


# ---------------------------------------------------------------- the C-ABI library
def test_library_exports_every_declared_symbol():
    lib_path = eqp.build.build_library()
    lib = ctypes.CDLL(lib_path)
    header = open(os.path.join(ROOT, "include", "eqp_count.h")).read()
    declared = set(re.findall(r"\b(eqp_[a-z_]+)\s*\(", header))
    assert set(eqp.EXPORTS) <= declared
    for sym in declared:
        assert hasattr(lib, sym), "libeqpcount.so does not export " + sym
    assert lib.eqp_version() == 2
    assert declared <= set(eqp.EXPORTS) | {"eqp_ctx"}, declared - set(eqp.EXPORTS)      # the ctypes binding covers the whole header


def test_params_struct_layout_and_defaults():
    p = eqp.default_params()
    assert (p.gene_overlap, p.exon_overlap, p.junction_overlap, p.modes) == (1, 5, 8, 7)   # compute-counts.sh:136-139
    assert p.read_weight_threshold == 0.01 and p.distance_slack == 0.25 and p.stranded == 0
    assert ctypes.sizeof(eqp.Params) == 64


def test_no_cpu_fallback():
    """Without a CUDA device the product path must fail loudly, never compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    with pytest.raises(eqp.EqpError) as e:
        eqp.CountContext(0)
    assert e.value.code == -2 and "ERROR" in str(e.value)


def test_product_sources_do_not_reference_the_oracle():
    pkg = os.path.join(ROOT, "eqp-cluster_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".cu", ".cuh", ".h", ".hpp", ".cpp")):
                assert "oracle/" not in open(os.path.join(d, f), errors="ignore").read(), f
