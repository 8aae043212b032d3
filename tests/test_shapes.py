"""CPU suite: the annotation shapes of SURVEY.md 8d (transcript/genome exon length differences, tr- filler exons, .inv-strand,
versioned ids, 3-column maps) and the strand-specific mode — the device functions (host emulation, tests/emu) against the oracle
on one dense world per shape; plus a brute-force cross-check of the oracle's bedtools restatement."""
import os

import pytest

from conftest import SHAPES, assert_rows_match, read_rows, read_stats, read_weights


def check_emu(world, name, flags=()):
    c = world.chunks[name]
    o = world.oracle_run(name, flags) if tuple(flags) != c["flags"] else c["oracle"]
    emu = world.emu_run(name, flags)
    st = read_stats(os.path.join(o, "stats.txt"))
    for k in ("NUMBER_MAPPED_READS", "NUMBER_EXPRESSED_READS", "NUM_WEIGHT_OBJECTS"):
        assert emu["stats"][k] == st[k], k
    for m in ("GENE", "EXON", "JUNCTION"):                                # "Sum of weights contributing to count objects" (ComputeCounts.java:989)
        assert abs(float(emu["stats"]["SUM_WEIGHTS_" + m]) - float(st["SUM_WEIGHTS_" + m])) <= 0.01, m
    for m in ("gene", "exon", "junction"):
        assert_rows_match(emu[m + "_rows"], read_rows(os.path.join(o, m + ".cnt")), what="%s %s %s" % (name, flags, m))
    assert emu["weights"] == read_weights(o)
    return st


@pytest.mark.parametrize("layout,seed", [("se", 1001), ("pe", 1003)])
@pytest.mark.parametrize("shape", sorted(SHAPES))
def test_shape_matches_oracle(shape_worlds, shape, layout, seed):
    w = shape_worlds(shape)
    w.chunk(layout, seed, 1, 6000, layout)
    st = check_emu(w, layout)
    if shape in ("ld", "mixed"):
        assert int(st["SHAPE_SNAPS"]) > 0                                 # a17 really moved intervals (GenomicInterval.java:330-334)
    if shape in ("tr", "trlong", "mixed"):
        assert any("/tr-" in l for l in open(os.path.join(w.chunks[layout]["oracle"], "intersection.bed")))


@pytest.mark.parametrize("direction", ["forward", "backward"])
@pytest.mark.parametrize("layout,seed", [("se", 1001), ("pe", 1003)])
def test_strand_specific_mode(shape_worlds, layout, seed, direction):
    """-s <direction> end to end: fragment strand in the blocks (SamRecord.java:431-436,859-873), bedtools -s, BedRecord.java:650-656."""
    w = shape_worlds("mixed")
    w.chunk(layout, seed, 1, 6000, layout)
    flag = "--stranded=" + direction
    st = check_emu(w, layout, (flag,))
    base = read_stats(os.path.join(w.chunks[layout]["oracle"], "stats.txt"))
    assert int(st["NUMBER_EXPRESSED_READS"]) < int(base["NUMBER_EXPRESSED_READS"])   # the mode really filters


@pytest.mark.parametrize("flags", [("--unweighted",), ("--unambig",), ("--containment", "--nonsplice")])
def test_switches_on_the_mixed_world(shape_worlds, flags):
    w = shape_worlds("mixed")
    w.chunk("pe", 1003, 1, 6000, "pe")
    check_emu(w, "pe", flags)


@pytest.mark.parametrize("variant", ["_chain", "_retry"])
def test_alternate_paths_on_the_mixed_world(shape_worlds, variant):
    """The duplicate-chain walk for every fragment (EQP_FLAT_SCAN=0) and the large-table retry of the counting kernel (small tables of
    2 / 3 entries): the paths that only big fragments take in the product build, forced for all."""
    w = shape_worlds("mixed")
    c = w.chunk("pe", 1003, 1, 6000, "pe")
    emu = w.emu_run("pe", variant=variant)
    for m in ("gene", "exon", "junction"):
        assert_rows_match(emu[m + "_rows"], read_rows(os.path.join(c["oracle"], m + ".cnt")), what="%s %s" % (variant, m))
    st = read_stats(os.path.join(c["oracle"], "stats.txt"))
    for k in ("NUMBER_EXPRESSED_READS", "NUM_WEIGHT_OBJECTS"):
        assert emu["stats"][k] == st[k], k


@pytest.mark.parametrize("layout,seed", [("se", 1201), ("pe", 1203)])
def test_genomic_aligner_input(shape_worlds, layout, seed):
    """SURVEY 8f rank 2 (compute-genomic-counts.sh:1040-1058,1126-1152): one SAM of a spliced genome aligner — N CIGARs of 2-4 segments,
    indels, HI:i hit indices, secondary alignments — joined with the genome exon rows: BedRecord.computeGenomicInterval(ga) (:555-578) and
    the first / inner / last splice-conforming rules (:696-733) with maxSeg > 0."""
    w = shape_worlds("mixed")
    c = w.genomic_chunk("genomic_" + layout, seed, 1, 6000, layout)
    emu = w.emu_run("genomic_" + layout, ("--genomic",))
    st = read_stats(os.path.join(c["oracle"], "stats.txt"))
    for k in ("NUMBER_MAPPED_READS", "NUMBER_EXPRESSED_READS", "NUM_WEIGHT_OBJECTS"):
        assert emu["stats"][k] == st[k], k
    for m in ("gene", "exon"):
        assert_rows_match(emu[m + "_rows"], read_rows(os.path.join(c["oracle"], m + ".cnt")), what="genomic %s %s" % (layout, m))
    cigars = [l.split("\t")[5] for l in open(c["sam"]) if not l.startswith("@")]
    assert sum(x.count("N") >= 2 for x in cigars) > 20 and any("I" in x for x in cigars) and any("D" in x for x in cigars)
    inter = open(os.path.join(c["oracle"], "intersection.bed")).read()
    assert "-F1-2/" in inter and "-F2-2/" in inter                        # inner and last segments of three-segment alignments reach the counting


def test_default_world_has_the_survey_mix(world):
    """The default world (what bench.py counts, scaled down) carries every shape at the SURVEY 8d rates."""
    bed = open(os.path.join(world.anno, "ensembl_rna_hs_100_combined_exons.bed")).read()
    assert "/tr-" in bed and ".inv-strand" in bed and "KI27" in bed
    names = {l.split("\t")[0] for l in open(os.path.join(world.anno, "ensembl_rna_hs_transcript_exons.bed"))}
    base = {}
    for n in names:
        base.setdefault(n.split(".")[0], set()).add(n)
    assert any(len(v) > 1 for v in base.values()) and any("." not in n for n in names)   # two versions, and unversioned ids


def test_join_against_brute_force(world):
    """The oracle's `bedtools intersect -wo` restatement (oracle/convert.hpp::intersect_wo: sorted rows + running max end, the same
    idea the device uses) against an O(|A| x |B|) restatement of the documented semantics: 0-based half-open intervals, every A x B pair
    on the same chromosome string with at least 1 bp in common, A order kept, overlap = min(end) - max(start)."""
    c = world.chunk("se_small", 1001, 1, 3000, "se")
    b_rows = {}
    for k, l in enumerate(open(os.path.join(world.anno, "ensembl_rna_hs_100_combined_exons.bed"))):
        t = l.rstrip("\n").split("\t")
        b_rows.setdefault(t[0], []).append((int(t[1]), int(t[2]), k, l.rstrip("\n")))
    want = []
    for l in list(open(os.path.join(c["oracle"], "combined.bed")))[:4000]:
        t = l.rstrip("\n").split("\t")
        s, e = int(t[1]), int(t[2])
        for bs, be, k, bl in b_rows.get(t[0], ()):                        # file order = the order the oracle emits B hits in
            ov = min(e, be) - max(s, bs)
            if ov >= 1:
                want.append("%s\t%s\t%d" % (l.rstrip("\n"), bl, ov))
    got = [l.rstrip("\n") for l in open(os.path.join(c["oracle"], "intersection.bed"))][:len(want)]
    assert len(want) > 3000 and got == want
