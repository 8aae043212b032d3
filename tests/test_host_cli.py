"""The C++ host drop-in (eqp-cluster_b200/host/compute_counts_main.cpp): compute-counts.sh's CLI, file names and log lines."""
import gzip
import os
import shutil
import subprocess

import pytest

import eqp_cluster_b200 as eqp
from conftest import read_cnt, read_stats

ANNO_LAYOUT = {  # reference layout under <project>/exon-pipeline-files (CIB:519-572, filter-junction-map-files.sh:207-240); True = stored gzipped
    "ensembl_rna_hs_100_combined_exons.bed": ("bed-files", False), "ensembl_rna_hs_transcript_exons.bed": ("bed-files", False),
    "ensembl_rna_hs_exon_gene.map": ("map-files", False), "ensembl_rna_hs_exon_exon.map": ("map-files", False),
    "ensembl_rna_hs_exon_junction.map": ("map-files", True), "ensembl_rna_hs_exon_junction_transcript.map": ("map-files", False),
    "ensembl_rna_hs_100_junction_junction_coordinate.map": ("map-files", True), "ensembl_rna_hs_equivalent_junctions.map": ("map-files", False),
    "genome.fa.fai": ("genome-files", False)}


def make_project(world, chunk, root, sample="s1", chunk_id="C001"):
    ex = os.path.join(root, "exon-pipeline-files")
    for name, (sub, gz) in ANNO_LAYOUT.items():
        os.makedirs(os.path.join(ex, sub), exist_ok=True)
        src = os.path.join(world.anno, name)
        if gz:
            with open(src, "rb") as f, gzip.open(os.path.join(ex, sub, name + ".gz"), "wb", compresslevel=1) as g:
                shutil.copyfileobj(f, g)
        else:
            shutil.copy(src, os.path.join(ex, sub, name))
    sam_dir = os.path.join(root, "samples", sample, chunk_id, "sam-files")
    os.makedirs(sam_dir)
    sfx = "pe" if chunk["layout"] == "pe" else "sr"
    paths = []
    for kind in ("transcript", "junction", "genome"):
        dst = os.path.join(sam_dir, "%s-%s-bowtie2-%s-%s.sam.gz" % (sample, chunk_id, kind, sfx))
        with open(os.path.join(chunk["dir"], "%s-%s.sam" % (kind, sfx)), "rb") as f, gzip.open(dst, "wb", compresslevel=1) as g:
            shutil.copyfileobj(f, g)
        paths.append(dst)
    return paths, os.path.join(root, "samples", sample, chunk_id, "count-files", "%s-%s-bowtie2-combined-%s" % (sample, chunk_id, sfx))


def test_cli_fails_loudly_without_gpu(world, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    cli = eqp.build.build_cli()
    eqp.build.build_library()
    c = world.chunk("se_small", 1001, 1, 3000, "se")
    paths, _ = make_project(world, c, str(tmp_path))
    r = subprocess.run([cli, "-a", "-nj", "-sr", str(tmp_path)] + paths, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 1 and "ERROR" in r.stderr and "successfully completed" not in r.stdout
    assert "READ_WEIGHT_THRESHOLD=0.01" in r.stdout and "JUNCTION_OVERLAP=8" in r.stdout and "PAIRED_END=FALSE" in r.stdout   # CC:641-646


@pytest.mark.gpu
@pytest.mark.parametrize("name,seed,layout,flags", [("se_small", 1001, "se", []), ("pe_small", 1003, "pe", ["--unweighted"])])
def test_cli_writes_reference_count_files(world, tmp_path, name, seed, layout, flags):
    cli = eqp.build.build_cli()
    c = world.chunk(name, seed, 1, 3000, layout)
    o = world.oracle_run(name, flags) if flags else c["oracle"]
    paths, out_base = make_project(world, c, str(tmp_path))
    r = subprocess.run([cli, "-a", "-nj"] + (["-sr"] if layout == "se" else []) + flags + [str(tmp_path)] + paths,
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 0, r.stderr
    st = read_stats(os.path.join(o, "stats.txt"))
    assert "NUMBER_MAPPED_READS=%s\n" % st["NUMBER_MAPPED_READS"] in r.stdout
    assert "NUMBER_EXPRESSED_READS=%s\n" % st["NUMBER_EXPRESSED_READS"] in r.stdout
    assert r.stdout.rstrip().endswith("Compute counts successfully completed.")
    for m in ("gene", "exon", "junction"):
        got_text, want_text = open("%s-%s.cnt" % (out_base, m)).read(), open(os.path.join(o, m + ".cnt")).read()
        if flags:                                                      # integer counts: the files are byte-identical
            assert got_text == want_text, m
            continue
        got_lines, want_lines = got_text.splitlines(), want_text.splitlines()
        assert [l.split("\t")[0] for l in got_lines] == [l.split("\t")[0] for l in want_lines], m      # same ids, same order
        for g, w in zip(got_lines, want_lines):
            gv, wv = g.split("\t")[1], w.split("\t")[1]
            assert gv == wv or (wv != "0" and gv != "0" and abs(float(gv) - float(wv)) <= 1e-9 * abs(float(wv))), (m, g, w)
