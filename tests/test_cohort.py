"""configs 4 and 5 in miniature through tools/cohort.py (the shipped sharding + accumulate + library reduce + writers), one rank."""
import json
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def run_cohort(tmp_path, extra):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "cohort.py"), "--genes", "300", "--workdir", str(tmp_path)] + extra,
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    return json.loads(r.stdout.strip().splitlines()[-1])


def test_cohort_matrix_against_oracle(tmp_path):
    line = run_cohort(tmp_path, ["--samples", "3", "--chunks", "2", "--reads", "20000", "--verify", "2", "--verify-reads", "5000"])
    assert line["chunks"] == 6 and len(line["verified_against_oracle"]) == 6           # 2 samples x 3 count files
    for c in line["verified_against_oracle"]:
        assert c["ids_equal"] and c["mismatch_gt_1e-9"] == 0, c
    matrix = open(os.path.join(str(tmp_path), "counts", "project-bowtie2-sr-gene.cnt")).read().splitlines()
    assert matrix[0].split("\t") == ["Id", "s01", "s02", "s03"] and len(matrix) == 301    # header + every gene of the map, zero-filled


def test_junction_sample_sharded_by_chunk(tmp_path):
    line = run_cohort(tmp_path, ["--workload", "junction", "--chunks", "2", "--reads", "30000", "--verify-reads", "6000"])
    assert len(line["chunk_heads_against_oracle"]) == 2
    for c in line["chunk_heads_against_oracle"]:
        assert c["ids_equal"] and c["mismatch_gt_1e-9"] == 0 and c["nonzero_junction_rows"] > 100, c
