"""combine-counts writer (eqp-cluster_b200/combine.py) against hand-traced outputs of combineCounts.py (Python 2 semantics:
str(float) = 12 significant digits, bytewise id order, zero fill).  Python 2 is not available here, so these goldens are traced
by hand from the reference source (combineCounts.py:122-299) — parity unpinned, like the rest of the path."""
import gzip
import os

import numpy as np
import pytest

import eqp_cluster_b200 as eqp

cmb = eqp.combine


def put(tmp_path, name, text):
    p = os.path.join(str(tmp_path), name)
    with open(p, "w", newline="\n") as f:
        f.write(text)
    return p


def test_py2_float_str():
    cases = {0.0: "0.0", 1.0: "1.0", 2.5: "2.5", 1 / 3: "0.333333333333", 1e-5: "1e-05", 0.1 + 0.2: "0.3", 123456789012.0: "123456789012.0",
             1234567890123.0: "1.23456789012e+12", 1e16: "1e+16", -3.0: "-3.0", 0.0001: "0.0001", 0.00001234: "1.234e-05", 1234.5678901234567: "1234.56789012"}
    for v, s in cases.items():
        assert cmb.py2_float_str(v) == s, v
    assert cmb.py2_float_str(float("inf")) == "inf" and cmb.py2_float_str(float("nan")) == "nan"


def test_add_up_chunks_to_sample(tmp_path):
    a = put(tmp_path, "s-C001-gene.cnt", "ENSG2\t0\nENSG1\t1.5\nENSG3\t0.3333333333333333\n")
    b = put(tmp_path, "s-C002-gene.cnt", "ENSG1\t2\nENSG3\t0.16666666666666666\nENSG10\t7\n\n")
    out = os.path.join(str(tmp_path), "s-gene.cnt")
    n = cmb.combine_counts([a, b], out, add_up=True)
    assert n == 4
    # ids sorted bytewise (ENSG10 < ENSG2); ENSG2 only in the first file, ENSG10 only in the second; no header row with -A
    assert open(out).read() == "ENSG1\t3.5\nENSG10\t7.0\nENSG2\t0.0\nENSG3\t0.5\n"


def test_matrix_of_samples(tmp_path):
    sfx = "EQP-pe-gene.cnt"
    a = put(tmp_path, "S1-" + sfx, "G1\t3.5\nG2\t0.0\n")
    b = put(tmp_path, "S2_" + sfx, "G2\t4.25\nG3\t1e-05\n")
    c = put(tmp_path, "S3-" + sfx, "G1\t1.0\n")
    out = os.path.join(str(tmp_path), "P-" + sfx)
    cmb.main(["-i", a, b, c, "-o", out, "-s", sfx])
    assert open(out).read() == ("Id\tS1\tS2\tS3\n"
                                "G1\t3.5\t0.0\t1.0\n"           # missing in S2: appended 0
                                "G2\t0.0\t4.25\t0.0\n"
                                "G3\t0.0\t1e-05\t0.0\n")        # new in S2: zero-filled on the left


def test_id_file_orders_and_zero_fills(tmp_path):
    a = put(tmp_path, "A.cnt", "J2\t2\nJ1\t1\n")
    ids = os.path.join(str(tmp_path), "junction_exon.map.gz")
    with gzip.open(ids, "wt") as f:
        f.write("J3\tE1\nJ1\tE2\nJ2\tE3\n")
    out = os.path.join(str(tmp_path), "o.cnt")
    cmb.combine_counts([a], out, id_file=ids)
    assert open(out).read() == "Id\tA.cnt\nJ3\t0.0\nJ1\t1.0\nJ2\t2.0\n"


def test_filter_keeps_rows_with_a_nonzero_value(tmp_path):
    a = put(tmp_path, "A.cnt", "J1\t0.0\nJ2\t0.05\nJ3\t12\nJ4\t0\nJ5\t0.5\n")
    out = os.path.join(str(tmp_path), "o.cnt")
    cmb.combine_counts([a], out, add_up=True, use_filter=True)
    assert open(out).read() == "J2\t0.05\nJ3\t12.0\nJ5\t0.5\n"


def test_errors(tmp_path):
    a = put(tmp_path, "A.cnt", "G1\t1\t2\n")
    b = put(tmp_path, "B.cnt", "G1\t1\n")
    e = put(tmp_path, "E.cnt", "")
    out = os.path.join(str(tmp_path), "o.cnt")
    with pytest.raises(ValueError):
        cmb.combine_counts([a, b], out, add_up=True)             # column count mismatch
    with pytest.raises(ValueError):
        cmb.combine_counts([e, b], out, add_up=True)             # first file without rows
    c = put(tmp_path, "C.cnt", "G2\t1\t2\n")
    with pytest.raises(ValueError):
        cmb.combine_counts([a, c], out, multi_column=True)       # -M: every id must be in every file


def test_write_counts_equals_add_up_of_chunk_files(tmp_path):
    rng = np.random.default_rng(5)
    ids = ["ENSG%011d" % i for i in rng.permutation(3000)[:400]]
    fixed = [rng.integers(0, 1 << 42, size=len(ids), dtype=np.uint64) * (rng.random(len(ids)) < 0.6) for _ in range(3)]
    files = []
    for c, fx in enumerate(fixed):                                   # chunk files as the host CLI writes them: Double.toString of fixed / 2^38, "0" for zero
        vals = fx.astype(np.float64) / float(1 << eqp.EQP_FIXED_SHIFT)
        files.append(put(tmp_path, "s-C%03d-junction.cnt" % c, "".join("%s\t%s\n" % (i, repr(float(v))) for i, v in zip(ids, vals) if v != 0)))
    want = os.path.join(str(tmp_path), "want.cnt")
    cmb.combine_counts(files, want, add_up=True)
    total = (fixed[0] + fixed[1] + fixed[2]).astype(np.float64) / float(1 << eqp.EQP_FIXED_SHIFT)
    got = os.path.join(str(tmp_path), "got.cnt")
    cmb.write_counts(ids, total, got, nonzero_only=True)
    w, g = open(want).read().splitlines(), open(got).read().splitlines()
    assert [r.split("\t")[0] for r in w] == [r.split("\t")[0] for r in g]
    for rw, rg in zip(w, g):                                         # fp64 sum of chunk values vs exact fixed-point sum: same to 1e-9 (usually the same text)
        assert abs(float(rw.split("\t")[1]) - float(rg.split("\t")[1])) <= 1e-9 * float(rw.split("\t")[1])


def test_fpkm_hand_traced(tmp_path):
    """computeFpkm.py semantics traced by hand: lengths from the gene-length file first, then from the GTF (union of all rows
    of the gene per chromosome), ids without any length dropped, reads summed per sample over chunk rows."""
    fp = eqp.fpkm
    cnt = put(tmp_path, "P-gene.cnt", "Id\tS1\tS2\nG1\t10.0\t0.0\nG2\t3.0\t6.0\nG3\t1.0\t1.0\nENST9\t1.0\t1.0\n")
    lens = put(tmp_path, "lengths.txt", "Gene Id\tLength\nG1\t2000\n")
    reads = put(tmp_path, "reads.txt", "Sample\tChunk\tAligned reads\nS1\tC001\t1500000\nS1\tC002\t500000\nS2\t4000000\n")
    gtf = put(tmp_path, "a.gtf",
              'chr1\tx\texon\t101\t200\t.\t+\t.\tgene_id "G2"; transcript_id "T1";\n'
              'chr1\tx\texon\t151\t300\t.\t+\t.\tgene_id "G2"; transcript_id "T2";\n'          # overlaps the first: union 101-300 = 200
              'chr1\tx\texon\t301\t400\t.\t+\t.\tgene_id "G2"; transcript_id "T2";\n'          # starts after 300: separate, +100
              'chr2\tx\texon\t1\t50\t.\t+\t.\tgene_id "G2"; transcript_id "T3";\n')            # other chromosome: 50 < 300, ignored
    out = os.path.join(str(tmp_path), "P-gene.fpkm")
    fp.main(["-c", cnt, "-g", gtf, "-n", reads, "-o", out, "-l", lens])
    # G1: 10 / 2.0 / 2.0 = 2.5 ; 0 ; G2: length 300 -> 3 / 0.3 / 2.0 = 5.0 ; 6 / 0.3 / 4.0 = 5.0 ; G3 and ENST9 have no length
    assert open(out).read() == "Gene Id\tS1\tS2\nG1\t2.5\t0.0\nG2\t5.0\t5.0\n"
    assert fp.union_length([(5, 9), (1, 5), (20, 29)]) == 9 + 10 and fp.union_length([]) == 0
    fp.main(["-c", cnt, "-n", reads, "-o", out, "-l", lens, "-m", "-C", "S2"])
    assert open(out).read() == "Gene Id\tMean control FPKM\nG1\t0.0\n"
    lens2 = put(tmp_path, "lengths2.txt", "G1\t1000\nG2\t1000\nG3\t1000\nENST9\t1000\n")
    fp.main(["-c", cnt, "-o", out, "-l", lens2])                         # no -n: normalised by the column sums (15 and 8)
    rows = dict(r.split("\t", 1) for r in open(out).read().splitlines())
    assert rows["G1"] == "%s\t0.0" % cmb.py2_float_str(10.0 / 1.0 / (15.0 / 1e6)) and rows["ENST9"].split("\t")[1] == "125000.0"
