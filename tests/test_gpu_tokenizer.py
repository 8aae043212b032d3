"""GPU tests (-m gpu) of the one-pass tokenizer kernel (eqp_tokenize.cuh) on its own, through the C ABI tap eqp_debug_decode:
every line's record must equal what the per-line device decoder produces in the host emulation (tests/emu, itself checked
against a Python model of SamRecord in test_decode_lines.py), for all tile-boundary / alignment / slow-path situations."""
import os
import random

import pytest

import eqp_cluster_b200 as eqp
from conftest import GOLDEN, run
from test_decode_lines import mk

pytestmark = pytest.mark.gpu

ANNO = os.path.join(GOLDEN, "appB", "anno")
TILE = 30704                                        # TK_OWNED: bytes of line starts per CTA


@pytest.fixture(scope="module")
def ctx():
    c = eqp.CountContext(0).load_annotation(ANNO)
    yield c
    c.close()


def emu_rows(tools, tmp_path, data):
    p = os.path.join(str(tmp_path), "t.sam")
    with open(p, "wb") as f:
        f.write(data)
    return run([tools["emu"], "decode", ANNO, p]).splitlines()


def check(ctx, tools, tmp_path, data, misaligns=(0,)):
    want = emu_rows(tools, tmp_path, data)
    assert not any("\tERR\t" in r for r in want), [r for r in want if "\tERR\t" in r][:3]
    for mis in misaligns:
        rows, ec, ed = ctx.debug_decode(data, mis)
        assert ec == 0, (mis, ec, ed)
        assert len(rows) == len(want), (mis, len(rows), len(want))
        for i, (g, w) in enumerate(zip(rows, want)):
            assert g == w, (mis, i, g, w)


def valid_lines(n, seed, seq_len=(1, 3, 16, 17, 33, 101)):
    rng = random.Random(seed)
    tags_ok = ["AS:i:-5", "XN:i:0", "XM:i:1", "NM:i:3", "nM:i:8", "MD:Z:55A45", "YT:Z:UU", "YS:i:-3", "ZS:i:12"]
    odd = ["foo", "NM:i:x", "ZZ:Z:a:b", "X", "zNM:i:4", "XS:Z:nM:i:12"]
    out = []
    for i in range(n):
        tags = [rng.choice(tags_ok) for _ in range(rng.randint(0, 9))]
        if rng.random() < 0.2:
            tags.insert(rng.randint(0, len(tags)), rng.choice(odd))
            tags = [t for t in tags if ":" in t or t == tags[-1]][:12]
            if sum(1 for t in tags if ":" not in t) > 1:
                tags = [t for t in tags if ":" in t]
        k = rng.choice(seq_len)
        seq = "".join(rng.choice("ACGT") for _ in range(k))
        qual = "".join(rng.choice("#5:;I") for _ in range(k))
        line = mk(qname="F%09d" % (i + 1), flag=rng.choice([0, 16, 4, 256, 272]), ref=rng.choice(["T1", "T2", "chr1", "chr2", "unknown", "*"]), pos=rng.randint(1, 10 ** 8),
                  rnext=rng.choice(["*", "=", "T2"]), pnext=rng.randint(0, 10 ** 6), seq=seq, qual=qual, cigar="%dM" % k, tags=tuple(tags))
        r = rng.random()
        if r < 0.05:
            line += "\r"
        elif r < 0.08:
            line += "\t"
        out.append(line)
        if rng.random() < 0.02:
            out.append("")
        if rng.random() < 0.02:
            out.append("@CO\tcomment %d" % i)
    return out


def test_small_inputs(ctx, tools, tmp_path):
    one = mk().encode()
    for data in (b"", b"\n", b"\n\n\n", one, one + b"\n", one + b"\n" + one, b"@HD\tVN:1.0\n" + one + b"\r\n", b"\n" + one):
        if data:
            check(ctx, tools, tmp_path, data, misaligns=(0, 1, 15))
        else:
            rows, ec, _ = ctx.debug_decode(data, 0)
            assert rows == [] and ec == 0


def test_many_tiles_all_alignments(ctx, tools, tmp_path):
    data = ("\n".join(valid_lines(12000, 11)) + "\n").encode()
    assert len(data) > 40 * TILE
    check(ctx, tools, tmp_path, data, misaligns=(0, 3, 8, 15))
    check(ctx, tools, tmp_path, data[:-1], misaligns=(0, 7))          # no final newline


def test_text_end_on_tile_boundary(ctx, tools, tmp_path):
    lines = valid_lines(400, 12, seq_len=(101,))
    body = ("\n".join(lines) + "\n").encode()
    n_tiles = len(body) // TILE
    for mis in (0, 9):
        for delta in (-1, 0, 1, 15, 16, 17):
            target = n_tiles * TILE - mis + delta                  # text length such that (length + misalignment) sits around a tile boundary
            cut = body[:target]
            cut = cut[:cut.rfind(b"\n") + 1]
            pad = target - len(cut)                                # refill with one record whose SEQ/QUAL make the length exact
            base = mk(qname="F000999999", seq="", qual="", cigar="1M", tags=()).encode()
            k = pad - len(base) - 1
            if k < 2 or k % 2:
                continue
            rec = mk(qname="F000999999", seq="A" * (k // 2), qual="I" * (k // 2), cigar="1M", tags=()).encode() + b"\n"
            data = cut + rec
            assert len(data) == target
            check(ctx, tools, tmp_path, data, misaligns=(mis,))
            check(ctx, tools, tmp_path, data[:-1], misaligns=(mis,))


def test_long_lines_take_the_slow_path(ctx, tools, tmp_path):
    lines = valid_lines(300, 13)
    for at in (5, 120, 299):
        lines[at] = mk(qname="F%09d" % (at + 1), seq="ACGT" * 1200, cigar="4800M")     # 9.6 KB line: longer than the 2 KB overhang
    data = ("\n".join(lines)).encode()
    check(ctx, tools, tmp_path, data, misaligns=(0, 5))
    check(ctx, tools, tmp_path, data + b"\n", misaligns=(0,))


def test_dense_headers_take_the_tile_slow_path(ctx, tools, tmp_path):
    hdr = ["@SQ\tSN:T%d\tLN:5" % i for i in range(9000)]               # > 2048 lines per tile
    lines = hdr + valid_lines(500, 14) + [""] * 5000 + valid_lines(200, 15)
    data = ("\n".join(lines) + "\n").encode()
    check(ctx, tools, tmp_path, data, misaligns=(0, 11))
    tabs = mk(tags=tuple(["AS:i:1"] * 3)) + "\t" * 9000                # > 6144 events in one tile: still one line, too many fields
    rows, ec, _ = ctx.debug_decode(("\n".join(valid_lines(50, 16) + [tabs]) + "\n").encode(), 0)
    want = emu_rows(tools, tmp_path, (tabs + "\n").encode())
    assert ec != 0 and want[0].split("\t")[1] == "ERR" and int(want[0].split("\t")[2]) == ec


def test_bad_lines_raise_the_same_code(ctx, tools, tmp_path):
    bad = [mk(tags=("plain", "plain2")), mk(tags=("HI:i:1",)), mk(qual="III"), mk(qname="F000000002/1", flag=128 | 1), mk(qname="G000000001"),
           mk(qname="F00000001"), mk(pos="x1"), mk(seq="A\tC"), "F000000001\t0\tT1"]
    good = valid_lines(200, 17)
    for b in bad:
        code = int(emu_rows(tools, tmp_path, (b + "\n").encode())[0].split("\t")[2])
        rows, ec, ed = ctx.debug_decode(("\n".join(good[:100] + [b] + good[100:]) + "\n").encode(), 4)
        assert ec == code, (b, ec, code)
        assert ed == emu_line_index(good[:100])


def emu_line_index(before):
    return len(before)


def test_short_lines_take_the_second_pass(ctx):
    """The record tables are sized for len/160 lines before the pass; texts whose lines average less take the pass twice
    (eqp_cuda_backend.cuh::decode_fused).  20,000 unaligned records of 45 bytes: every line decoded, nothing mapped; and the
    golden chunk with 60,000 such records behind its genome input counts to the same bits as without them."""
    import numpy as np
    from conftest import read_sams

    def unaligned(first, n):
        return b"".join(b"F%09d\t4\t*\t0\t0\t*\t*\t0\t0\tACGT\tIIII\n" % (first + i) for i in range(n))
    short = unaligned(1, 20000)
    assert len(short) / 160 + 1024 < 20000
    res = ctx.count_chunk(short, b"", b"")
    assert res.n_records[0] == 20000 and res.number_mapped_reads == 0
    rows, code, _ = ctx.debug_decode(short)
    assert code == 0 and len(rows) == 20000 and rows[19999].split("\t")[1:4] == ["REC", "20000", "4"]
    t, j, g = read_sams(os.path.join(GOLDEN, "appB", "chunk"), "se")
    ref = ctx.count_chunk(t, j, g)
    mixed = ctx.count_chunk(t, j, g + unaligned(1000, 60000))
    assert ref.number_mapped_reads == mixed.number_mapped_reads == 3
    for m in eqp.MODES:
        assert np.array_equal(ref.fixed[m], mixed.fixed[m]), m
