"""GPU tests (-m gpu) of the device inflate through the C ABI: eqp_debug_inflate must return the bytes zlib returns, and a chunk counted
from three gzip members (eqp_count_chunk_gz) must give the counts of the same chunk counted from its text."""
import gzip
import os

import numpy as np
import pytest

import eqp_cluster_b200 as eqp
from conftest import GOLDEN
from test_inflate import pigz_like, sam_text

pytestmark = pytest.mark.gpu
ANNO = os.path.join(GOLDEN, "appB", "anno")


@pytest.fixture(scope="module")
def ctx():
    c = eqp.CountContext(0).load_annotation(ANNO)
    yield c
    c.close()


@pytest.fixture(scope="module")
def text():
    return sam_text(60000, 9)                                             # 18 MB, ~150 deflate blocks


def test_bytes_equal_zlib(ctx, text, monkeypatch):
    cases = [gzip.compress(text, 6), gzip.compress(text, 1), gzip.compress(text[:700000], 0), pigz_like(text), gzip.compress(b"", 6), gzip.compress(b"x\n", 6)]
    want = [text, text, text[:700000], text, b"", b"x\n"]
    for seg in ("65536", "8192"):
        monkeypatch.setenv("EQP_INFLATE_SEG", seg)
        for z, w in zip(cases, want):
            got = ctx.debug_inflate(z)
            assert got == w, (seg, len(z), len(got), len(w))
    info = ctx.last_inflate(0)
    assert info["bogus_starts"] == 0


def test_small_scratch_runs_in_waves(ctx, text, monkeypatch):
    monkeypatch.setenv("EQP_INFLATE_SEG", "8192")
    monkeypatch.setenv("EQP_INFLATE_SCRATCH_GB", "0.02")                  # 10 M symbols of scratch for 72 MB of text: several waves
    big = text * 4
    got = ctx.debug_inflate(gzip.compress(big, 6))
    assert got == big
    assert ctx.last_inflate(0)["waves"] > 1


def test_refusals(ctx, text):
    with pytest.raises(eqp.EqpError, match="after the first gzip member"):
        ctx.debug_inflate(gzip.compress(text[:100000]) + gzip.compress(text[100000:200000]))
    with pytest.raises(eqp.EqpError):
        ctx.debug_inflate(b"plain text, not gzip at all, but long enough\n")
    z = bytearray(gzip.compress(text, 6))
    z[-6] ^= 0x01                                                          # the trailer's CRC-32: the text itself inflates fine, the check must catch it
    with pytest.raises(eqp.EqpError, match="CRC"):
        ctx.debug_inflate(bytes(z))
    z = bytearray(gzip.compress(text[:2_000_000], 0))                      # stored blocks: a flipped payload byte changes nothing but the CRC
    z[len(z) // 2] ^= 0x20
    with pytest.raises(eqp.EqpError, match="CRC"):
        ctx.debug_inflate(bytes(z))
    z = bytearray(gzip.compress(text, 6))
    z[len(z) // 2] ^= 0x55
    try:
        assert ctx.debug_inflate(bytes(z)) != text
    except eqp.EqpError:
        pass


def test_chunk_from_gzip_equals_chunk_from_text(ctx):
    d = os.path.join(GOLDEN, "appB", "chunk")
    sams = [open(os.path.join(d, "%s-sr.sam" % k), "rb").read() for k in ("transcript", "junction", "genome")]
    a = ctx.count_chunk(*sams)
    b = ctx.count_chunk_gz(*[gzip.compress(s, 6) for s in sams])
    for m in eqp.MODES:
        assert np.array_equal(a.fixed[m], b.fixed[m]), m
    assert a.number_mapped_reads == b.number_mapped_reads


def test_wrong_block_starts_and_multi_segment_groups(ctx, text, monkeypatch):
    """The retry path (a chain member passes a claimed start: it is dropped, the member re-run through it — by ONE thread, hence the small
    inputs) and resolve groups of several segments (the suite's small inputs otherwise give every segment its own group)."""
    monkeypatch.setenv("EQP_INFLATE_SEG", "8192")
    monkeypatch.setenv("EQP_INFLATE_GROUPS", "7")
    assert ctx.debug_inflate(gzip.compress(text, 6)) == text
    assert ctx.last_inflate(0)["bogus_starts"] == 0
    part = text[:1_500_000]
    monkeypatch.setenv("EQP_INFLATE_TEST_BOGUS", "2")
    assert ctx.debug_inflate(gzip.compress(part, 6)) == part
    assert ctx.last_inflate(0)["bogus_starts"] >= 3
    part = text[:300_000]
    monkeypatch.setenv("EQP_INFLATE_TEST_BOGUS", "1")
    assert ctx.debug_inflate(gzip.compress(part, 6)) == part
    assert ctx.last_inflate(0)["chain"] == 1
