"""The device inflate's per-thread functions (eqp-cluster_b200/csrc/eqp_inflate.h) run on the host (tests/emu/inflate_emu.cpp: every kernel's
loop one "thread" after the other, the driver's chain walk and scratch waves restated) against zlib: the output must be the bytes zlib
gives, for every block type, segment size and wave size; files the decoder does not take must be refused, not mis-decoded."""
import gzip
import os
import random
import subprocess
import zlib

import pytest

import eqp_cluster_b200 as eqp


def sam_text(n, seed):
    rng = random.Random(seed)
    out = []
    for i in range(n):
        seq = "".join(rng.choice("ACGT") for _ in range(101))
        q = "".join(rng.choice("#5:;I") for _ in range(101))
        out.append("F%09d\t%d\tENST%08d\t%d\t42\t101M\t*\t0\t0\t%s\t%s\tAS:i:-%d\tXN:i:0\tXM:i:%d\tXO:i:0\tXG:i:0\tNM:i:%d\tMD:Z:101\tYT:Z:UU"
                   % (i + 1, rng.choice([0, 16, 256]), rng.randint(1, 20000), rng.randint(1, 5000), seq, q, rng.randint(0, 30), rng.randint(0, 3), rng.randint(0, 3)))
    return ("\n".join(out) + "\n").encode()


def pigz_like(data, piece=1 << 20, level=6):
    """One gzip member made of independently compressed pieces (what pigz writes): raw deflate per piece, sync-flushed, concatenated."""
    out = [b"\x1f\x8b\x08\x00\x00\x00\x00\x00\x00\x03"]
    for i in range(0, len(data), piece):
        c = zlib.compressobj(level, zlib.DEFLATED, -15)
        out.append(c.compress(data[i:i + piece]) + c.flush(zlib.Z_SYNC_FLUSH))
    c = zlib.compressobj(level, zlib.DEFLATED, -15)
    out.append(c.flush())
    out.append((zlib.crc32(data) & 0xFFFFFFFF).to_bytes(4, "little") + (len(data) & 0xFFFFFFFF).to_bytes(4, "little"))
    return b"".join(out)


@pytest.fixture(scope="module")
def emu():
    eqp.build.build_tools()
    return eqp.build.INFLATE_EMU_BIN


@pytest.fixture(scope="module")
def text():
    return sam_text(12000, 5)


def run_emu(emu, tmp_path, z, args=()):
    zp, op = os.path.join(str(tmp_path), "t.gz"), os.path.join(str(tmp_path), "t.out")
    with open(zp, "wb") as f:
        f.write(z)
    p = subprocess.run([emu, zp, op] + [str(a) for a in args], capture_output=True, text=True)
    return p, (open(op, "rb").read() if p.returncode == 0 else None)


@pytest.mark.parametrize("level", [1, 6, 9])
def test_levels_small_segments_and_waves(emu, tmp_path, text, level):
    p, out = run_emu(emu, tmp_path, gzip.compress(text, level), (8192, 8, 7))
    assert p.returncode == 0, p.stderr
    assert out == text
    assert "bogus 0" in p.stdout


@pytest.mark.parametrize("args", [(4096, 8, 50), (1 << 20, 8), (32768, 8, 2), (65536, 8)])
def test_segment_sizes(emu, tmp_path, text, args):
    p, out = run_emu(emu, tmp_path, gzip.compress(text, 6), args)
    assert p.returncode == 0, p.stderr
    assert out == text


def test_block_types_and_flushes(emu, tmp_path, text):
    cases = {"stored": gzip.compress(text[:500000], 0), "tiny (fixed block)": gzip.compress(b"hello\n", 6), "empty": gzip.compress(b"", 6),
             "one line": gzip.compress(text[:300], 6), "pigz-like": pigz_like(text, 1 << 18)}
    c = zlib.compressobj(6, zlib.DEFLATED, 31)
    z = b""
    for i in range(0, len(text), 250000):
        z += c.compress(text[i:i + 250000]) + c.flush(zlib.Z_SYNC_FLUSH if (i // 250000) % 2 else zlib.Z_FULL_FLUSH)
    cases["flushes"] = z + c.flush()
    want = {"stored": text[:500000], "tiny (fixed block)": b"hello\n", "empty": b"", "one line": text[:300], "pigz-like": text, "flushes": text}
    for name, z in cases.items():
        p, out = run_emu(emu, tmp_path, z, (8192, 8, 11))
        assert p.returncode == 0, (name, p.stderr)
        assert out == want[name], name


def test_header_fields_and_compressible_region(emu, tmp_path, text):
    hdr = "".join("@SQ\tSN:chr%d\tLN:%d\n" % (i, 1000 + i) for i in range(100000)).encode()
    body = hdr + text[:800000]
    z = gzip.compress(body, 6)
    named = b"\x1f\x8b\x08\x08" + z[4:10] + b"reads.sam\x00" + z[10:]               # FNAME set, as the gzip command writes it
    p, out = run_emu(emu, tmp_path, named, (8192, 16, 9))
    assert p.returncode == 0, p.stderr
    assert out == body
    p, out = run_emu(emu, tmp_path, named, (8192, 2, 1))                             # scratch smaller than one segment's output: one segment per wave
    assert p.returncode == 0 and out == body


def test_refusals(emu, tmp_path, text):
    p, _ = run_emu(emu, tmp_path, gzip.compress(text[:100000]) + gzip.compress(text[100000:200000]))
    assert p.returncode != 0 and "after the first gzip member" in p.stderr
    z = bytearray(gzip.compress(text, 6))
    z[len(z) // 2] ^= 0x55                                                            # corrupt: must fail, whatever the message
    p, out = run_emu(emu, tmp_path, bytes(z), (8192, 8))
    assert p.returncode != 0 or out != text
    p, _ = run_emu(emu, tmp_path, b"plain text, not gzip at all\n")
    assert p.returncode != 0
    z = bytearray(gzip.compress(text, 6))
    z[-6] ^= 0x01                                                                     # the trailer's CRC-32
    p, _ = run_emu(emu, tmp_path, bytes(z), (8192, 8))
    assert p.returncode != 0 and "CRC" in p.stderr
    z = bytearray(gzip.compress(text[:500000], 0))                                    # stored blocks: a flipped payload byte changes nothing but the CRC
    z[len(z) // 2] ^= 0x20
    p, _ = run_emu(emu, tmp_path, bytes(z), (8192, 8))
    assert p.returncode != 0 and "CRC" in p.stderr


@pytest.mark.parametrize("every", [1, 3])
def test_wrong_block_starts_are_dropped_by_the_chain_walk(emu, tmp_path, text, every):
    """Test hook: every k-th found start is moved one bit on.  The decoder in front of it passes it without landing on it, the host marks it
    bogus and re-runs that decoder through it: the output must not change (every = 1: no found start survives, one thread decodes everything)."""
    p, out = run_emu(emu, tmp_path, gzip.compress(text, 6), (8192, 8, 1 << 30, every))
    assert p.returncode == 0, p.stderr
    assert out == text
    assert "bogus 0" not in p.stdout


@pytest.mark.parametrize("groups", [1, 3, 7])
def test_window_resolve_by_groups(emu, tmp_path, text, groups):
    """The markers in every segment's last 32 K symbols are resolved per group of segments against a window of 16-bit entries, the groups'
    windows are then composed in order: any group size (and several scratch waves, 9 segments each) must give zlib's bytes."""
    p, out = run_emu(emu, tmp_path, gzip.compress(text, 6), (8192, 8, 9, 0, groups))
    assert p.returncode == 0, p.stderr
    assert out == text


def test_fuzz_against_zlib(emu, tmp_path, text):
    """Seeded sweep over content kinds, sizes, levels, strategies (fixed / Huffman-only / RLE blocks), memLevels, flush points, segment sizes down to
    512 bytes, scratch waves, forced wrong starts and resolve group sizes (1100 such cases were run once by hand; 60 stay in the suite)."""
    rng = random.Random(7)

    def content(kind, n):
        if kind == 0:
            o = rng.randrange(0, max(1, len(text) - n))
            return text[o:o + n]
        if kind == 1:
            return bytes(rng.choice(b"ACGT\t\n") for _ in range(n))
        if kind == 2:
            return (b"@SQ\tSN:chr%d\tLN:1000\n" % rng.randrange(99)) * (n // 24 + 1)
        if kind == 3:
            return bytes(rng.randrange(32, 127) for _ in range(n))
        return b"A" * n
    for it in range(60):
        n = rng.choice([0, 1, 2, 5, 100, 1000, 5000, 70000, 300000])
        data = content(rng.randrange(5), n)[:n]
        c = zlib.compressobj(rng.choice([1, 3, 6, 9]), zlib.DEFLATED, 31, rng.choice([1, 4, 8, 9]),
                             rng.choice([zlib.Z_DEFAULT_STRATEGY, zlib.Z_FILTERED, zlib.Z_HUFFMAN_ONLY, zlib.Z_RLE, zlib.Z_FIXED]))
        z, i = b"", 0
        while i < len(data):
            k = rng.choice([len(data), 1000, 50000])
            z += c.compress(data[i:i + k])
            i += k
            if rng.random() < 0.2:
                z += c.flush(rng.choice([zlib.Z_SYNC_FLUSH, zlib.Z_FULL_FLUSH]))
        z += c.flush()
        args = (rng.choice([512, 4096, 8192, 65536]), 8, rng.choice([1, 3, 50, 1 << 20]), rng.choice([0, 0, 0, 2, 5]), rng.choice([1, 4, 296]))
        p, out = run_emu(emu, tmp_path, z, args)
        assert p.returncode == 0, (it, args, p.stderr)
        assert out == data, (it, args, len(data))
