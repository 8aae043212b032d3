"""GPU test (-m gpu) of eqp_tokenize_noalloc, the allocation-free tokenizer entry point (include/eqp_count.h; no counterpart in the
reference).  Its output — the decoded-record columns in a buffer this test owns — must equal, bit for bit, what the existing tap
eqp_debug_decode reports for the same text, on two calls in a row (the second on a DIFFERENT, shorter text: the shared scratch array
keeps the first call's tile states and must have been cleared), and the call must leave the device's free memory unchanged.
Collected last (test_zz_*)."""
import numpy as np
import pytest

import eqp_cluster_b200 as eqp
from test_gpu_tokenizer import ANNO, TILE, valid_lines

pytestmark = pytest.mark.gpu

DT = {"line_off": np.uint64, "frag": np.uint32, "ref": np.int32, "pos": np.int32, "rnext": np.int32, "pnext": np.int32, "flag": np.uint16, "nm": np.int16,
      "cig_off": np.uint16, "cig_len": np.uint16, "seq_len": np.uint16, "cig_m": np.uint16, "rf": np.uint8}


@pytest.fixture(scope="module")
def ctx():
    c = eqp.CountContext(0).load_annotation(ANNO)
    yield c
    c.close()


def columns(out, cols, n):
    host = out.cpu().numpy()
    res = {}
    for k, dt in DT.items():
        off, cnt = getattr(cols, k), n + (1 if k == "line_off" else 0)
        res[k] = host[off:off + cnt * np.dtype(dt).itemsize].view(dt)
    return res


def rows_from_columns(c, n, data, names):
    """The rows eqp_debug_decode formats (eqp_capi.cu), rebuilt from the columns; `names`: reference index -> RNAME."""
    rows = []
    for i in range(n):
        b = int(c["line_off"][i])
        if not c["rf"][i] & 1:
            rows.append("%d\tSKIP" % i)
            continue
        cig = data[b + int(c["cig_off"][i]): b + int(c["cig_off"][i]) + int(c["cig_len"][i])].decode()
        rows.append("%d\tREC\t%d\t%d\t%s\t%d\t%d\t%d\t%s\t%d\t%d" % (i, c["frag"][i], c["flag"][i], names[int(c["ref"][i])], c["pos"][i], c["pnext"][i], c["nm"][i],
                                                                  cig, c["seq_len"][i], c["rf"][i]))
    return rows


def reference_names(c, n, want):
    """reference index -> RNAME as the existing entry point prints it; must be a function of the index."""
    names = {}
    for i in range(n):
        f = want[i].split("\t")
        if f[1] == "REC":
            assert names.setdefault(int(c["ref"][i]), f[4]) == f[4], (i, c["ref"][i], f[4])
    assert names.get(-1, "*") == "*"
    return names


def run_one(ctx, torch, data, max_lines, out, check_memory):
    text = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
    want, ec, _ = ctx.debug_decode(data, 0)                            # the existing entry point (allocates its own buffers)
    assert ec == 0
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    cols = ctx.tokenize_noalloc(text.data_ptr(), len(data), max_lines, out.data_ptr(), out.numel())
    free1 = torch.cuda.mem_get_info()[0]
    if check_memory:
        assert free1 == free0, (free0, free1)                          # (the first call of a process may load the kernel's module)
    n = int(cols.n_lines)
    assert cols.err_code == 0 and n == len(want), (cols.err_code, cols.err_detail, n, len(want))
    c = columns(out, cols, n)
    assert int(c["line_off"][n]) == (len(data) if data.endswith(b"\n") else len(data) + 1)     # entry i + 1 is one past line i's newline (real or virtual)
    got = rows_from_columns(c, n, data, reference_names(c, n, want))
    for i, (g, w) in enumerate(zip(got, want)):
        assert g == w, (i, g, w)                                       # integers and bytes: bit-exact, the bar of every tokenizer test here
    return n


def test_two_calls_in_a_row_match_the_existing_entry_point(ctx):
    import torch
    big = ("\n".join(valid_lines(12000, 21)) + "\n").encode()          # > 40 tiles
    small = ("@HD\tVN:1.0\n" + "\n".join(valid_lines(700, 22))).encode()   # 3 tiles, header line, no final newline
    assert len(big) > 40 * TILE and len(small) < len(big) // 8
    max_lines = 20000
    out = torch.empty(eqp.native.noalloc_out_bytes(max_lines), dtype=torch.uint8, device="cuda")
    n1 = run_one(ctx, torch, big, max_lines, out, check_memory=False)
    out.fill_(0xA5)                                                    # nothing of the first result may be needed by the second
    n2 = run_one(ctx, torch, small, max_lines, out, check_memory=True)
    n3 = run_one(ctx, torch, big, max_lines, out, check_memory=True)
    assert n1 == n3 and n2 < n1


def test_limits_are_refused(ctx):
    import torch
    data = ("\n".join(valid_lines(50, 23)) + "\n").encode()
    text = torch.frombuffer(bytearray(data), dtype=torch.uint8).cuda()
    out = torch.empty(eqp.native.noalloc_out_bytes(64), dtype=torch.uint8, device="cuda")
    with pytest.raises(eqp.EqpError, match="EQP_NOALLOC_MAX_TEXT_BYTES"):
        ctx.tokenize_noalloc(text.data_ptr(), eqp.native.NOALLOC_MAX_TEXT_BYTES + 1, 64, out.data_ptr(), out.numel())
    with pytest.raises(eqp.EqpError, match="EQP_NOALLOC_OUT_BYTES"):
        ctx.tokenize_noalloc(text.data_ptr(), len(data), 65, out.data_ptr(), out.numel())
    n_lines = len(ctx.debug_decode(data, 0)[0])
    assert n_lines > 8
    with pytest.raises(eqp.EqpError, match="has %d lines" % n_lines):
        ctx.tokenize_noalloc(text.data_ptr(), len(data), 8, out.data_ptr(), out.numel())
    cols = ctx.tokenize_noalloc(text.data_ptr(), len(data), 64, out.data_ptr(), out.numel())
    assert cols.n_lines == n_lines and cols.err_code == 0
