"""ctypes binding of libeqpcount.so (include/eqp_count.h).  No compute happens in Python and there is no fallback:
if the library is missing or no CUDA device is present, the calls raise."""
import ctypes as C
import os

import numpy as np

from . import build as _build

EQP_FIXED_SHIFT = 38
MODES = ("gene", "exon", "junction")
ANNOTATION_FILES = (          # field order of eqp_annotation_files; names under <project>/exon-pipeline-files (CC:584-603)
    ("combined_exons_bed", "{base}_{radius}_combined_exons.bed"),
    ("transcript_exons_bed", "{base}_transcript_exons.bed"),
    ("exon_gene_map", "{base}_exon_gene.map"),
    ("exon_exon_map", "{base}_exon_exon.map"),
    ("exon_junction_map", "{base}_exon_junction.map"),
    ("exon_junction_transcript_map", "{base}_exon_junction_transcript.map"),
    ("junction_junction_coordinate_map", "{base}_{radius}_junction_junction_coordinate.map"),
    ("equivalent_junctions_map", "{base}_equivalent_junctions.map"),
    ("genome_fai", None),
)


class EqpError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("ERROR: %s (eqp code %d)" % (msg, code))
        self.code = code


class AnnotationFiles(C.Structure):
    _fields_ = [(n, C.c_char_p) for n, _ in ANNOTATION_FILES]


class Params(C.Structure):
    _fields_ = [("paired_end", C.c_int32), ("stranded", C.c_int32), ("gene_overlap", C.c_int32), ("exon_overlap", C.c_int32),
                ("junction_overlap", C.c_int32), ("read_weight_threshold", C.c_double), ("unweighted", C.c_int32),
                ("unambiguous", C.c_int32), ("containment", C.c_int32), ("nonsplice_conforming", C.c_int32),
                ("distance_slack", C.c_double), ("modes", C.c_int32), ("genomic_input", C.c_int32)]


class Counts(C.Structure):
    _fields_ = [("n_gene", C.c_int64), ("n_exon", C.c_int64), ("n_junction", C.c_int64),
                ("gene", C.POINTER(C.c_double)), ("exon", C.POINTER(C.c_double)), ("junction", C.POINTER(C.c_double)),
                ("gene_fixed", C.POINTER(C.c_uint64)), ("exon_fixed", C.POINTER(C.c_uint64)), ("junction_fixed", C.POINTER(C.c_uint64)),
                ("junction_active", C.POINTER(C.c_uint8)), ("n_junction_act", C.c_int64),
                ("number_mapped_reads", C.c_int64), ("number_expressed_reads", C.c_int64),
                ("num_weight_objects", C.c_int64 * 3), ("num_weight_objects_included", C.c_int64 * 3),
                ("num_weight_objects_excluded", C.c_int64 * 3), ("n_records", C.c_int64 * 3), ("sum_weights", C.c_double * 3),
                ("device_ms", C.c_double), ("dominant_kernel_ms", C.c_double), ("kernel_launches", C.c_int64)]


class OutputRow(C.Structure):
    _fields_ = [("id", C.c_char_p), ("feature", C.c_int64), ("act", C.c_int64), ("always", C.c_int32), ("reserved", C.c_int32)]


class InflateInfo(C.Structure):
    """eqp_inflate_info: what the device inflate of one gzip member did."""
    _fields_ = [("gz_bytes", C.c_uint64), ("text_bytes", C.c_uint64), ("segments", C.c_uint32), ("chain", C.c_uint32), ("bogus_starts", C.c_uint32), ("waves", C.c_uint32),
                ("search_ms", C.c_float), ("decode_ms", C.c_float), ("resolve_ms", C.c_float), ("emit_ms", C.c_float)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class RecordColumns(C.Structure):
    """eqp_record_columns: byte offsets of the decoded-record columns inside the caller's device buffer."""
    _fields_ = [(k, C.c_uint64) for k in ("line_off", "frag", "ref", "pos", "rnext", "pnext", "flag", "nm", "cig_off", "cig_len", "seq_len", "cig_m", "rf", "n_lines")] + \
               [("err_code", C.c_uint32), ("err_detail", C.c_uint32)]


EXPORTS = ("eqp_version", "eqp_default_params", "eqp_create", "eqp_load_annotation", "eqp_annotation_bytes", "eqp_count_chunk", "eqp_prefetch_chunk",
           "eqp_feature_count", "eqp_feature_id", "eqp_map_columns", "eqp_output_row_count", "eqp_get_output_row", "eqp_device_vector", "eqp_accumulate",
           "eqp_wide_to_double", "eqp_comm_id", "eqp_comm_init", "eqp_reduce_sample", "eqp_reduce_sample_local", "eqp_chunk_begin", "eqp_chunk_feed",
           "eqp_chunk_end", "eqp_debug_fetch", "eqp_debug_decode", "eqp_tokenize_noalloc", "eqp_chunk_feed_gz", "eqp_count_chunk_gz",
           "eqp_last_inflate", "eqp_debug_inflate", "eqp_last_error", "eqp_destroy")
NOALLOC_MAX_TEXT_BYTES = 128 << 20


def noalloc_out_bytes(max_lines):
    """EQP_NOALLOC_OUT_BYTES(max_lines) of include/eqp_count.h."""
    return max_lines * 41 + 8 + 13 * 256
COMM_ID_BYTES = 128

_lib = None


def load_library(path=None):
    """dlopen libeqpcount.so (building it in-tree if needed) and declare the prototypes of include/eqp_count.h.
    `path` (or $EQP_LIB) names another build of the library: tools/variants.py times several builds side by side."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or os.environ.get("EQP_LIB") or _build.LIB_PATH
    if not os.path.exists(path):
        _build.build_library()
    lib = C.CDLL(path)
    vp = C.c_void_p
    lib.eqp_version.restype = C.c_int
    lib.eqp_default_params.argtypes = [C.POINTER(Params)]; lib.eqp_default_params.restype = None
    lib.eqp_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.eqp_load_annotation.argtypes = [vp, C.POINTER(AnnotationFiles)]
    lib.eqp_count_chunk.argtypes = [vp, C.POINTER(Params), vp, C.c_size_t, vp, C.c_size_t, vp, C.c_size_t, C.c_int, C.POINTER(Counts)]
    lib.eqp_prefetch_chunk.argtypes = [vp, vp, C.c_size_t, vp, C.c_size_t, vp, C.c_size_t]
    lib.eqp_prefetch_chunk.restype = C.c_int
    lib.eqp_feature_count.argtypes = [vp, C.c_int]; lib.eqp_feature_count.restype = C.c_int64
    lib.eqp_feature_id.argtypes = [vp, C.c_int, C.c_int64]; lib.eqp_feature_id.restype = C.c_char_p
    lib.eqp_device_vector.argtypes = [vp, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
    lib.eqp_accumulate.argtypes = [vp, vp]
    if not hasattr(lib, "eqp_map_columns"):                              # an ABI-1 build (tools/variants.py times old builds): the common entry points only
        lib.eqp_last_error.argtypes = [vp]; lib.eqp_last_error.restype = C.c_char_p
        lib.eqp_destroy.argtypes = [vp]; lib.eqp_destroy.restype = None
        lib.eqp_debug_fetch.argtypes = [vp, C.c_char_p, vp, C.c_size_t, C.POINTER(C.c_size_t)]
        return lib
    lib.eqp_map_columns.argtypes = [vp, C.c_int]
    lib.eqp_output_row_count.argtypes = [vp, C.c_int]; lib.eqp_output_row_count.restype = C.c_int64
    lib.eqp_get_output_row.argtypes = [vp, C.c_int, C.c_int64, C.POINTER(OutputRow)]
    lib.eqp_wide_to_double.argtypes = [vp, C.c_int64, vp]; lib.eqp_wide_to_double.restype = None
    lib.eqp_comm_id.argtypes = [vp]
    lib.eqp_comm_init.argtypes = [vp, vp, C.c_int, C.c_int]
    lib.eqp_reduce_sample.argtypes = [vp, vp, C.c_int]
    lib.eqp_reduce_sample_local.argtypes = [C.POINTER(vp), C.c_int, C.POINTER(vp), C.c_int]
    lib.eqp_chunk_begin.argtypes = [vp, C.c_size_t, C.c_size_t, C.c_size_t]
    lib.eqp_chunk_feed.argtypes = [vp, C.c_int, vp, C.c_size_t]
    lib.eqp_chunk_end.argtypes = [vp, C.POINTER(Params), C.POINTER(Counts)]
    lib.eqp_debug_fetch.argtypes = [vp, C.c_char_p, vp, C.c_size_t, C.POINTER(C.c_size_t)]
    lib.eqp_debug_decode.argtypes = [vp, vp, C.c_size_t, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    if hasattr(lib, "eqp_count_chunk_gz"):
        lib.eqp_chunk_feed_gz.argtypes = [vp, C.c_int, vp, C.c_size_t]
        lib.eqp_count_chunk_gz.argtypes = [vp, C.POINTER(Params), vp, C.c_size_t, vp, C.c_size_t, vp, C.c_size_t, C.POINTER(Counts)]
        lib.eqp_last_inflate.argtypes = [vp, C.c_int, C.POINTER(InflateInfo)]
        lib.eqp_debug_inflate.argtypes = [vp, vp, C.c_size_t, vp, C.c_size_t, C.POINTER(C.c_size_t)]
    if hasattr(lib, "eqp_tokenize_noalloc"):
        lib.eqp_tokenize_noalloc.argtypes = [vp, vp, C.c_size_t, C.c_uint64, vp, C.c_size_t, C.POINTER(RecordColumns)]
    lib.eqp_last_error.argtypes = [vp]; lib.eqp_last_error.restype = C.c_char_p
    lib.eqp_destroy.argtypes = [vp]; lib.eqp_destroy.restype = None
    lib.eqp_annotation_bytes.argtypes = [vp]; lib.eqp_annotation_bytes.restype = C.c_int64
    if _lib is None:
        _lib = lib
    return lib


def default_params():
    p = Params()
    load_library().eqp_default_params(C.byref(p))
    return p


def annotation_paths(anno_dir, base="ensembl_rna_hs", radius=100):
    out = {}
    for field, pattern in ANNOTATION_FILES:
        out[field] = os.path.join(anno_dir, "genome.fa.fai" if pattern is None else pattern.format(base=base, radius=radius))
    return out


def _buf(x):
    """(address, length, keepalive) of a bytes / numpy uint8 array / (int device_ptr, length) tuple."""
    if isinstance(x, tuple):
        return int(x[0]), int(x[1]), None
    if isinstance(x, (bytes, bytearray)):
        a = np.frombuffer(x, dtype=np.uint8)
        return (a.ctypes.data if a.size else 0), a.size, a
    a = np.ascontiguousarray(x, dtype=np.uint8)
    return (a.ctypes.data if a.size else 0), a.size, a


class ChunkCounts:
    """Host copy of one chunk's eqp_counts."""

    def __init__(self, c, feature_ids, rows=None, columns=None):
        self.output_rows, self.map_columns = rows, columns
        self.sum_weights = list(c.sum_weights)

        def arr(ptr, n, dt):
            return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dt, copy=True) if n else np.zeros(0, dt)
        n = (c.n_gene, c.n_exon, c.n_junction)
        self.counts = {m: arr(getattr(c, m), n[i], np.float64) for i, m in enumerate(MODES)}
        self.fixed = {m: arr(getattr(c, m + "_fixed"), n[i], np.uint64) for i, m in enumerate(MODES)}
        self.junction_active = arr(c.junction_active, c.n_junction_act, np.uint8)
        self.feature_ids = feature_ids
        self.number_mapped_reads = c.number_mapped_reads
        self.number_expressed_reads = c.number_expressed_reads
        self.num_weight_objects = list(c.num_weight_objects)
        self.num_weight_objects_included = list(c.num_weight_objects_included)
        self.num_weight_objects_excluded = list(c.num_weight_objects_excluded)
        self.n_records = list(c.n_records)
        self.device_ms = c.device_ms
        self.dominant_kernel_ms = c.dominant_kernel_ms
        self.kernel_launches = c.kernel_launches

    def as_dict(self, mode, nonzero_only=False):
        ids, v = self.feature_ids[mode], self.counts[mode]
        return {ids[i]: float(v[i]) for i in range(len(ids)) if not nonzero_only or self.fixed[mode][i]}

    def rows(self, mode, nonzero_only=None):
        """(printed id, value) in the reference's print order (ComputeCounts.java:875-950): every row of the map for the gene and exon
        runs, the rows of this chunk's filtered junction map for the junction run (non-zero ones only under -n, the default flow)."""
        if nonzero_only is None:
            nonzero_only = mode == "junction"
        out, v, fx, act = [], self.counts[mode], self.fixed[mode], self.junction_active
        for rid, feat, a, always in self.output_rows[mode]:
            if mode == "junction" and not (always or act[a]):
                continue
            if nonzero_only and not fx[feat]:
                continue
            out.append((rid, float(v[feat])))
        return out


class CountContext:
    """One eqp_ctx: bound to one GPU, one chunk in flight (include/eqp_count.h)."""

    def __init__(self, device=0, lib_path=None):
        self.lib = load_library(lib_path)
        self.h = C.c_void_p()
        rc = self.lib.eqp_create(device, C.byref(self.h))
        if rc:
            raise EqpError(rc, (self.lib.eqp_last_error(None) or b"").decode())
        self.feature_ids = None
        self._prefetch_keep = []

    def _check(self, rc):
        if rc:
            raise EqpError(rc, (self.lib.eqp_last_error(self.h) or b"").decode())

    def load_annotation(self, anno_dir=None, paths=None, base="ensembl_rna_hs", radius=100):
        paths = paths or annotation_paths(anno_dir, base, radius)
        f = AnnotationFiles(**{k: (v.encode() if v is not None else None) for k, v in paths.items()})
        self._check(self.lib.eqp_load_annotation(self.h, C.byref(f)))
        self.feature_ids, self.output_rows, self.map_columns = {}, {}, {}
        row = OutputRow()
        for i, m in enumerate(MODES):
            n = self.lib.eqp_feature_count(self.h, i)
            self.feature_ids[m] = [self.lib.eqp_feature_id(self.h, i, k).decode() for k in range(n)]
            if not hasattr(self.lib, "eqp_map_columns"):
                continue
            self.map_columns[m] = self.lib.eqp_map_columns(self.h, i)
            rows = []
            for k in range(self.lib.eqp_output_row_count(self.h, i)):
                self._check(self.lib.eqp_get_output_row(self.h, i, k, C.byref(row)))
                rows.append((row.id.decode(), row.feature, row.act, row.always))
            self.output_rows[m] = rows
        return self

    @property
    def vector_length(self):
        return sum(len(self.feature_ids[m]) for m in MODES)

    @property
    def annotation_bytes(self):
        return self.lib.eqp_annotation_bytes(self.h)

    def count_chunk(self, t_sam, j_sam, g_sam, params=None, on_device=False, fetch=True):
        p = params or default_params()
        bufs = [_buf(x) for x in (t_sam, j_sam, g_sam)]
        c = Counts()
        self._check(self.lib.eqp_count_chunk(self.h, C.byref(p), bufs[0][0], bufs[0][1], bufs[1][0], bufs[1][1], bufs[2][0], bufs[2][1],
                                              1 if on_device else 0, C.byref(c)))
        if self._prefetch_keep and not on_device:
            self._prefetch_keep.pop(0)
        return ChunkCounts(c, self.feature_ids, self.output_rows, self.map_columns) if fetch else c

    def count_stream(self, pieces, params=None, hints=(0, 0, 0)):
        """Streaming ingest (eqp_chunk_begin/feed/end): `pieces` yields (input, bytes) with input 0/1/2 = transcript/junction/genome."""
        p = params or default_params()
        self._check(self.lib.eqp_chunk_begin(self.h, *[int(h) for h in hints]))
        for which, data in pieces:
            addr, n, keep = _buf(data)
            self._check(self.lib.eqp_chunk_feed(self.h, int(which), addr, n))
        c = Counts()
        self._check(self.lib.eqp_chunk_end(self.h, C.byref(p), C.byref(c)))
        return ChunkCounts(c, self.feature_ids, self.output_rows, self.map_columns)

    def count_chunk_gz(self, t_gz, j_gz, g_gz, params=None, fetch=True):
        """eqp_count_chunk_gz: three gzip members (host memory) are copied compressed and inflated on the device."""
        p = params or default_params()
        bufs = [_buf(x) for x in (t_gz, j_gz, g_gz)]
        c = Counts()
        self._check(self.lib.eqp_count_chunk_gz(self.h, C.byref(p), bufs[0][0], bufs[0][1], bufs[1][0], bufs[1][1], bufs[2][0], bufs[2][1], C.byref(c)))
        return ChunkCounts(c, self.feature_ids, self.output_rows, self.map_columns) if fetch else c

    def last_inflate(self, which):
        info = InflateInfo()
        self._check(self.lib.eqp_last_inflate(self.h, int(which), C.byref(info)))
        return info.as_dict()

    def debug_inflate(self, gz):
        """Test tap: the text the device inflate makes of one gzip member (bytes)."""
        addr, n, keep = _buf(gz)
        size = C.c_size_t()
        self._check(self.lib.eqp_debug_inflate(self.h, addr, n, None, 0, C.byref(size)))
        out = np.empty(size.value, dtype=np.uint8)
        if size.value:
            self._check(self.lib.eqp_debug_inflate(self.h, addr, n, out.ctypes.data, out.size, C.byref(size)))
        return out.tobytes()

    def prefetch_chunk(self, t_sam, j_sam, g_sam):
        """Start the H2D copy of a later chunk's host buffers; the count_chunk call with the same buffers picks it up."""
        bufs = [_buf(x) for x in (t_sam, j_sam, g_sam)]
        self._check(self.lib.eqp_prefetch_chunk(self.h, bufs[0][0], bufs[0][1], bufs[1][0], bufs[1][1], bufs[2][0], bufs[2][1]))
        self._prefetch_keep.append(bufs)                                 # the host buffers must outlive the copy

    def device_vector(self):
        ptr, n = C.c_void_p(), C.c_int64()
        self._check(self.lib.eqp_device_vector(self.h, C.byref(ptr), C.byref(n)))
        return ptr.value, n.value

    def accumulate(self, device_ptr):
        """Add the last chunk's vector into a WIDE (2 * vector_length uint64) device accumulator."""
        self._check(self.lib.eqp_accumulate(self.h, C.c_void_p(int(device_ptr))))

    def comm_init(self, comm_id, n_ranks, rank):
        self._check(self.lib.eqp_comm_init(self.h, comm_id, n_ranks, rank))

    def reduce_sample(self, device_ptr, root=0):
        """In-place ncclReduce(sum) of a wide accumulator onto `root`, NCCL called inside libeqpcount.so."""
        self._check(self.lib.eqp_reduce_sample(self.h, C.c_void_p(int(device_ptr)), root))

    def debug_fetch(self, what, dtype):
        n = C.c_size_t()
        self._check(self.lib.eqp_debug_fetch(self.h, what.encode(), None, 0, C.byref(n)))
        out = np.zeros(n.value // np.dtype(dtype).itemsize, dtype=dtype)
        if n.value:
            self._check(self.lib.eqp_debug_fetch(self.h, what.encode(), out.ctypes.data, out.nbytes, C.byref(n)))
        return out

    def bucket_names(self):
        return self.debug_fetch("anno:buckets", np.uint8).tobytes().decode().split("\n")[:-1]

    def row_table(self):
        return self.debug_fetch("anno:rows", np.int32).reshape(-1, 2)

    def debug_decode(self, sam, misalign=0):
        """Run only the tokenizer kernel on one SAM text: (rows, err_code, err_detail); rows as `eqp_emu decode` prints them."""
        addr, n, keep = _buf(sam)
        text, tlen, ec, ed = C.c_char_p(), C.c_size_t(), C.c_uint32(), C.c_uint32()
        self._check(self.lib.eqp_debug_decode(self.h, addr, n, misalign, C.byref(text), C.byref(tlen), C.byref(ec), C.byref(ed)))
        rows = C.string_at(text, tlen.value).decode().splitlines() if tlen.value else []
        return rows, ec.value, ed.value

    def tokenize_noalloc(self, dev_sam_ptr, length, max_lines, dev_out_ptr, dev_out_bytes):
        """eqp_tokenize_noalloc: device text in, decoded-record columns in the caller's device buffer out; performs no allocation on the
        device.  Returns the RecordColumns (offsets, n_lines, err_code, err_detail)."""
        cols = RecordColumns()
        self._check(self.lib.eqp_tokenize_noalloc(self.h, C.c_void_p(dev_sam_ptr), length, max_lines, C.c_void_p(dev_out_ptr), dev_out_bytes, C.byref(cols)))
        return cols

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def close(self):
        if self.h:
            self.lib.eqp_destroy(self.h)
            self.h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def comm_id():
    """ncclGetUniqueId through the library: 128 bytes for rank 0 to ship to the other ranks."""
    lib = load_library()
    buf = C.create_string_buffer(COMM_ID_BYTES)
    rc = lib.eqp_comm_id(buf)
    if rc:
        raise EqpError(rc, (lib.eqp_last_error(None) or b"").decode())
    return buf.raw


def reduce_sample_local(contexts, device_ptrs, root=0):
    """One process, several GPUs: grouped ncclReduce over this process's contexts (eqp_reduce_sample_local)."""
    lib = load_library()
    n = len(contexts)
    hs = (C.c_void_p * n)(*[c.h for c in contexts])
    ps = (C.c_void_p * n)(*[C.c_void_p(int(p)) for p in device_ptrs])
    rc = lib.eqp_reduce_sample_local(hs, n, ps, root)
    if rc:
        raise EqpError(rc, (lib.eqp_last_error(contexts[0].h) or b"").decode())


def wide_to_double(wide):
    """Host copy of a wide accumulator (2n uint64: fraction limbs, then integer limbs) -> n float64 (eqp_wide_to_double)."""
    w = np.ascontiguousarray(wide).view(np.uint64)
    n = w.size // 2
    out = np.zeros(n, dtype=np.float64)
    load_library().eqp_wide_to_double(w.ctypes.data, n, out.ctypes.data)
    return out
