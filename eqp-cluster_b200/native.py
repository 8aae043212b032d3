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
                ("distance_slack", C.c_double), ("modes", C.c_int32), ("reserved", C.c_int32)]


class Counts(C.Structure):
    _fields_ = [("n_gene", C.c_int64), ("n_exon", C.c_int64), ("n_junction", C.c_int64),
                ("gene", C.POINTER(C.c_double)), ("exon", C.POINTER(C.c_double)), ("junction", C.POINTER(C.c_double)),
                ("gene_fixed", C.POINTER(C.c_uint64)), ("exon_fixed", C.POINTER(C.c_uint64)), ("junction_fixed", C.POINTER(C.c_uint64)),
                ("junction_active", C.POINTER(C.c_uint8)),
                ("number_mapped_reads", C.c_int64), ("number_expressed_reads", C.c_int64),
                ("num_weight_objects", C.c_int64 * 3), ("num_weight_objects_included", C.c_int64 * 3),
                ("num_weight_objects_excluded", C.c_int64 * 3), ("n_records", C.c_int64 * 3),
                ("device_ms", C.c_double), ("dominant_kernel_ms", C.c_double), ("kernel_launches", C.c_int64)]


EXPORTS = ("eqp_version", "eqp_default_params", "eqp_create", "eqp_load_annotation", "eqp_count_chunk", "eqp_prefetch_chunk", "eqp_feature_count",
           "eqp_feature_id", "eqp_device_vector", "eqp_accumulate", "eqp_debug_fetch", "eqp_debug_decode", "eqp_last_error", "eqp_destroy")

_lib = None


def load_library(path=None):
    """dlopen libeqpcount.so (building it in-tree if needed) and declare the prototypes of include/eqp_count.h."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or _build.LIB_PATH
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
    lib.eqp_debug_fetch.argtypes = [vp, C.c_char_p, vp, C.c_size_t, C.POINTER(C.c_size_t)]
    lib.eqp_debug_decode.argtypes = [vp, vp, C.c_size_t, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    lib.eqp_last_error.argtypes = [vp]; lib.eqp_last_error.restype = C.c_char_p
    lib.eqp_destroy.argtypes = [vp]; lib.eqp_destroy.restype = None
    lib.eqp_annotation_bytes.argtypes = [vp]; lib.eqp_annotation_bytes.restype = C.c_int64
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

    def __init__(self, c, feature_ids):
        def arr(ptr, n, dt):
            return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dt, copy=True) if n else np.zeros(0, dt)
        n = (c.n_gene, c.n_exon, c.n_junction)
        self.counts = {m: arr(getattr(c, m), n[i], np.float64) for i, m in enumerate(MODES)}
        self.fixed = {m: arr(getattr(c, m + "_fixed"), n[i], np.uint64) for i, m in enumerate(MODES)}
        self.junction_active = arr(c.junction_active, n[2], np.uint8)
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


class CountContext:
    """One eqp_ctx: bound to one GPU, one chunk in flight (include/eqp_count.h)."""

    def __init__(self, device=0):
        self.lib = load_library()
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
        self.feature_ids = {}
        for i, m in enumerate(MODES):
            n = self.lib.eqp_feature_count(self.h, i)
            self.feature_ids[m] = [self.lib.eqp_feature_id(self.h, i, k).decode() for k in range(n)]
        return self

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
        return ChunkCounts(c, self.feature_ids) if fetch else c

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
        self._check(self.lib.eqp_accumulate(self.h, C.c_void_p(int(device_ptr))))

    def debug_fetch(self, what, dtype):
        n = C.c_size_t()
        self._check(self.lib.eqp_debug_fetch(self.h, what.encode(), None, 0, C.byref(n)))
        out = np.zeros(n.value // np.dtype(dtype).itemsize, dtype=dtype)
        if n.value:
            self._check(self.lib.eqp_debug_fetch(self.h, what.encode(), out.ctypes.data, out.nbytes, C.byref(n)))
        return out

    def debug_decode(self, sam, misalign=0):
        """Run only the tokenizer kernel on one SAM text: (rows, err_code, err_detail); rows as `eqp_emu decode` prints them."""
        addr, n, keep = _buf(sam)
        text, tlen, ec, ed = C.c_char_p(), C.c_size_t(), C.c_uint32(), C.c_uint32()
        self._check(self.lib.eqp_debug_decode(self.h, addr, n, misalign, C.byref(text), C.byref(tlen), C.byref(ec), C.byref(ed)))
        rows = C.string_at(text, tlen.value).decode().splitlines() if tlen.value else []
        return rows, ec.value, ed.value

    def close(self):
        if self.h:
            self.lib.eqp_destroy(self.h)
            self.h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
