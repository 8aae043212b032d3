/* eqp_count.h — C ABI of libeqpcount.so: the B200 replacement for EQP-cluster's compute-counts hot path.
 *
 * The reference has no FFI for this path; its seam is "process + files" (SURVEY.md §8b):
 *   exon-pipeline-scripts/bin/process-scripts/compute-counts.sh:733-1037 spawns
 *     ComputeReadWeightsSam x3, CombineReadWeightFiles x2, CombineSamFiles x2, ConvertSamBed,
 *     bedtools intersect, ComputeReadWeights and ComputeCounts x3
 *   (exon-pipeline-scripts/java/source/{sam-source,bed-source,util-source}).
 * One eqp_count_chunk() call replaces that whole process tree for one chunk; the Java side keeps
 * option parsing, inflate, the .cnt writers and the log lines (INTEGRATION.md shows the JNI / Panama stub).
 *
 * All functions return 0 on success or a negative EQP_ERR_* code; eqp_last_error() gives the message
 * the Java main prints as "ERROR: <msg>" before System.exit(1) (ComputeCounts.java:996-999).
 * There is NO CPU fallback: without a CUDA device every compute entry point fails with EQP_ERR_CUDA.
 * An eqp_ctx is bound to one GPU, is not thread-safe and runs one chunk at a time.
 */
#ifndef EQP_COUNT_H
#define EQP_COUNT_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define EQP_ABI_VERSION 2

enum {
  EQP_OK = 0,
  EQP_ERR_ARG = -1,          /* bad argument / call order */
  EQP_ERR_CUDA = -2,         /* no device, driver or allocation failure */
  EQP_ERR_IO = -3,           /* annotation file missing or malformed */
  EQP_ERR_FORMAT = -4,       /* input violates a check the reference aborts on */
  EQP_ERR_UNSUPPORTED = -5,  /* valid for the reference, not handled on the device (DESIGN.md "out of scope"): one corner, an alignment
                                with both a tr- exon record and a transcript/genome exon length difference */
  EQP_ERR_CAPACITY = -6,     /* a per-fragment device table overflowed, or a chunk has 2^26 or more fragments */
  EQP_ERR_NCCL = -7          /* NCCL missing (libnccl.so.2 is loaded at run time) or a collective failed */
};

typedef struct eqp_ctx eqp_ctx;

/* Annotation inputs: the files compute-counts.sh reads from <project>/exon-pipeline-files
 * (compute-counts.sh:584-603, compute-intersection-bed-file.sh:551-572, filter-junction-map-files.sh:207-236).
 * Plain (uncompressed) paths; the Java caller inflates *.gz maps. */
typedef struct eqp_annotation_files {
  const char* combined_exons_bed;              /* <base>_<R>_combined_exons.bed  (bedtools -b file) */
  const char* transcript_exons_bed;            /* <base>_transcript_exons.bed    (ConvertSamBed -T) */
  const char* exon_gene_map;                   /* ComputeCounts -g -m */
  const char* exon_exon_map;                   /* ComputeCounts -e -m */
  const char* exon_junction_map;               /* unfiltered; filtered per chunk on the device */
  const char* exon_junction_transcript_map;
  const char* junction_junction_coordinate_map;
  const char* equivalent_junctions_map;
  const char* genome_fai;                      /* ComputeReadWeightsSam -c */
} eqp_annotation_files;

/* The flags process-samples.sh passes (process-samples.sh:1480-1481) and compute-counts.sh forwards
 * (compute-counts.sh:932-935,959-960,1034-1035); replaces the UtilLib statics (UtilLib.java:101-226). */
typedef struct eqp_params {
  int32_t paired_end;            /* informational: layout is read from the SAM flags */
  int32_t stranded;              /* -s <direction>: 0 none, 1 forward, 2 backward (compute-counts.sh:372-387; ConvertSamBed -S, bedtools -s, ComputeCounts -s) */
  int32_t gene_overlap;          /* 1 */
  int32_t exon_overlap;          /* -E, default 5  (compute-counts.sh:136) */
  int32_t junction_overlap;      /* -J, default 8  (compute-counts.sh:137,391-401) */
  double  read_weight_threshold; /* -W, default 0.01 (compute-counts.sh:139) */
  int32_t unweighted;            /* --unweighted -> ComputeCounts -u */
  int32_t unambiguous;           /* --unambig    -> ComputeCounts -U (gene and exon runs) */
  int32_t containment;           /* -C (gene run) */
  int32_t nonsplice_conforming;  /* -N (gene and exon runs) */
  double  distance_slack;        /* CombineSamFiles -s, default 0.25 */
  int32_t modes;                 /* bit0 gene, bit1 exon, bit2 junction */
  int32_t genomic_input;         /* 1: t_sam is the name-sorted SAM of a spliced GENOME aligner (compute-genomic-counts.sh:1040-1058,1126-1152): N
                                    CIGARs, HI:i hit indices (SamRecord.java:658-660), weights = that SAM's own alignment counts; j_sam / g_sam
                                    empty, modes 1 | 2; the context's combined_exons_bed is <base>_genome_exons.bed (or any B file with genome rows) */
} eqp_params;

/* Results of one chunk.  Vectors are indexed by the rank of the feature id in the map's sorted id set
 * (ComputeCounts.java:877-897 prints in TreeSet order); eqp_feature_id() gives the id strings.
 * `*_fixed` are the exact accumulators: value = fixed / 2^EQP_FIXED_SHIFT.  They are what eqp_reduce adds. */
#define EQP_FIXED_SHIFT 38
typedef struct eqp_counts {
  int64_t n_gene, n_exon, n_junction;        /* distinct count-object ids per map (column 2, or column 3 of a 3-column map) */
  const double* gene; const double* exon; const double* junction;
  const uint64_t* gene_fixed; const uint64_t* exon_fixed; const uint64_t* junction_fixed;
  const uint8_t* junction_active;            /* n_junction_act flags: 1 = the column-2 id survives filter-junction-map-files.sh for this chunk */
  int64_t n_junction_act;                    /* == n_junction for a 2-column junction map */
  int64_t number_mapped_reads;               /* NUMBER_MAPPED_READS=   (ConvertSamBed.java:266) */
  int64_t number_expressed_reads;            /* NUMBER_EXPRESSED_READS= (ComputeCounts.java:991-993, gene run) */
  int64_t num_weight_objects[3];             /* per mode: gene, exon, junction (ComputeCounts.java:986-989) */
  int64_t num_weight_objects_included[3];
  int64_t num_weight_objects_excluded[3];
  int64_t n_records[3];                      /* SAM records decoded per input: transcript, junction, genome */
  double  sum_weights[3];                    /* "Sum of weights contributing to count objects" per run (ComputeCounts.java:989) */
  double  device_ms;                         /* first kernel launch -> counts ready, CUDA events */
  double  dominant_kernel_ms;                /* the SAM decode kernel, summed over the three inputs */
  int64_t kernel_launches;
} eqp_counts;

int  eqp_version(void);
void eqp_default_params(eqp_params* p);
int  eqp_create(int device, eqp_ctx** out);
int  eqp_load_annotation(eqp_ctx* ctx, const eqp_annotation_files* files);
/* Device-resident bytes of the annotation index (the |annotation index| term of the algorithmic bytes). */
int64_t eqp_annotation_bytes(eqp_ctx* ctx);

/* Count one chunk.  t_sam/j_sam/g_sam are UNCOMPRESSED SAM text (what `zcat` feeds the Java mains,
 * compute-intersection-bed-file.sh:644-662).  *_on_device != 0 means the pointer is device memory already
 * resident in HBM; otherwise it is host memory (pinned or pageable) and the call copies it.
 * `out` stays valid until the next eqp_count_chunk / eqp_destroy on this ctx. */
int  eqp_count_chunk(eqp_ctx* ctx, const eqp_params* params,
                     const uint8_t* t_sam, size_t t_len, const uint8_t* j_sam, size_t j_len,
                     const uint8_t* g_sam, size_t g_len, int on_device, eqp_counts* out);

/* Double-buffered ingest: start the host-to-device copy of a LATER chunk's three SAM buffers now, on the
 * copy stream, so that it overlaps the counting of the chunk before it.  Call it before the eqp_count_chunk of the previous
 * chunk (that call blocks); the eqp_count_chunk(on_device = 0) that follows with the same pointers and lengths picks the copy up.
 * At most two chunks may be announced; the host buffers must stay valid (and should be page-locked) until their chunk has been
 * counted.  An announcement that does not match the next eqp_count_chunk call is dropped.  This replaces the
 * `zcat file | java ...` hand-over of compute-intersection-bed-file.sh:644-662 for callers that hold several chunks in memory. */
int  eqp_prefetch_chunk(eqp_ctx* ctx, const uint8_t* t_sam, size_t t_len, const uint8_t* j_sam, size_t j_len,
                        const uint8_t* g_sam, size_t g_len);

/* Feature ids, mode = 0 gene, 1 exon, 2 junction. */
int64_t     eqp_feature_count(eqp_ctx* ctx, int mode);
const char* eqp_feature_id(eqp_ctx* ctx, int mode, int64_t index);

/* Rows of a count file in the reference's print order (ComputeCounts.java:875-950).  2-column map: one row per id, sorted
 * (TreeSet order).  3-column map (Exon.java:107-125): one row per map line, printed id = column 2, value = the count of
 * the column-3 id; the junction map's lines in the order filter-junction-map-files.sh:298-299 leaves them in.  A junction row
 * is part of this chunk's filtered map iff `always` (it is in exon_junction_transcript.map) or junction_active[act] is set.
 * eqp_map_columns tells the writer which zero to print: "0" (2 columns) or "0.0" (3 columns). */
typedef struct eqp_output_row { const char* id; int64_t feature; int64_t act; int32_t always; int32_t reserved; } eqp_output_row;
int     eqp_map_columns(eqp_ctx* ctx, int mode);
int64_t eqp_output_row_count(eqp_ctx* ctx, int mode);
int     eqp_get_output_row(eqp_ctx* ctx, int mode, int64_t index, eqp_output_row* out);

/* ---- sample level: the sum over a sample's chunks (replaces `combineCounts.py -A`, combineCounts.py:199-210) ----
 * A chunk vector is the concatenated gene|exon|junction accumulators, n = n_gene + n_exon + n_junction uint64 in 2^-38
 * units.  One chunk cannot overflow them (a fragment adds at most 1.0 to a feature and eqp_count_chunk refuses chunks of
 * 2^26 or more fragments), but a sample can: 2^26 = 67.1 M fragments' worth on ONE feature wraps a u64.  A sample
 * accumulator is therefore WIDE: 2n uint64, element i = fraction limb (sum of the chunks' low 38 bits), element n + i =
 * integer limb (sum of the chunks' bits 38..63); value_i = acc[n + i] + acc[i] / 2^38.  Plain integer sums, so NCCL's
 * reduction order cannot change a bit, and the limbs hold 2^26 chunks. */
int  eqp_device_vector(eqp_ctx* ctx, uint64_t** dev_ptr, int64_t* n_elements);       /* the last chunk's vector (n elements) */
/* Add the last chunk's vector into a caller-owned, zero-initialised WIDE device accumulator (2n elements).  The add runs on
 * the context's stream after the chunk's counting; work the caller queued on the legacy default stream (torch's default) is
 * ordered before it, work on any other stream must be synchronised by the caller first.  Returns when the add has finished. */
int  eqp_accumulate(eqp_ctx* ctx, uint64_t* dev_accumulator_wide);
/* Host helper: wide accumulator (host copy, 2n elements) -> n doubles. */
void eqp_wide_to_double(const uint64_t* wide, int64_t n, double* out);

/* The reduce itself, with NCCL called inside the library (libnccl.so.2, loaded at run time), so that a host that cannot
 * call torch.distributed — the Java mains — gets the combine-counts replacement.  One process per GPU:
 *   rank 0: eqp_comm_id(id) -> ship the 128 bytes to the other ranks (file, socket, MPI ...);
 *   every rank: eqp_comm_init(ctx, id, n_ranks, rank) once; then per sample eqp_reduce_sample(ctx, acc_wide, root):
 *   in-place ncclReduce(sum, uint64) of the 2n elements onto `root`, ordered after this context's pending work.
 * One process driving several GPUs (threads of one JVM): eqp_reduce_sample_local over that process's contexts. */
#define EQP_COMM_ID_BYTES 128
int  eqp_comm_id(void* id_out);
int  eqp_comm_init(eqp_ctx* ctx, const void* id, int n_ranks, int rank);
int  eqp_reduce_sample(eqp_ctx* ctx, uint64_t* dev_accumulator_wide, int root);
int  eqp_reduce_sample_local(eqp_ctx** ctxs, int n, uint64_t** dev_accumulators_wide, int root);

/* ---- streaming ingest: feed the three SAM texts piecewise while they are inflated (compute-intersection-bed-file.sh:644-662
 * pipes `zcat` into the Java mains; a host need not hold 45 GB of text).  begin -> any number of feed calls per input, in
 * text order per input, inputs interleaved freely -> end counts what was fed.  feed copies `len` host bytes through a
 * page-locked staging ring and returns as soon as `buf` may be reused; the device buffers grow as needed (`*_hint` bytes
 * are reserved up front, 0 = unknown). */
int  eqp_chunk_begin(eqp_ctx* ctx, size_t t_hint, size_t j_hint, size_t g_hint);
int  eqp_chunk_feed(eqp_ctx* ctx, int input /* 0 transcript, 1 junction, 2 genome */, const uint8_t* buf, size_t len);
int  eqp_chunk_end(eqp_ctx* ctx, const eqp_params* params, eqp_counts* out);

/* ---- `.sam.gz` input inflated on the device (replaces the `zcat` of compute-intersection-bed-file.sh:644-662 and the
 * GZIPInputStream of the Java mains).  `gz` is ONE complete gzip member in host memory (what `gzip` / `bowtie2 | gzip` write): the compressed
 * bytes cross PCIe, a speculative block-parallel DEFLATE decoder (eqp_inflate.h) rebuilds the text in HBM and appends it to the input's feed
 * buffer.  eqp_chunk_feed_gz belongs to the begin / feed / end family (one member per call, may be mixed with eqp_chunk_feed);
 * eqp_count_chunk_gz = begin + three members (their copies overlap the inflate of the one before) + end.  Checked: header, block chain
 * (every accepted block start is landed on exactly by the decoder in front of it), ISIZE and the trailer's CRC-32 (computed on the device
 * over the inflated text; EQP_INFLATE_NO_CRC=1 skips it).
 * EQP_ERR_UNSUPPORTED: a valid file this decoder does not take — several members / BGZF, or a large member without dynamic blocks that
 * start with text (stored blocks, binary content) — inflate such a file on the host and use eqp_chunk_feed.  EQP_ERR_FORMAT: not gzip / corrupt.
 * Tunables (environment): EQP_INFLATE_SEG (bytes of compressed data per decoder thread, default 16384), EQP_INFLATE_SCRATCH_GB
 * (16-bit symbol scratch, default 24, at most half of the free device memory; a larger text is written in several waves). */
typedef struct eqp_inflate_info {
  uint64_t gz_bytes, text_bytes;
  uint32_t segments, chain, bogus_starts, waves;   /* segments cut, segments on the verified chain, block starts proven wrong, scratch refills */
  float    search_ms, decode_ms, resolve_ms, emit_ms;
} eqp_inflate_info;
int  eqp_chunk_feed_gz(eqp_ctx* ctx, int input, const uint8_t* gz, size_t len);
int  eqp_count_chunk_gz(eqp_ctx* ctx, const eqp_params* params, const uint8_t* t_gz, size_t t_len, const uint8_t* j_gz, size_t j_len,
                        const uint8_t* g_gz, size_t g_len, eqp_counts* out);
int  eqp_last_inflate(eqp_ctx* ctx, int input, eqp_inflate_info* out);   /* what the last inflate of that input did */
/* Test tap: inflate one member on the device and copy the text to dst (dst == NULL: only *n_text is set). */
int  eqp_debug_inflate(eqp_ctx* ctx, const uint8_t* gz, size_t len, uint8_t* dst, size_t dst_bytes, size_t* n_text);

/* Debug/parity taps: copy a device intermediate of the last chunk to the host.
 * what: "weights:<T|J|G|TJ|final>", "selection", "blocks", "hits", "a20", "meta"; "anno:buckets" (newline-separated reference names in
 * bucket order) and "anno:rows" (int32 start,end per B row) decode the bucket / row indices.  Layouts are documented in DESIGN.md. */
int  eqp_debug_fetch(eqp_ctx* ctx, const char* what, void* dst, size_t dst_bytes, size_t* n_written);

/* Test tap for the a1/a2 tokenizer kernel alone: decodes one SAM text, placed `misalign` (0-15) bytes after a 16-byte boundary in
 * device memory, and formats one row per text line: "<line>\tREC\t<frag>\t<flag>\t<rname>\t<pos>\t<pnext>\t<nm>\t<cigar>\t<seqlen>\t<rf>",
 * "<line>\tSKIP" (header / empty line), or "<line>\tLINE\t<offset>" for every line when the device raised an error (err_code != 0,
 * the codes of SamRecord.java:182-186,246-257).  *text stays valid until the next call on this ctx. */
int  eqp_debug_decode(eqp_ctx* ctx, const uint8_t* sam, size_t len, int misalign, const char** text, size_t* text_len,
                      uint32_t* err_code, uint32_t* err_detail);

/* ---- allocation-free tokenizer call.  This entry point has NO counterpart in the reference; it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  It runs the a1/a2 tokenizer kernel (the one eqp_count_chunk / eqp_debug_decode launch) on ONE SAM text
 * that is already in device memory and writes the decoded records into a device buffer the CALLER owns.  The kernel's scratch
 * (tile look-back states, tile ticket, newline total, error slot) lives in a file-scope device array of the library
 * (EQP_NOALLOC_SCRATCH_BYTES per GPU that loads the library, held until the process exits), so the call performs no cudaMalloc /
 * cudaFree / pool allocation.  Limits: len <= EQP_NOALLOC_MAX_TEXT_BYTES (larger texts are refused with EQP_ERR_ARG);
 * dev_out_bytes >= EQP_NOALLOC_OUT_BYTES(max_lines); a text with more than max_lines lines is refused with EQP_ERR_ARG
 * (cols->n_lines then tells how many it has).  HI:i: tags are rejected as in the default flow (err_code 2).
 * All calls on one GPU share the scratch array: the call clears it, runs on the context's stream and returns when the kernel
 * has finished, so calls in sequence are safe; calls that OVERLAP on one GPU (two contexts / two host threads) race.
 * cols receives the byte offset of every column inside dev_out (element types as named; a column has max_lines entries,
 * line_off max_lines + 1), the line count and the device error slot (0 = none; records are undefined after an error).
 * Columns other than frag / rf are only written for record lines (rf bit 0). */
#define EQP_NOALLOC_MAX_TEXT_BYTES ((size_t)128 << 20)
#define EQP_NOALLOC_SCRATCH_BYTES  36864
#define EQP_NOALLOC_OUT_BYTES(max_lines) ((size_t)(max_lines) * 41 + 8 + 13 * 256)
typedef struct eqp_record_columns {
  uint64_t line_off;                                    /* uint64: text offset of each line; entry i + 1 minus 1 = end of line i */
  uint64_t frag, ref, pos, rnext, pnext;                /* uint32 fragment id; int32 reference index (-1 "*"), POS, RNEXT index, PNEXT */
  uint64_t flag, nm, cig_off, cig_len, seq_len, cig_m;  /* uint16 FLAG; int16 NM (-1 none); uint16 CIGAR offset in line / length; SEQ length; n of "<n>M" or 0xFFFF */
  uint64_t rf;                                          /* uint8: bit0 record, bit1 mapped, bit2 has mate, bits3-4 read index */
  uint64_t n_lines;
  uint32_t err_code, err_detail;
} eqp_record_columns;
int  eqp_tokenize_noalloc(eqp_ctx* ctx, const uint8_t* dev_sam, size_t len, uint64_t max_lines,
                          void* dev_out, size_t dev_out_bytes, eqp_record_columns* cols);

const char* eqp_last_error(eqp_ctx* ctx);
void eqp_destroy(eqp_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif
