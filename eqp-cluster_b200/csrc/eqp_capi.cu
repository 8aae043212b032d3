// libeqpcount.so — the C ABI declared in include/eqp_count.h, implemented on the CUDA pipeline.
// There is no CPU path in this library: every compute entry point needs a CUDA device and fails with EQP_ERR_CUDA without one.
#include <cstdio>
#include <cstring>
#include <chrono>
#include <deque>
#include <memory>
#include <thread>
#include "../../include/eqp_count.h"
#include "eqp_cuda_backend.cuh"
#include "eqp_nccl.h"
#include "eqp_inflate.cuh"

using namespace eqpb;

namespace {

template <class T> struct DevArray {
  T* p = nullptr; size_t n = 0;
  void upload(const std::vector<T>& v) { upload(v.data(), v.size()); }
  void upload(const T* src, size_t count) {
    release(); n = count;
    EQP_CUDA(cudaMalloc((void**)&p, (count ? count : 1) * sizeof(T) + 16));
    if (count) EQP_CUDA(cudaMemcpy(p, src, count * sizeof(T), cudaMemcpyHostToDevice));
  }
  void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
  ~DevArray() { release(); }
};

struct AnnoDevice {
  DevArray<DictEntry> dict; DevArray<char> strPool; DevArray<uint32_t> refStrOff; DevArray<uint16_t> refStrLen;
  DevArray<uint8_t> refFlags; DevArray<uint32_t> refVerOff; DevArray<int32_t> refVerBucket; DevArray<uint32_t> refActOff; DevArray<int32_t> refAct;
  DevArray<uint32_t> bktRowOff; DevArray<uint8_t> bktFlags; DevArray<int32_t> bktRefId; DevArray<uint32_t> featAct; DevArray<RowA> row;
  DevArray<ExonA> ex; DevArray<ExonFeat> exf; DevArray<uint32_t> feat[3];
  size_t bytes = 0;
  AnnoView upload(const AnnoHost& H) {
    AnnoView v = H.view();
    dict.upload(H.dict); strPool.upload(H.strPool.data(), H.strPool.size()); refStrOff.upload(H.refStrOff);
    refStrLen.upload(H.refStrLen); refFlags.upload(H.refFlags); refVerOff.upload(H.refVerOff); refVerBucket.upload(H.refVerBucket);
    refActOff.upload(H.refActOff); refAct.upload(H.refAct); bktRowOff.upload(H.bktRowOff); bktFlags.upload(H.bktFlags); bktRefId.upload(H.bktRefId); featAct.upload(H.featAct);
    row.upload(H.rowA); ex.upload(H.exA); exf.upload(H.exF);
    v.dict = dict.p; v.strPool = strPool.p; v.refStrOff = refStrOff.p; v.refStrLen = refStrLen.p; v.refFlags = refFlags.p;
    v.refVerOff = refVerOff.p; v.refVerBucket = refVerBucket.p; v.refActOff = refActOff.p; v.refAct = refAct.p;
    v.bktRowOff = bktRowOff.p; v.bktFlags = bktFlags.p; v.bktRefId = bktRefId.p; v.featAct = H.featAct.empty() ? nullptr : featAct.p; v.row = row.p;
    v.ex = ex.p; v.exf = exf.p;
    for (int m = 0; m < 3; m++) { feat[m].upload(H.feat[m]); v.feat[m] = feat[m].p; }
    bytes = H.dict.size() * 16 + H.strPool.size() + H.refStrOff.size() * 7 + H.refVerOff.size() * 4 + H.refVerBucket.size() * 4 + H.refActOff.size() * 4 + H.refAct.size() * 4 +
            H.bktRowOff.size() * 9 + H.rowA.size() * 16 + H.exA.size() * 16 + H.featAct.size() * 4;
    bytes += H.exF.size() * 16; for (int m = 0; m < 3; m++) bytes += H.feat[m].size() * 4;
    return v;
  }
};

// chunk vector (n elements, 2^-38 units) into a wide sample accumulator (2n elements): low 38 bits into the fraction limb, the rest into
// the integer limb (include/eqp_count.h, "sample level")
__global__ void k_add_wide(unsigned long long* __restrict__ dst, const unsigned long long* __restrict__ src, int64_t n) {
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    unsigned long long v = src[i];
    if (v) { dst[i] += v & ((1ull << EQP_FIXED_SHIFT_D) - 1); dst[n + i] += v >> EQP_FIXED_SHIFT_D; }
  }
}

const char* derr_text(uint32_t code) {
  switch (code) {
    case DERR_QNAME: return "read name is not of the form Fnnnnnnnnn[/1|/2] (SamReader.java:157)";
    case DERR_FIELDS: return "SAM record does not have 11 mandatory fields (+ tags) or carries an HI tag (SamRecord.java:251-257)";
    case DERR_SEQ_QUAL: return "sequence and quality lengths differ (SamRecord.java:246-249)";
    case DERR_READ_INDEX: return "read index of the flag differs from the read name (SamRecord.java:182-186)";
    case DERR_ORDER: return "SAM records are not ordered by fragment id (CombineSamFiles.java:191-193)";
    case DERR_NULL_WEIGHT: return "fragment has only non-whitelisted genome alignments: the reference prints 'null' and aborts";
    case DERR_CIGAR_OP: return "unknown CIGAR operation (SamRecord.java:419-421)";
    case DERR_CIGAR_LEN: return "read length differs from the CIGAR's M+I+S (SamRecord.java:463-466)";
    case DERR_MATE_REF: return "Differing referenceName mate found (SamRecord.java:567-571)";
    case DERR_TR_LENDIFF: return "an alignment has both a tr- exon record and a transcript/genome exon length difference: its coordinates differ between the junction run and the others (GenomicAlignment.java:233-234, BedRecord.java:417-452), not handled on the device";
    case DERR_ACC_OVERFLOW: return "chunk has 2^26 or more fragments: a 64-bit fixed-point accumulator could overflow";
    case DERR_CAPACITY: return "a per-fragment device table overflowed";
    case DERR_GENOME_COORD: return "genomic record does not agree with its exon (BedRecord.java:669-673,683-687)";
    case DERR_INCONSISTENT: return "inconsistent records within one alignment (GenomicAlignment.java:123-148,245-300)";
    case DERR_BAD_JUNCTION_ID: return "junction reference id is not <gene>-junction-<i>-<j>:<l>:<r> (createExtendedJunctionIds.py:93-114)";
    case DERR_STRANDED: return "strand-specific direction must be 0 (none), 1 (forward) or 2 (backward)";
    case DERR_INT: return "integer SAM field does not parse";
    default: return "unknown device error";
  }
}
int derr_class(uint32_t code) {
  switch (code) {
    case DERR_TR_LENDIFF: return EQP_ERR_UNSUPPORTED;
    case DERR_STRANDED: return EQP_ERR_ARG;
    case DERR_CAPACITY: case DERR_ACC_OVERFLOW: return EQP_ERR_CAPACITY;
    default: return EQP_ERR_FORMAT;
  }
}

}  // namespace

struct eqp_ctx {
  int device = 0; int nSm = 0; size_t totalMem = 0; cudaStream_t stream = nullptr;
  AnnoHost H; AnnoDevice D; AnnoView A{}; bool annoLoaded = false;
  std::unique_ptr<CudaBackend> be; std::unique_ptr<Pipeline<CudaBackend>> pl;
  ChunkResult R{}; bool haveChunk = false;
  unsigned long long* vec = nullptr; int64_t vecLen = 0;           // gene|exon|junction fixed-point vector of the last chunk
  cudaEvent_t ev0 = nullptr, ev1 = nullptr; cudaStream_t copyStream = nullptr;
  std::string err, debugText;
  std::vector<void*> pinned;
  // Host chunks whose H2D copy was started ahead of their eqp_count_chunk call (eqp_prefetch_chunk), oldest first; `current` = the
  // device text (and copy-done events) of the chunk being counted.
  struct HostChunk { const uint8_t* src[3]; size_t len[3]; char* dev[3]; cudaEvent_t ready[3]; };
  std::deque<HostChunk> prefetched; HostChunk current{};
  void retire(HostChunk& c, cudaStream_t s) {
    for (int f = 0; f < 3; f++) { if (c.dev[f]) cudaFreeAsync(c.dev[f], s); if (c.ready[f]) cudaEventDestroy(c.ready[f]); }
    c = HostChunk{};
  }
  void start_copy(HostChunk& c) {                                   // allocation + copy + event per input, all ordered on the copy stream
    if (!copyStream) EQP_CUDA(cudaStreamCreateWithFlags(&copyStream, cudaStreamNonBlocking));
    for (int f = 0; f < 3; f++) {
      void* d = nullptr; EQP_CUDA(cudaMallocAsync(&d, c.len[f] + 16, copyStream)); c.dev[f] = (char*)d;
      if (c.len[f]) EQP_CUDA(cudaMemcpyAsync(d, c.src[f], c.len[f], cudaMemcpyHostToDevice, copyStream));
      EQP_CUDA(cudaEventCreateWithFlags(&c.ready[f], cudaEventDisableTiming));
      EQP_CUDA(cudaEventRecord(c.ready[f], copyStream));
    }
  }
  // streaming ingest (eqp_chunk_begin/feed/end): growing device text per input + a page-locked staging ring
  static const int kStages = 4; static const size_t kStageBytes = 16u << 20;
  struct Feed { bool active = false; char* dev[3] = {nullptr, nullptr, nullptr}; size_t len[3] = {0, 0, 0}, cap[3] = {0, 0, 0}; } feed;
  uint8_t* stage[kStages] = {nullptr, nullptr, nullptr, nullptr}; cudaEvent_t stageFree[kStages] = {nullptr, nullptr, nullptr, nullptr}; int nextStage = 0;
  void feed_reserve(int f, size_t want) { feed_reserve_on(f, want, feed.len[f], copyStream); }   // grow on the copy stream: earlier pieces are copied over in stream order
  void feed_reserve_on(int f, size_t want, size_t keep, cudaStream_t st) {                       // keep: bytes of the old buffer that hold text
    if (want <= feed.cap[f]) return;
    size_t cap = feed.cap[f] ? feed.cap[f] * 2 : (size_t)64 << 20; if (cap < want) cap = want;
    void* d = nullptr; EQP_CUDA(cudaMallocAsync(&d, cap + 16, st));
    if (keep) EQP_CUDA(cudaMemcpyAsync(d, feed.dev[f], keep, cudaMemcpyDeviceToDevice, st));
    if (feed.dev[f]) EQP_CUDA(cudaFreeAsync(feed.dev[f], st));
    feed.dev[f] = (char*)d; feed.cap[f] = cap;
  }
  void feed_drop() { for (int f = 0; f < 3; f++) { if (feed.dev[f]) cudaFreeAsync(feed.dev[f], copyStream ? copyStream : stream); } feed = Feed(); }
  // device inflate (eqp_chunk_feed_gz / eqp_count_chunk_gz): compressed bytes travel on their own stream, the inflate runs on the copy stream
  cudaStream_t gzStream = nullptr, inflateStream[3] = {nullptr, nullptr, nullptr}; eqp_inflate_info lastInflate[3] = {};
  // NCCL communicator of this context (one rank per context)
  nccl_comm_t comm = nullptr; int commRanks = 0, commRank = -1;
  // device address of g_noalloc_scratch on this context's device (eqp_tokenize_noalloc), resolved in eqp_create
  unsigned long long* noallocScratch = nullptr; cudaError_t noallocRc = cudaErrorNotReady;
  void pin(void* p, size_t bytes) { if (bytes && cudaHostRegister(p, bytes, cudaHostRegisterDefault) == cudaSuccess) pinned.push_back(p); else cudaGetLastError(); }
  void unpin() { for (void* p : pinned) cudaHostUnregister(p); pinned.clear(); }
};

static thread_local std::string g_err;                              // errors raised before a ctx exists

// Scratch of eqp_tokenize_noalloc: the tokenizer's tile states [nTiles], newline total, tile ticket, and (last two words) the error
// slot + fragment id range.  One copy per GPU that loads the library.  In host code this NAME is only a stand-in for the symbol:
// the device address comes from cudaGetSymbolAddress (eqp_create, with the context's device current) and is what the kernel gets.
static const size_t kNoallocWords = EQP_NOALLOC_SCRATCH_BYTES / 8;
static __device__ __align__(256) unsigned long long g_noalloc_scratch[EQP_NOALLOC_SCRATCH_BYTES / 8];
static_assert((EQP_NOALLOC_MAX_TEXT_BYTES + 15) / TK_OWNED + 1 + 2 + 2 <= EQP_NOALLOC_SCRATCH_BYTES / 8, "scratch array holds the largest accepted text's tile states");

static int fail(eqp_ctx* ctx, int code, const std::string& msg) { if (ctx) ctx->err = msg; else g_err = msg; return code; }

extern "C" {

int eqp_version(void) { return EQP_ABI_VERSION; }

void eqp_default_params(eqp_params* p) {
  if (!p) return;
  memset(p, 0, sizeof *p);
  p->paired_end = 1; p->stranded = 0; p->gene_overlap = 1; p->exon_overlap = 5; p->junction_overlap = 8;
  p->read_weight_threshold = 0.01; p->distance_slack = 0.25; p->modes = 7;
}

int eqp_create(int device, eqp_ctx** out) {
  if (!out) return fail(nullptr, EQP_ERR_ARG, "eqp_create: out is null");
  *out = nullptr;
  int n = 0; cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) return fail(nullptr, EQP_ERR_CUDA, std::string("no CUDA device: ") + cudaGetErrorString(e) + " (this library has no CPU path)");
  if (device < 0 || device >= n) return fail(nullptr, EQP_ERR_ARG, "eqp_create: device index out of range");
  try {
    std::unique_ptr<eqp_ctx> c(new eqp_ctx); c->device = device;
    EQP_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop; EQP_CUDA(cudaGetDeviceProperties(&prop, device)); c->nSm = prop.multiProcessorCount; c->totalMem = prop.totalGlobalMem;
    // a blocking stream: ordered against the legacy default stream, on which a caller (torch) typically prepares device-resident
    // inputs and accumulators (include/eqp_count.h, eqp_accumulate)
    EQP_CUDA(cudaStreamCreate(&c->stream));
    cudaMemPool_t pool; EQP_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
    uint64_t keep = ~0ULL; EQP_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));   // keep chunk buffers cached between chunks
    EQP_CUDA(cudaEventCreate(&c->ev0)); EQP_CUDA(cudaEventCreate(&c->ev1));
    c->noallocRc = cudaGetSymbolAddress((void**)&c->noallocScratch, g_noalloc_scratch);   // this device's copy; a failure is reported by eqp_tokenize_noalloc
    if (c->noallocRc != cudaSuccess) { c->noallocScratch = nullptr; cudaGetLastError(); }
    *out = c.release(); return EQP_OK;
  } catch (const std::exception& ex) { return fail(nullptr, EQP_ERR_CUDA, ex.what()); }
}

int eqp_load_annotation(eqp_ctx* ctx, const eqp_annotation_files* f) {
  if (!ctx || !f) return fail(ctx, EQP_ERR_ARG, "eqp_load_annotation: null argument");
  try {
    EQP_CUDA(cudaSetDevice(ctx->device));
    ctx->H = AnnoHost();
    ctx->H.load(f->combined_exons_bed, f->transcript_exons_bed, f->exon_gene_map, f->exon_exon_map, f->exon_junction_map, f->exon_junction_transcript_map,
                f->junction_junction_coordinate_map, f->equivalent_junctions_map, f->genome_fai);
    ctx->A = ctx->D.upload(ctx->H);
    if (ctx->vec) { cudaFree(ctx->vec); ctx->vec = nullptr; }
    ctx->vecLen = (int64_t)ctx->A.nFeat[0] + ctx->A.nFeat[1] + ctx->A.nFeat[2];
    EQP_CUDA(cudaMalloc((void**)&ctx->vec, (size_t)(ctx->vecLen + 1) * 8)); EQP_CUDA(cudaMemset(ctx->vec, 0, (size_t)(ctx->vecLen + 1) * 8));
    // result buffers: sized once per annotation and page-locked, so the per-chunk D2H of the count vectors runs at PCIe speed
    ctx->unpin();
    for (int m = 0; m < 3; m++) { ctx->R.fixed[m].assign((size_t)ctx->A.nFeat[m], 0); ctx->R.counts[m].assign((size_t)ctx->A.nFeat[m], 0.0); }
    ctx->R.junctionActive.assign((size_t)ctx->A.nAct, 0);
    for (int m = 0; m < 3; m++) ctx->pin(ctx->R.fixed[m].data(), ctx->R.fixed[m].size() * 8);
    ctx->pin(ctx->R.junctionActive.data(), ctx->R.junctionActive.size());
    ctx->annoLoaded = true; ctx->haveChunk = false; return EQP_OK;
  } catch (const AnnoError& ex) { return fail(ctx, ex.code == -5 ? EQP_ERR_UNSUPPORTED : ex.code == -1 ? EQP_ERR_ARG : EQP_ERR_IO, ex.what()); }
  catch (const CudaError& ex) { return fail(ctx, EQP_ERR_CUDA, ex.what()); }
  catch (const std::exception& ex) { return fail(ctx, EQP_ERR_IO, ex.what()); }
}

int64_t eqp_annotation_bytes(eqp_ctx* ctx) { return ctx && ctx->annoLoaded ? (int64_t)ctx->D.bytes : 0; }

// Count the chunk whose three texts are in device memory (text/len); ready[f] (may be null) = event after which text f is complete.
static int run_chunk(eqp_ctx* ctx, const eqp_params* p, const char* const text[3], const uint64_t len[3], cudaEvent_t const ready[3], eqp_counts* out) {
  auto wall = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  double w0 = wall();
  ctx->haveChunk = false; ctx->pl.reset(); ctx->be.reset();          // frees the previous chunk's device buffers (stream-ordered)
  double w1 = wall();
  ctx->be.reset(new CudaBackend(ctx->stream, ctx->nSm));
  RunParams P{};
  P.overlap[0] = p->gene_overlap; P.overlap[1] = p->exon_overlap; P.overlap[2] = p->junction_overlap;
  P.spliceConforming[0] = P.spliceConforming[1] = p->nonsplice_conforming ? 0 : 1; P.spliceConforming[2] = 1;   // CC:932-935,959-960,1034-1035
  P.containment[0] = p->containment ? 1 : 0; P.containment[1] = P.containment[2] = 0;
  P.unambiguous[0] = P.unambiguous[1] = p->unambiguous ? 1 : 0; P.unambiguous[2] = 0;
  P.unweighted = p->unweighted ? 1 : 0; P.modes = (uint8_t)(p->modes & 7); P.weightThreshold = p->read_weight_threshold; P.slack = p->distance_slack;
  P.stranded = (uint8_t)p->stranded; P.genomic = p->genomic_input ? 1 : 0;
  if (P.genomic) P.modes &= 3;
  for (int f = 0; f < 3; f++) ctx->be->inputReady[f] = ready ? ready[f] : nullptr;
  ctx->be->wantHi = P.genomic != 0;
  ctx->pl.reset(new Pipeline<CudaBackend>(*ctx->be, ctx->A, P));
  double w2 = wall();
  EQP_CUDA(cudaEventRecord(ctx->ev0, ctx->stream));
  memset(ctx->R.stats, 0, sizeof ctx->R.stats); memset(ctx->R.err, 0, sizeof ctx->R.err); memset(ctx->R.nRecords, 0, sizeof ctx->R.nRecords);   // the count vectors keep their (page-locked) storage
  int rc = ctx->pl->run(text, len, ctx->R);
  if (rc == 0) {
    int64_t off = 0;
    for (int m = 0; m < 3; m++) { if (ctx->A.nFeat[m]) EQP_CUDA(cudaMemcpyAsync(ctx->vec + off, ctx->pl->V.acc[m], (size_t)ctx->A.nFeat[m] * 8, cudaMemcpyDeviceToDevice, ctx->stream)); off += ctx->A.nFeat[m]; }
  }
  EQP_CUDA(cudaEventRecord(ctx->ev1, ctx->stream)); EQP_CUDA(cudaEventSynchronize(ctx->ev1));
  float ms = 0; EQP_CUDA(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
  ctx->retire(ctx->current, ctx->stream);                           // the text is not needed once the results are on the host
  ctx->be->finish_timing(); ctx->be->report_stages();
  if (ctx->be->timing) fprintf(stderr, "[eqp host ] release %.3f ms, setup %.3f ms, run (events %.3f) %.3f ms wall\n", w1 - w0, w2 - w1, ms, wall() - w2);
  if (rc) {
    char buf[512]; snprintf(buf, sizeof buf, "%s [code %u, detail %u]", derr_text(ctx->R.err[0]), ctx->R.err[0], ctx->R.err[1]);
    return fail(ctx, derr_class(ctx->R.err[0]), buf);
  }
  const ChunkResult& R = ctx->R;
  out->n_gene = ctx->A.nFeat[0]; out->n_exon = ctx->A.nFeat[1]; out->n_junction = ctx->A.nFeat[2]; out->n_junction_act = ctx->A.nAct;
  out->gene = R.counts[0].data(); out->exon = R.counts[1].data(); out->junction = R.counts[2].data();
  out->gene_fixed = (const uint64_t*)R.fixed[0].data(); out->exon_fixed = (const uint64_t*)R.fixed[1].data(); out->junction_fixed = (const uint64_t*)R.fixed[2].data();
  out->junction_active = R.junctionActive.data();
  out->number_mapped_reads = (int64_t)R.stats[9]; out->number_expressed_reads = (int64_t)R.stats[1];
  for (int m = 0; m < 3; m++) { out->num_weight_objects[m] = (int64_t)R.stats[m * 3]; out->num_weight_objects_included[m] = (int64_t)R.stats[m * 3 + 1]; out->num_weight_objects_excluded[m] = (int64_t)R.stats[m * 3 + 2]; out->n_records[m] = R.nRecords[m];
    out->sum_weights[m] = (double)R.stats[10 + m] / 4294967296.0; }
  out->device_ms = ms; out->dominant_kernel_ms = ctx->be->decodeMs; out->kernel_launches = ctx->be->launches;
  ctx->haveChunk = true; return EQP_OK;
}

static int check_params(eqp_ctx* ctx, const eqp_params* p) {
  if (p->stranded < 0 || p->stranded > 2) return fail(ctx, EQP_ERR_ARG, derr_text(DERR_STRANDED));
  if (p->genomic_input && (p->modes & 4)) return fail(ctx, EQP_ERR_UNSUPPORTED, "genomic_input: the junction run of compute-genomic-counts.sh (map filtered through ExtractSplicedExonExonIds and a GTF, :1210-1290) is not built; ask for modes 1 | 2");
  return EQP_OK;
}

int eqp_count_chunk(eqp_ctx* ctx, const eqp_params* p, const uint8_t* t_sam, size_t t_len, const uint8_t* j_sam, size_t j_len,
                    const uint8_t* g_sam, size_t g_len, int on_device, eqp_counts* out) {
  if (!ctx || !p || !out) return fail(ctx, EQP_ERR_ARG, "eqp_count_chunk: null argument");
  if (!ctx->annoLoaded) return fail(ctx, EQP_ERR_ARG, "eqp_count_chunk: call eqp_load_annotation first");
  if ((t_len && !t_sam) || (j_len && !j_sam) || (g_len && !g_sam)) return fail(ctx, EQP_ERR_ARG, "eqp_count_chunk: null SAM buffer with non-zero length");
  if (ctx->feed.active) return fail(ctx, EQP_ERR_ARG, "eqp_count_chunk: a streamed chunk is open (eqp_chunk_begin without eqp_chunk_end)");
  if (int rc = check_params(ctx, p)) return rc;
  if (p->genomic_input && (j_len || g_len)) return fail(ctx, EQP_ERR_ARG, "genomic_input: the one SAM goes into the first input, the other two stay empty");
  memset(out, 0, sizeof *out);
  try {
    EQP_CUDA(cudaSetDevice(ctx->device));
    const uint8_t* src[3] = {t_sam, j_sam, g_sam}; uint64_t len[3] = {t_len, j_len, g_len}; const char* text[3];
    ctx->retire(ctx->current, ctx->stream);
    if (on_device) { for (int f = 0; f < 3; f++) text[f] = (const char*)src[f]; return run_chunk(ctx, p, text, len, nullptr, out); }
    // host buffers: the copies run on the copy stream and the tokenizer of input f waits for copy f only, so tokenizing overlaps the
    // later copies.  A chunk announced with eqp_prefetch_chunk is already on its way (or resident).
    bool pre = false;
    if (!ctx->prefetched.empty()) {
      const eqp_ctx::HostChunk& q = ctx->prefetched.front(); pre = true;
      for (int f = 0; f < 3; f++) pre = pre && q.src[f] == src[f] && q.len[f] == len[f];
      if (pre) { ctx->current = q; ctx->prefetched.pop_front(); }
      else {                                                      // not the announced chunk: the announcements are dropped, this chunk is copied now
        EQP_CUDA(cudaStreamSynchronize(ctx->copyStream));
        for (auto& c : ctx->prefetched) ctx->retire(c, ctx->copyStream);
        ctx->prefetched.clear();
      }
    }
    if (!pre) { for (int f = 0; f < 3; f++) { ctx->current.src[f] = src[f]; ctx->current.len[f] = len[f]; } ctx->start_copy(ctx->current); }
    for (int f = 0; f < 3; f++) text[f] = ctx->current.dev[f];
    return run_chunk(ctx, p, text, len, ctx->current.ready, out);
  } catch (const CudaError& ex) { return fail(ctx, EQP_ERR_CUDA, ex.what()); }
  catch (const std::exception& ex) { return fail(ctx, EQP_ERR_CUDA, ex.what()); }
}

int eqp_chunk_begin(eqp_ctx* ctx, size_t t_hint, size_t j_hint, size_t g_hint) {
  if (!ctx) return fail(ctx, EQP_ERR_ARG, "eqp_chunk_begin: null context");
  if (!ctx->annoLoaded) return fail(ctx, EQP_ERR_ARG, "eqp_chunk_begin: call eqp_load_annotation first");
  try {
    EQP_CUDA(cudaSetDevice(ctx->device));
    if (!ctx->copyStream) EQP_CUDA(cudaStreamCreateWithFlags(&ctx->copyStream, cudaStreamNonBlocking));
    if (ctx->feed.active) ctx->feed_drop();                         // an abandoned stream
    for (int k = 0; k < eqp_ctx::kStages; k++) if (!ctx->stage[k]) {
      EQP_CUDA(cudaHostAlloc((void**)&ctx->stage[k], eqp_ctx::kStageBytes, cudaHostAllocDefault));
      EQP_CUDA(cudaEventCreateWithFlags(&ctx->stageFree[k], cudaEventDisableTiming));
    }
    ctx->feed.active = true;
    size_t hint[3] = {t_hint, j_hint, g_hint};
    for (int f = 0; f < 3; f++) ctx->feed_reserve(f, hint[f] ? hint[f] : 1);
    return EQP_OK;
  } catch (const std::exception& ex) { return fail(ctx, EQP_ERR_CUDA, ex.what()); }
}

int eqp_chunk_feed(eqp_ctx* ctx, int input, const uint8_t* buf, size_t len) {
  if (!ctx || input < 0 || input > 2 || (len && !buf)) return fail(ctx, EQP_ERR_ARG, "eqp_chunk_feed: bad argument");
  if (!ctx->feed.active) return fail(ctx, EQP_ERR_ARG, "eqp_chunk_feed: call eqp_chunk_begin first");
  try {
    EQP_CUDA(cudaSetDevice(ctx->device));
    eqp_ctx::Feed& F = ctx->feed;
    ctx->feed_reserve(input, F.len[input] + len);
    for (size_t off = 0; off < len;) {
      size_t piece = std::min(len - off, eqp_ctx::kStageBytes); int k = ctx->nextStage; ctx->nextStage = (k + 1) % eqp_ctx::kStages;
      EQP_CUDA(cudaEventSynchronize(ctx->stageFree[k]));              // the copy that last used this staging buffer has finished
      memcpy(ctx->stage[k], buf + off, piece);
      EQP_CUDA(cudaMemcpyAsync(F.dev[input] + F.len[input], ctx->stage[k], piece, cudaMemcpyHostToDevice, ctx->copyStream));
      EQP_CUDA(cudaEventRecord(ctx->stageFree[k], ctx->copyStream));
      F.len[input] += piece; off += piece;
    }
    return EQP_OK;
  } catch (const std::exception& ex) { return fail(ctx, EQP_ERR_CUDA, ex.what()); }
}

int eqp_chunk_end(eqp_ctx* ctx, const eqp_params* p, eqp_counts* out) {
  if (!ctx || !p || !out) return fail(ctx, EQP_ERR_ARG, "eqp_chunk_end: null argument");
  if (!ctx->feed.active) return fail(ctx, EQP_ERR_ARG, "eqp_chunk_end: call eqp_chunk_begin first");
  if (int rc = check_params(ctx, p)) return rc;
  memset(out, 0, sizeof *out);
  try {
    EQP_CUDA(cudaSetDevice(ctx->device));
    ctx->retire(ctx->current, ctx->stream);
    eqp_ctx::HostChunk c{}; const char* text[3]; uint64_t len[3];
    for (int f = 0; f < 3; f++) { c.dev[f] = ctx->feed.dev[f]; c.len[f] = ctx->feed.len[f]; text[f] = c.dev[f]; len[f] = c.len[f]; }
    EQP_CUDA(cudaEventCreateWithFlags(&c.ready[0], cudaEventDisableTiming)); EQP_CUDA(cudaEventRecord(c.ready[0], ctx->copyStream));
    ctx->feed = eqp_ctx::Feed();                                    // ownership of the device texts moves to `current` (freed when the chunk is done)
    ctx->current = c;
    cudaEvent_t ready[3] = {c.ready[0], nullptr, nullptr};          // one event after the last piece covers all three inputs
    return run_chunk(ctx, p, text, len, ready, out);
  } catch (const CudaError& ex) { return fail(ctx, EQP_ERR_CUDA, ex.what()); }
  catch (const std::exception& ex) { return fail(ctx, EQP_ERR_CUDA, ex.what()); }
}

// ---- .sam.gz input inflated on the device (eqp_inflate.cuh).  The compressed bytes of one gzip member go to the device on gzStream; the
// inflate runs on the copy stream, on which the feed buffers live, and appends its text to the input's feed buffer.
static size_t inflate_scratch_budget(int sharers = 1) {                // per concurrently inflated member
  size_t budget = (size_t)24 << 30;
  if (const char* e = getenv("EQP_INFLATE_SCRATCH_GB")) { double g = atof(e); if (g >= 0.001) budget = (size_t)(g * (double)((size_t)1 << 30)); }
  size_t freeB = 0, totalB = 0;
  if (cudaMemGetInfo(&freeB, &totalB) == cudaSuccess) {
    int dev = 0; cudaMemPool_t pool; uint64_t reserved = 0, used = 0;   // memory the stream-ordered pool holds but does not use is ours to take as well
    if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess &&
        cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &reserved) == cudaSuccess && cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemCurrent, &used) == cudaSuccess && reserved > used)
      freeB += (size_t)(reserved - used);
    if (freeB > ((size_t)4 << 30)) budget = std::min(budget, (freeB - ((size_t)2 << 30)) / (2 * (size_t)sharers));
  }
  return budget;
}
static uint64_t env_u64(const char* name, uint64_t dflt) { const char* e = getenv(name); return e && *e ? strtoull(e, nullptr, 10) : dflt; }
static char* gz_upload(eqp_ctx* ctx, const uint8_t* gz, size_t len, cudaEvent_t* ready) {
  if (!ctx->gzStream) EQP_CUDA(cudaStreamCreateWithFlags(&ctx->gzStream, cudaStreamNonBlocking));
  void* d = nullptr; EQP_CUDA(cudaMallocAsync(&d, len + 16, ctx->gzStream));
  EQP_CUDA(cudaMemsetAsync((char*)d + (len & ~(size_t)3), 0, len + 16 - (len & ~(size_t)3), ctx->gzStream));       // the words past the last byte read as zero
  EQP_CUDA(cudaMemcpyAsync(d, gz, len, cudaMemcpyHostToDevice, ctx->gzStream));
  EQP_CUDA(cudaEventCreateWithFlags(ready, cudaEventDisableTiming)); EQP_CUDA(cudaEventRecord(*ready, ctx->gzStream));
  return (char*)d;
}
static int gz_inflate_into_feed(eqp_ctx* ctx, int f, const uint8_t* gz, const char* zDev, size_t len, cudaEvent_t ready) {
  eqp_ctx::Feed& F = ctx->feed; const size_t base = F.len[f];
  EQP_CUDA(cudaStreamWaitEvent(ctx->copyStream, ready, 0));
  try {
    ctx->feed_reserve(f, base + std::max<size_t>(3 * len, 1));
    zf::InflateInfo info = zf::device_inflate(gz, (const uint8_t*)zDev, len, ctx->copyStream,
        [&](uint64_t n) { ctx->feed_reserve(f, base + (size_t)n); return (uint8_t*)F.dev[f] + base; },
        env_u64("EQP_INFLATE_SEG", 16384), 0, inflate_scratch_budget(), (uint32_t)env_u64("EQP_INFLATE_GROUPS", 296), (uint32_t)env_u64("EQP_INFLATE_TEST_BOGUS", 0), !getenv("EQP_INFLATE_NO_CRC"));
    F.len[f] = base + (size_t)info.outBytes;
    eqp_inflate_info& o = ctx->lastInflate[f];
    o.gz_bytes = len; o.text_bytes = info.outBytes; o.segments = info.nSeg; o.chain = info.nChain; o.bogus_starts = info.nBogus; o.waves = info.nWaves;
    o.search_ms = info.msSearch; o.decode_ms = info.msDecode; o.resolve_ms = info.msResolve; o.emit_ms = info.msEmit;
    return EQP_OK;
  } catch (const zf::InflateError& e) { return fail(ctx, e.unsupported ? EQP_ERR_UNSUPPORTED : EQP_ERR_FORMAT, "gzip input " + std::to_string(f) + ": " + e.msg); }
}

int eqp_chunk_feed_gz(eqp_ctx* ctx, int input, const uint8_t* gz, size_t len) {
  if (!ctx || input < 0 || input > 2 || !gz || len < 18) return fail(ctx, EQP_ERR_ARG, "eqp_chunk_feed_gz: bad argument");
  if (!ctx->feed.active) return fail(ctx, EQP_ERR_ARG, "eqp_chunk_feed_gz: call eqp_chunk_begin first");
  cudaEvent_t ready = nullptr; char* d = nullptr;
  try {
    EQP_CUDA(cudaSetDevice(ctx->device));
    d = gz_upload(ctx, gz, len, &ready);
    int rc = gz_inflate_into_feed(ctx, input, gz, d, len, ready);
    cudaFreeAsync(d, ctx->copyStream); cudaEventDestroy(ready);
    return rc;
  } catch (const std::exception& ex) { if (d) cudaFreeAsync(d, ctx->copyStream); if (ready) cudaEventDestroy(ready); return fail(ctx, EQP_ERR_CUDA, ex.what()); }
}

// The three members are inflated side by side: one host thread and one CUDA stream per input (copy of the compressed bytes, then the
// inflate's kernels in order on that stream), so the copies overlap the inflates and the kernels of one input fill the SMs another one's
// launch leaves idle (a stream of a few million reads has fewer deflate blocks than the device has thread slots).
int eqp_count_chunk_gz(eqp_ctx* ctx, const eqp_params* p, const uint8_t* t_gz, size_t t_len, const uint8_t* j_gz, size_t j_len, const uint8_t* g_gz, size_t g_len, eqp_counts* out) {
  if (!ctx || !p || !out) return fail(ctx, EQP_ERR_ARG, "eqp_count_chunk_gz: null argument");
  const uint8_t* gz[3] = {t_gz, j_gz, g_gz}; const size_t len[3] = {t_len, j_len, g_len};
  for (int f = 0; f < 3; f++) if (len[f] && (!gz[f] || len[f] < 18)) return fail(ctx, EQP_ERR_ARG, "eqp_count_chunk_gz: bad gzip buffer");
  if (int rc = eqp_chunk_begin(ctx, 0, 0, 0)) return rc;
  int rcs[3] = {EQP_OK, EQP_OK, EQP_OK}; std::string errs[3];
  try {
    // everything queued so far has finished before the three threads allocate: memory that earlier calls freed on the context's streams is then
    // reusable by the stream-ordered pool from any stream (otherwise it may map new memory, which costs hundreds of milliseconds per 10 GB)
    EQP_CUDA(cudaStreamSynchronize(ctx->stream)); EQP_CUDA(cudaStreamSynchronize(ctx->copyStream));
    cudaEvent_t begun; EQP_CUDA(cudaEventCreateWithFlags(&begun, cudaEventDisableTiming)); EQP_CUDA(cudaEventRecord(begun, ctx->copyStream));
    for (int f = 0; f < 3; f++) {
      if (!ctx->inflateStream[f]) EQP_CUDA(cudaStreamCreateWithFlags(&ctx->inflateStream[f], cudaStreamNonBlocking));
      EQP_CUDA(cudaStreamWaitEvent(ctx->inflateStream[f], begun, 0));           // the feed buffers were reserved on the copy stream
    }
    cudaEventDestroy(begun);
    const size_t budget = inflate_scratch_budget(3); const uint64_t seg = env_u64("EQP_INFLATE_SEG", 16384);
    auto work = [&](int f) {
      cudaStream_t st = ctx->inflateStream[f]; void* d = nullptr;
      try {
        EQP_CUDA(cudaSetDevice(ctx->device));
        EQP_CUDA(cudaMallocAsync(&d, len[f] + 16, st));
        EQP_CUDA(cudaMemsetAsync((char*)d + (len[f] & ~(size_t)3), 0, len[f] + 16 - (len[f] & ~(size_t)3), st));
        EQP_CUDA(cudaMemcpyAsync(d, gz[f], len[f], cudaMemcpyHostToDevice, st));
        eqp_ctx::Feed& F = ctx->feed; size_t written = 0;
        ctx->feed_reserve_on(f, std::max<size_t>(3 * len[f], 1), 0, st);
        zf::InflateInfo info = zf::device_inflate(gz[f], (const uint8_t*)d, len[f], st,
            [&](uint64_t n) { ctx->feed_reserve_on(f, (size_t)n, written, st); written = (size_t)n; return (uint8_t*)F.dev[f]; }, seg, 0, budget, (uint32_t)env_u64("EQP_INFLATE_GROUPS", 296), (uint32_t)env_u64("EQP_INFLATE_TEST_BOGUS", 0), !getenv("EQP_INFLATE_NO_CRC"));
        F.len[f] = (size_t)info.outBytes;
        eqp_inflate_info& o = ctx->lastInflate[f];
        o.gz_bytes = len[f]; o.text_bytes = info.outBytes; o.segments = info.nSeg; o.chain = info.nChain; o.bogus_starts = info.nBogus; o.waves = info.nWaves;
        o.search_ms = info.msSearch; o.decode_ms = info.msDecode; o.resolve_ms = info.msResolve; o.emit_ms = info.msEmit;
      } catch (const zf::InflateError& e) { rcs[f] = e.unsupported ? EQP_ERR_UNSUPPORTED : EQP_ERR_FORMAT; errs[f] = "gzip input " + std::to_string(f) + ": " + e.msg; }
      catch (const std::exception& ex) { rcs[f] = EQP_ERR_CUDA; errs[f] = ex.what(); }
      if (d) cudaFreeAsync(d, st);
      cudaStreamSynchronize(st);
    };
    std::vector<std::thread> th;
    for (int f = 0; f < 3; f++) if (len[f]) th.emplace_back(work, f);
    for (auto& t : th) t.join();
  } catch (const std::exception& ex) { ctx->feed_drop(); return fail(ctx, EQP_ERR_CUDA, ex.what()); }
  for (int f = 0; f < 3; f++) if (rcs[f] != EQP_OK) { ctx->feed_drop(); return fail(ctx, rcs[f], errs[f]); }
  return eqp_chunk_end(ctx, p, out);
}

// test tap: inflate one member on the device and copy the text back (dst may be null to ask for the size)
int eqp_debug_inflate(eqp_ctx* ctx, const uint8_t* gz, size_t len, uint8_t* dst, size_t dst_bytes, size_t* n_text) {
  if (!ctx || !gz || !n_text) return fail(ctx, EQP_ERR_ARG, "eqp_debug_inflate: null argument");
  if (int rc = eqp_chunk_begin(ctx, 0, 0, 0)) return rc;
  int rc = eqp_chunk_feed_gz(ctx, 0, gz, len);
  if (rc == EQP_OK) {
    *n_text = ctx->feed.len[0];
    if (dst && dst_bytes >= ctx->feed.len[0] && ctx->feed.len[0]) {
      if (cudaMemcpyAsync(dst, ctx->feed.dev[0], ctx->feed.len[0], cudaMemcpyDeviceToHost, ctx->copyStream) != cudaSuccess || cudaStreamSynchronize(ctx->copyStream) != cudaSuccess)
        rc = fail(ctx, EQP_ERR_CUDA, "eqp_debug_inflate: copy to host failed");
    } else if (dst && dst_bytes < ctx->feed.len[0]) rc = fail(ctx, EQP_ERR_ARG, "eqp_debug_inflate: dst too small");
  }
  cudaStreamSynchronize(ctx->copyStream); ctx->feed_drop();
  return rc;
}

int eqp_last_inflate(eqp_ctx* ctx, int input, eqp_inflate_info* out) {
  if (!ctx || !out || input < 0 || input > 2) return fail(ctx, EQP_ERR_ARG, "eqp_last_inflate: bad argument");
  *out = ctx->lastInflate[input]; return EQP_OK;
}

int eqp_prefetch_chunk(eqp_ctx* ctx, const uint8_t* t_sam, size_t t_len, const uint8_t* j_sam, size_t j_len, const uint8_t* g_sam, size_t g_len) {
  if (!ctx) return fail(ctx, EQP_ERR_ARG, "eqp_prefetch_chunk: null context");
  if ((t_len && !t_sam) || (j_len && !j_sam) || (g_len && !g_sam)) return fail(ctx, EQP_ERR_ARG, "eqp_prefetch_chunk: null SAM buffer with non-zero length");
  if (ctx->prefetched.size() >= 2) return fail(ctx, EQP_ERR_ARG, "eqp_prefetch_chunk: two chunks are already in flight");
  { // never let announced texts crowd out the counting stage: resident texts + its work buffers (about half a text) must fit the device
    size_t resident = t_len + j_len + g_len, largest = resident;
    for (auto& c : ctx->prefetched) { size_t b = c.len[0] + c.len[1] + c.len[2]; resident += b; largest = std::max(largest, b); }
    { size_t b = ctx->current.len[0] + ctx->current.len[1] + ctx->current.len[2]; if (ctx->current.dev[0]) { resident += b; largest = std::max(largest, b); } }
    if ((double)resident + 0.6 * (double)largest > 0.92 * (double)ctx->totalMem)
      return fail(ctx, EQP_ERR_CAPACITY, "eqp_prefetch_chunk: the announced chunk does not fit next to the resident ones (count a chunk first)"); }
  eqp_ctx::HostChunk c{};
  try {
    EQP_CUDA(cudaSetDevice(ctx->device));
    c.src[0] = t_sam; c.src[1] = j_sam; c.src[2] = g_sam; c.len[0] = t_len; c.len[1] = j_len; c.len[2] = g_len;
    ctx->start_copy(c);
    ctx->prefetched.push_back(c); return EQP_OK;
  } catch (const std::exception& ex) {
    cudaGetLastError(); if (ctx->copyStream) { cudaStreamSynchronize(ctx->copyStream); ctx->retire(c, ctx->copyStream); }
    return fail(ctx, EQP_ERR_CUDA, ex.what());
  }
}

int64_t eqp_feature_count(eqp_ctx* ctx, int mode) { return (ctx && ctx->annoLoaded && mode >= 0 && mode < 3) ? (int64_t)ctx->H.featIds[mode].size() : -1; }
const char* eqp_feature_id(eqp_ctx* ctx, int mode, int64_t i) {
  if (!ctx || !ctx->annoLoaded || mode < 0 || mode > 2 || i < 0 || i >= (int64_t)ctx->H.featIds[mode].size()) return nullptr;
  return ctx->H.featIds[mode][(size_t)i].c_str();
}

int eqp_map_columns(eqp_ctx* ctx, int mode) { return (ctx && ctx->annoLoaded && mode >= 0 && mode < 3) ? ctx->H.mapColumns[mode] : -1; }
int64_t eqp_output_row_count(eqp_ctx* ctx, int mode) { return (ctx && ctx->annoLoaded && mode >= 0 && mode < 3) ? (int64_t)ctx->H.outRows[mode].size() : -1; }
int eqp_get_output_row(eqp_ctx* ctx, int mode, int64_t i, eqp_output_row* out) {
  if (!ctx || !out || !ctx->annoLoaded || mode < 0 || mode > 2 || i < 0 || i >= (int64_t)ctx->H.outRows[mode].size()) return fail(ctx, EQP_ERR_ARG, "eqp_get_output_row: bad argument");
  const AnnoHost::OutRow& r = ctx->H.outRows[mode][(size_t)i];
  out->id = r.id.c_str(); out->feature = r.feature; out->act = r.act; out->always = r.always; out->reserved = 0; return EQP_OK;
}

int eqp_device_vector(eqp_ctx* ctx, uint64_t** dev_ptr, int64_t* n) {
  if (!ctx || !dev_ptr || !n) return fail(ctx, EQP_ERR_ARG, "eqp_device_vector: null argument");
  if (!ctx->haveChunk) return fail(ctx, EQP_ERR_ARG, "eqp_device_vector: no chunk counted yet");
  *dev_ptr = (uint64_t*)ctx->vec; *n = ctx->vecLen; return EQP_OK;
}
int eqp_accumulate(eqp_ctx* ctx, uint64_t* acc) {
  if (!ctx || !acc) return fail(ctx, EQP_ERR_ARG, "eqp_accumulate: null argument");
  if (!ctx->haveChunk) return fail(ctx, EQP_ERR_ARG, "eqp_accumulate: no chunk counted yet");
  try {
    EQP_CUDA(cudaSetDevice(ctx->device));
    if (ctx->vecLen) k_add_wide<<<ctx->nSm * 4, 256, 0, ctx->stream>>>((unsigned long long*)acc, ctx->vec, ctx->vecLen);
    EQP_CUDA(cudaGetLastError()); EQP_CUDA(cudaStreamSynchronize(ctx->stream)); return EQP_OK;
  } catch (const std::exception& ex) { return fail(ctx, EQP_ERR_CUDA, ex.what()); }
}
void eqp_wide_to_double(const uint64_t* wide, int64_t n, double* out) {
  if (!wide || !out) return;
  const double scale = 1.0 / (double)(1ULL << EQP_FIXED_SHIFT_D);
  for (int64_t i = 0; i < n; i++) out[i] = (double)wide[n + i] + (double)wide[i] * scale;
}

// ---- NCCL inside the library (eqp_nccl.h)
int eqp_comm_id(void* id_out) {
  if (!id_out) return fail(nullptr, EQP_ERR_ARG, "eqp_comm_id: null argument");
  NcclApi& N = nccl_api(); if (!N.load()) return fail(nullptr, EQP_ERR_NCCL, N.error);
  nccl_unique_id id; int rc = N.GetUniqueId(&id);
  if (rc != NCCL_SUCCESS) return fail(nullptr, EQP_ERR_NCCL, "ncclGetUniqueId: " + N.text(rc));
  static_assert(sizeof id == EQP_COMM_ID_BYTES, "ncclUniqueId is 128 bytes");
  memcpy(id_out, &id, sizeof id); return EQP_OK;
}
int eqp_comm_init(eqp_ctx* ctx, const void* id, int n_ranks, int rank) {
  if (!ctx || !id || n_ranks < 1 || rank < 0 || rank >= n_ranks) return fail(ctx, EQP_ERR_ARG, "eqp_comm_init: bad argument");
  NcclApi& N = nccl_api(); if (!N.load()) return fail(ctx, EQP_ERR_NCCL, N.error);
  try {
    EQP_CUDA(cudaSetDevice(ctx->device));
    if (ctx->comm) { N.CommDestroy(ctx->comm); ctx->comm = nullptr; }
    nccl_unique_id uid; memcpy(&uid, id, sizeof uid);
    int rc = N.CommInitRank(&ctx->comm, n_ranks, uid, rank);
    if (rc != NCCL_SUCCESS) { ctx->comm = nullptr; return fail(ctx, EQP_ERR_NCCL, "ncclCommInitRank: " + N.text(rc)); }
    ctx->commRanks = n_ranks; ctx->commRank = rank; return EQP_OK;
  } catch (const std::exception& ex) { return fail(ctx, EQP_ERR_CUDA, ex.what()); }
}
int eqp_reduce_sample(eqp_ctx* ctx, uint64_t* acc, int root) {
  if (!ctx || !acc) return fail(ctx, EQP_ERR_ARG, "eqp_reduce_sample: null argument");
  if (!ctx->comm) return fail(ctx, EQP_ERR_ARG, "eqp_reduce_sample: call eqp_comm_init first");
  if (root < 0 || root >= ctx->commRanks) return fail(ctx, EQP_ERR_ARG, "eqp_reduce_sample: root out of range");
  if (!ctx->annoLoaded) return fail(ctx, EQP_ERR_ARG, "eqp_reduce_sample: call eqp_load_annotation first");
  NcclApi& N = nccl_api();
  try {
    EQP_CUDA(cudaSetDevice(ctx->device));
    int rc = N.Reduce(acc, acc, (size_t)ctx->vecLen * 2, NCCL_UINT64, NCCL_SUM, root, ctx->comm, ctx->stream);
    if (rc != NCCL_SUCCESS) return fail(ctx, EQP_ERR_NCCL, "ncclReduce: " + N.text(rc));
    EQP_CUDA(cudaStreamSynchronize(ctx->stream)); return EQP_OK;
  } catch (const std::exception& ex) { return fail(ctx, EQP_ERR_CUDA, ex.what()); }
}
int eqp_reduce_sample_local(eqp_ctx** ctxs, int n, uint64_t** accs, int root) {
  if (!ctxs || !accs || n < 1 || root < 0 || root >= n) return fail(nullptr, EQP_ERR_ARG, "eqp_reduce_sample_local: bad argument");
  for (int i = 0; i < n; i++) if (!ctxs[i] || !accs[i] || !ctxs[i]->annoLoaded || ctxs[i]->vecLen != ctxs[0]->vecLen) return fail(ctxs[0], EQP_ERR_ARG, "eqp_reduce_sample_local: contexts must hold the same annotation");
  NcclApi& N = nccl_api(); if (!N.load()) return fail(ctxs[0], EQP_ERR_NCCL, N.error);
  try {
    bool have = true; for (int i = 0; i < n; i++) have = have && ctxs[i]->comm && ctxs[i]->commRanks == n && ctxs[i]->commRank == i;
    if (!have) {                                                     // one communicator clique over this process's contexts, kept for later samples
      std::vector<int> devs(n); std::vector<nccl_comm_t> comms(n);
      for (int i = 0; i < n; i++) { devs[i] = ctxs[i]->device; if (ctxs[i]->comm) { N.CommDestroy(ctxs[i]->comm); ctxs[i]->comm = nullptr; } }
      int rc = N.CommInitAll(comms.data(), n, devs.data());
      if (rc != NCCL_SUCCESS) return fail(ctxs[0], EQP_ERR_NCCL, "ncclCommInitAll: " + N.text(rc));
      for (int i = 0; i < n; i++) { ctxs[i]->comm = comms[i]; ctxs[i]->commRanks = n; ctxs[i]->commRank = i; }
    }
    int rc = N.GroupStart();
    for (int i = 0; i < n && rc == NCCL_SUCCESS; i++) { EQP_CUDA(cudaSetDevice(ctxs[i]->device)); rc = N.Reduce(accs[i], accs[i], (size_t)ctxs[i]->vecLen * 2, NCCL_UINT64, NCCL_SUM, root, ctxs[i]->comm, ctxs[i]->stream); }
    int rc2 = N.GroupEnd(); if (rc == NCCL_SUCCESS) rc = rc2;
    if (rc != NCCL_SUCCESS) return fail(ctxs[0], EQP_ERR_NCCL, "ncclReduce (group): " + N.text(rc));
    for (int i = 0; i < n; i++) { EQP_CUDA(cudaSetDevice(ctxs[i]->device)); EQP_CUDA(cudaStreamSynchronize(ctxs[i]->stream)); }
    return EQP_OK;
  } catch (const std::exception& ex) { return fail(ctxs[0], EQP_ERR_CUDA, ex.what()); }
}

int eqp_debug_fetch(eqp_ctx* ctx, const char* what, void* dst, size_t dstBytes, size_t* nWritten) {
  if (!ctx || !what || !nWritten) return fail(ctx, EQP_ERR_ARG, "eqp_debug_fetch: null argument");
  if (!ctx->haveChunk) return fail(ctx, EQP_ERR_ARG, "eqp_debug_fetch: no chunk counted yet");
  const Pipeline<CudaBackend>& pl = *ctx->pl; const void* src = nullptr; size_t bytes = 0; uint32_t meta[8];
  std::string w = what; size_t N = pl.C.nSlots;
  if (w == "meta") { meta[0] = pl.C.base; meta[1] = pl.C.nSlots; meta[2] = pl.nBlocks; meta[3] = pl.nAln; meta[4] = pl.nHits; meta[5] = pl.C.sam[0].nLines; meta[6] = pl.C.sam[1].nLines; meta[7] = pl.C.sam[2].nLines; }
  else if (w == "weights:T") { src = pl.C.st[0]; bytes = N * sizeof(FragStat); }
  else if (w == "weights:J") { src = pl.C.st[1]; bytes = N * sizeof(FragStat); }
  else if (w == "weights:G") { src = pl.C.st[2]; bytes = N * sizeof(FragStat); }
  else if (w == "weights:TJ") { src = pl.C.stTJ; bytes = N * sizeof(FragStat); }
  else if (w == "weights:final" || w == "a20") {                            // one field of the per-slot SlotPack
    bytes = N * 4; *nWritten = bytes; if (!dst) return EQP_OK;
    if (dstBytes < bytes) return fail(ctx, EQP_ERR_ARG, "eqp_debug_fetch: buffer too small");
    try { EQP_CUDA(cudaSetDevice(ctx->device)); std::vector<SlotPack> sp(N); if (N) ctx->be->to_host(sp.data(), pl.V.pack, N * sizeof(SlotPack));
      uint32_t* o = (uint32_t*)dst; for (size_t i = 0; i < N; i++) o[i] = w == "a20" ? sp[i].a20 : sp[i].nFinal; return EQP_OK; }
    catch (const std::exception& ex) { return fail(ctx, EQP_ERR_CUDA, ex.what()); }
  }
  else if (w == "selection") { src = pl.C.setC; bytes = N * sizeof(SetDesc); }
  else if (w == "blocks") { src = pl.C.blocks; bytes = (size_t)pl.nBlocks * sizeof(Block); }
  else if (w == "hits") { src = pl.C.hits; bytes = (size_t)pl.nHits * sizeof(Hit); }
  else if (w == "anno:buckets" || w == "anno:rows") {                      // host-side annotation tables, to decode the `blocks` / `hits` taps
    std::string names; std::vector<int32_t> se;
    if (w == "anno:buckets") { for (auto& n : ctx->H.bucketNames) { names += n; names += '\n'; } bytes = names.size(); }
    else { se.resize(ctx->H.rowStart.size() * 2); for (size_t i = 0; i < ctx->H.rowStart.size(); i++) { se[2 * i] = ctx->H.rowStart[i]; se[2 * i + 1] = ctx->H.rowEnd[i]; } bytes = se.size() * 4; }
    *nWritten = bytes; if (!dst) return EQP_OK;
    if (dstBytes < bytes) return fail(ctx, EQP_ERR_ARG, "eqp_debug_fetch: buffer too small");
    memcpy(dst, w == "anno:buckets" ? (const void*)names.data() : (const void*)se.data(), bytes); return EQP_OK;
  }
  else return fail(ctx, EQP_ERR_ARG, "eqp_debug_fetch: unknown tap " + w);
  if (w == "meta") { bytes = sizeof meta; *nWritten = bytes; if (!dst) return EQP_OK; if (dstBytes < bytes) return fail(ctx, EQP_ERR_ARG, "eqp_debug_fetch: buffer too small"); memcpy(dst, meta, bytes); return EQP_OK; }
  *nWritten = bytes;
  if (!dst) return EQP_OK;                                            // size query
  if (dstBytes < bytes) return fail(ctx, EQP_ERR_ARG, "eqp_debug_fetch: buffer too small");
  try { EQP_CUDA(cudaSetDevice(ctx->device)); if (bytes) ctx->be->to_host(dst, src, bytes); return EQP_OK; }
  catch (const std::exception& ex) { return fail(ctx, EQP_ERR_CUDA, ex.what()); }
}

int eqp_debug_decode(eqp_ctx* ctx, const uint8_t* sam, size_t len, int misalign, const char** text, size_t* textLen, uint32_t* errCode, uint32_t* errDetail) {
  if (!ctx || !text || !textLen || !errCode || !errDetail || (len && !sam) || misalign < 0 || misalign > 15) return fail(ctx, EQP_ERR_ARG, "eqp_debug_decode: bad argument");
  if (!ctx->annoLoaded) return fail(ctx, EQP_ERR_ARG, "eqp_debug_decode: call eqp_load_annotation first");
  try {
    EQP_CUDA(cudaSetDevice(ctx->device));
    CudaBackend be(ctx->stream, ctx->nSm);
    char* d = be.alloc<char>(len + 64); d += (16 - ((uintptr_t)d & 15)) % 16 + misalign;
    if (len) EQP_CUDA(cudaMemcpyAsync(d, sam, len, cudaMemcpyHostToDevice, ctx->stream));
    uint32_t* err = be.alloc<uint32_t>(4); be.memset(err, 0, 16);
    SamView S{}; uint8_t* proc = nullptr;
    be.decode_fused(ctx->A, S, proc, d, len, err, err + 2);
    uint32_t h[2]; be.to_host(h, err, 8); *errCode = h[0]; *errDetail = h[1];
    size_t n = S.nLines;
    std::vector<uint64_t> lo(n + 1); std::vector<uint32_t> frag(n); std::vector<int32_t> ref(n), pos(n), pnext(n); std::vector<uint16_t> flag(n), co(n), cl(n), sl(n); std::vector<int16_t> nm(n); std::vector<uint8_t> rf(n);
    be.to_host(lo.data(), S.lineOff, (n + 1) * 8);
    if (n) { be.to_host(frag.data(), S.frag, n * 4); be.to_host(ref.data(), S.ref, n * 4); be.to_host(pos.data(), S.pos, n * 4); be.to_host(pnext.data(), S.pnext, n * 4);
      be.to_host(flag.data(), S.flag, n * 2); be.to_host(co.data(), S.cigOff, n * 2); be.to_host(cl.data(), S.cigLen, n * 2); be.to_host(sl.data(), S.seqLen, n * 2);
      be.to_host(nm.data(), S.nm, n * 2); be.to_host(rf.data(), S.rf, n); }
    std::string& out = ctx->debugText; out.clear(); char buf[256];
    for (size_t i = 0; i < n; i++) {
      uint64_t b = lo[i], e = lo[i + 1] - 1;
      if (b > len || e > len || e < b) { snprintf(buf, sizeof buf, "%zu\tBADOFF\t%llu\t%llu\n", i, (unsigned long long)b, (unsigned long long)e); out += buf; continue; }
      if (h[0]) { snprintf(buf, sizeof buf, "%zu\tLINE\t%llu\n", i, (unsigned long long)b); out += buf; continue; }    // records are undefined after an error: offsets only
      if (!(rf[i] & RF_REC)) { snprintf(buf, sizeof buf, "%zu\tSKIP\n", i); out += buf; continue; }
      std::string r = ref[i] >= 0 ? ctx->H.strPool.substr(ctx->H.refStrOff[(size_t)ref[i]], ctx->H.refStrLen[(size_t)ref[i]]) : (ref[i] == -1 ? "*" : "?");
      snprintf(buf, sizeof buf, "%zu\tREC\t%u\t%u\t%s\t%d\t%d\t%d\t", i, frag[i], flag[i], r.c_str(), pos[i], pnext[i], nm[i]); out += buf;
      out.append((const char*)sam + b + co[i], cl[i]);
      snprintf(buf, sizeof buf, "\t%u\t%u\n", sl[i], rf[i]); out += buf;
    }
    *text = out.c_str(); *textLen = out.size(); return EQP_OK;
  } catch (const std::exception& ex) { return fail(ctx, EQP_ERR_CUDA, ex.what()); }
}

// No counterpart in the reference (include/eqp_count.h): the tokenizer kernel with its scratch in g_noalloc_scratch and its output in the
// caller's buffer.  Nothing here allocates: cudaMemsetAsync, the launch, two small copies to host stack variables, a stream synchronise.
int eqp_tokenize_noalloc(eqp_ctx* ctx, const uint8_t* devSam, size_t len, uint64_t maxLines, void* devOut, size_t devOutBytes, eqp_record_columns* cols) {
  if (!ctx || !cols || (len && !devSam)) return fail(ctx, EQP_ERR_ARG, "eqp_tokenize_noalloc: null argument");
  memset(cols, 0, sizeof *cols);
  if (!ctx->annoLoaded) return fail(ctx, EQP_ERR_ARG, "eqp_tokenize_noalloc: call eqp_load_annotation first");
  if (len > EQP_NOALLOC_MAX_TEXT_BYTES) return fail(ctx, EQP_ERR_ARG, "eqp_tokenize_noalloc: text of " + std::to_string(len) + " bytes exceeds EQP_NOALLOC_MAX_TEXT_BYTES (" + std::to_string(EQP_NOALLOC_MAX_TEXT_BYTES) + "): use eqp_count_chunk");
  if (maxLines == 0 || maxLines > 0xFFFFFFF0ull) return fail(ctx, EQP_ERR_ARG, "eqp_tokenize_noalloc: max_lines out of range");
  if (!devOut || devOutBytes < EQP_NOALLOC_OUT_BYTES(maxLines)) return fail(ctx, EQP_ERR_ARG, "eqp_tokenize_noalloc: dev_out is smaller than EQP_NOALLOC_OUT_BYTES(max_lines)");
  if (ctx->noallocRc != cudaSuccess || !ctx->noallocScratch) return fail(ctx, EQP_ERR_CUDA, std::string("eqp_tokenize_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(ctx->noallocRc));
  try {
    EQP_CUDA(cudaSetDevice(ctx->device));
    // columns inside the caller's buffer, each on a 256-byte boundary
    uintptr_t base = (uintptr_t)devOut, cur = base;
    auto col = [&](uint64_t& off, size_t bytes) { cur = (cur + 255) & ~(uintptr_t)255; off = cur - base; void* p = (void*)cur; cur += bytes; return p; };
    const size_t cap = (size_t)maxLines;
    SamView S{}; S.text = (const char*)devSam; S.len = len; S.hi = nullptr;
    S.lineOff = (uint64_t*)col(cols->line_off, (cap + 1) * 8);
    S.frag = (uint32_t*)col(cols->frag, cap * 4); S.ref = (int32_t*)col(cols->ref, cap * 4); S.pos = (int32_t*)col(cols->pos, cap * 4);
    S.rnext = (int32_t*)col(cols->rnext, cap * 4); S.pnext = (int32_t*)col(cols->pnext, cap * 4);
    S.flag = (uint16_t*)col(cols->flag, cap * 2); S.nm = (int16_t*)col(cols->nm, cap * 2); S.cigOff = (uint16_t*)col(cols->cig_off, cap * 2);
    S.cigLen = (uint16_t*)col(cols->cig_len, cap * 2); S.seqLen = (uint16_t*)col(cols->seq_len, cap * 2); S.cigM = (uint16_t*)col(cols->cig_m, cap * 2);
    S.rf = (uint8_t*)col(cols->rf, cap);
    if (cur - base > devOutBytes) return fail(ctx, EQP_ERR_ARG, "eqp_tokenize_noalloc: dev_out is too small for its alignment (give it EQP_NOALLOC_OUT_BYTES(max_lines), 256-byte aligned)");
    if (!len) return EQP_OK;
    const uint64_t vEnd = len + ((uintptr_t)devSam & 15), nTiles = vEnd / TK_OWNED + 1;
    if (nTiles + 4 > kNoallocWords) return fail(ctx, EQP_ERR_ARG, "eqp_tokenize_noalloc: text needs more tile states than the scratch array holds");
    unsigned long long* status = ctx->noallocScratch;                    // the device address, never &g_noalloc_scratch
    uint32_t* err = (uint32_t*)(status + kNoallocWords - 2);              // err[0..1] code/detail, err[2..3] fragment id min/max
    EQP_CUDA(cudaMemsetAsync(status, 0, kNoallocWords * 8, ctx->stream)); // the array keeps the previous call's states: everything starts at zero
    EQP_CUDA(cudaMemsetAsync(err + 2, 0xFF, 4, ctx->stream));
    EQP_CUDA(cudaFuncSetAttribute(k_tokenize_sam, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TokShared)));
    k_tokenize_sam<<<(unsigned)nTiles, TK_THREADS, sizeof(TokShared), ctx->stream>>>(ctx->A, S, (uint32_t)cap, (uint32_t)nTiles, status, (uint32_t*)(status + nTiles + 1), status + nTiles, err, err + 2);
    EQP_CUDA(cudaGetLastError());
    unsigned long long tot = 0; uint32_t h[2] = {0, 0}; char last = 0;
    EQP_CUDA(cudaMemcpyAsync(&tot, status + nTiles, 8, cudaMemcpyDeviceToHost, ctx->stream));
    EQP_CUDA(cudaMemcpyAsync(h, err, 8, cudaMemcpyDeviceToHost, ctx->stream));
    EQP_CUDA(cudaMemcpyAsync(&last, devSam + len - 1, 1, cudaMemcpyDeviceToHost, ctx->stream));
    EQP_CUDA(cudaStreamSynchronize(ctx->stream));
    const uint64_t nl = tot - 1, lines = last == '\n' ? nl : nl + 1;      // tot counts the virtual newline in front of text[0]
    cols->n_lines = lines; cols->err_code = h[0]; cols->err_detail = h[1];
    if (lines > cap) return fail(ctx, EQP_ERR_ARG, "eqp_tokenize_noalloc: the text has " + std::to_string(lines) + " lines, max_lines is " + std::to_string(cap));
    return EQP_OK;
  } catch (const std::exception& ex) { return fail(ctx, EQP_ERR_CUDA, ex.what()); }
}

const char* eqp_last_error(eqp_ctx* ctx) { return ctx ? ctx->err.c_str() : g_err.c_str(); }

void eqp_destroy(eqp_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  ctx->pl.reset(); ctx->be.reset();
  if (ctx->comm && nccl_api().lib) { nccl_api().CommDestroy(ctx->comm); ctx->comm = nullptr; }
  if (ctx->copyStream) cudaStreamSynchronize(ctx->copyStream);
  ctx->feed_drop();
  for (int k = 0; k < eqp_ctx::kStages; k++) { if (ctx->stage[k]) cudaFreeHost(ctx->stage[k]); if (ctx->stageFree[k]) cudaEventDestroy(ctx->stageFree[k]); }
  ctx->retire(ctx->current, ctx->stream); for (auto& c : ctx->prefetched) ctx->retire(c, ctx->stream);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  ctx->unpin();
  if (ctx->vec) cudaFree(ctx->vec);
  if (ctx->ev0) cudaEventDestroy(ctx->ev0);
  if (ctx->ev1) cudaEventDestroy(ctx->ev1);
  if (ctx->gzStream) { cudaStreamSynchronize(ctx->gzStream); cudaStreamDestroy(ctx->gzStream); }
  for (int f = 0; f < 3; f++) if (ctx->inflateStream[f]) { cudaStreamSynchronize(ctx->inflateStream[f]); cudaStreamDestroy(ctx->inflateStream[f]); }
  if (ctx->copyStream) cudaStreamDestroy(ctx->copyStream);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

}  // extern "C"
