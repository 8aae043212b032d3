// libeqpcount.so — the C ABI declared in include/eqp_count.h, implemented on the CUDA pipeline.
// There is no CPU path in this library: every compute entry point needs a CUDA device and fails with EQP_ERR_CUDA without one.
#include <cstdio>
#include <cstring>
#include <chrono>
#include <deque>
#include <memory>
#include "../../include/eqp_count.h"
#include "eqp_cuda_backend.cuh"

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
  DevArray<uint32_t> bktRowOff; DevArray<uint8_t> bktFlags; DevArray<int32_t> rowStart, rowEnd, rowMaxEnd, rowExon; DevArray<uint8_t> rowFlags;
  DevArray<int32_t> exChr, exStart, exEnd; DevArray<uint8_t> exStrand; DevArray<uint32_t> featOff[3], feat[3];
  size_t bytes = 0;
  AnnoView upload(const AnnoHost& H) {
    AnnoView v = H.view();
    dict.upload(H.dict); strPool.upload(H.strPool.data(), H.strPool.size()); refStrOff.upload(H.refStrOff);
    refStrLen.upload(H.refStrLen); refFlags.upload(H.refFlags); refVerOff.upload(H.refVerOff); refVerBucket.upload(H.refVerBucket);
    refActOff.upload(H.refActOff); refAct.upload(H.refAct); bktRowOff.upload(H.bktRowOff); bktFlags.upload(H.bktFlags);
    rowStart.upload(H.rowStart); rowEnd.upload(H.rowEnd); rowMaxEnd.upload(H.rowMaxEnd); rowExon.upload(H.rowExon); rowFlags.upload(H.rowFlags);
    exChr.upload(H.exChr); exStart.upload(H.exStart); exEnd.upload(H.exEnd); exStrand.upload(H.exStrand);
    v.dict = dict.p; v.strPool = strPool.p; v.refStrOff = refStrOff.p; v.refStrLen = refStrLen.p; v.refFlags = refFlags.p;
    v.refVerOff = refVerOff.p; v.refVerBucket = refVerBucket.p; v.refActOff = refActOff.p; v.refAct = refAct.p;
    v.bktRowOff = bktRowOff.p; v.bktFlags = bktFlags.p; v.rowStart = rowStart.p; v.rowEnd = rowEnd.p; v.rowMaxEnd = rowMaxEnd.p; v.rowExon = rowExon.p; v.rowFlags = rowFlags.p;
    v.exChr = exChr.p; v.exStart = exStart.p; v.exEnd = exEnd.p; v.exStrand = exStrand.p;
    for (int m = 0; m < 3; m++) { featOff[m].upload(H.featOff[m]); feat[m].upload(H.feat[m]); v.featOff[m] = featOff[m].p; v.feat[m] = feat[m].p; }
    bytes = H.dict.size() * 16 + H.strPool.size() + H.refStrOff.size() * 7 + H.refVerOff.size() * 4 + H.refVerBucket.size() * 4 + H.refActOff.size() * 4 + H.refAct.size() * 4 +
            H.bktRowOff.size() * 5 + H.rowStart.size() * 17 + H.exChr.size() * 13;
    for (int m = 0; m < 3; m++) bytes += (H.featOff[m].size() + H.feat[m].size()) * 4;
    return v;
  }
};

__global__ void k_add_u64(unsigned long long* __restrict__ dst, const unsigned long long* __restrict__ src, int64_t n) {
  int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] += src[i];
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
    case DERR_TR_EXON: return "tr- transcript exon hit: not handled on the device yet";
    case DERR_POS_LENDIFF: return "transcript exon longer than its genome exon: not handled on the device yet";
    case DERR_CAPACITY: return "a per-fragment device table overflowed";
    case DERR_GENOME_COORD: return "genomic record does not agree with its exon (BedRecord.java:669-673,683-687)";
    case DERR_INCONSISTENT: return "inconsistent records within one alignment (GenomicAlignment.java:123-148,245-300)";
    case DERR_BAD_JUNCTION_ID: return "junction reference id is not <gene>-junction-<i>-<j>:<l>:<r> (createExtendedJunctionIds.py:93-114)";
    case DERR_STRANDED: return "strand-specific counting is not handled on the device yet";
    case DERR_INT: return "integer SAM field does not parse";
    default: return "unknown device error";
  }
}
int derr_class(uint32_t code) {
  switch (code) {
    case DERR_TR_EXON: case DERR_POS_LENDIFF: case DERR_STRANDED: return EQP_ERR_UNSUPPORTED;
    case DERR_CAPACITY: return EQP_ERR_CAPACITY;
    default: return EQP_ERR_FORMAT;
  }
}

}  // namespace

struct eqp_ctx {
  int device = 0; int nSm = 0; cudaStream_t stream = nullptr;
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
  void pin(void* p, size_t bytes) { if (bytes && cudaHostRegister(p, bytes, cudaHostRegisterDefault) == cudaSuccess) pinned.push_back(p); else cudaGetLastError(); }
  void unpin() { for (void* p : pinned) cudaHostUnregister(p); pinned.clear(); }
};

static thread_local std::string g_err;                              // errors raised before a ctx exists

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
    cudaDeviceProp prop; EQP_CUDA(cudaGetDeviceProperties(&prop, device)); c->nSm = prop.multiProcessorCount;
    EQP_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    cudaMemPool_t pool; EQP_CUDA(cudaDeviceGetDefaultMemPool(&pool, device));
    uint64_t keep = ~0ULL; EQP_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));   // keep chunk buffers cached between chunks
    EQP_CUDA(cudaEventCreate(&c->ev0)); EQP_CUDA(cudaEventCreate(&c->ev1));
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
    ctx->R.junctionActive.assign((size_t)ctx->A.nFeat[2], 0);
    for (int m = 0; m < 3; m++) ctx->pin(ctx->R.fixed[m].data(), ctx->R.fixed[m].size() * 8);
    ctx->pin(ctx->R.junctionActive.data(), ctx->R.junctionActive.size());
    ctx->annoLoaded = true; ctx->haveChunk = false; return EQP_OK;
  } catch (const AnnoError& ex) { return fail(ctx, ex.code == -5 ? EQP_ERR_UNSUPPORTED : ex.code == -1 ? EQP_ERR_ARG : EQP_ERR_IO, ex.what()); }
  catch (const CudaError& ex) { return fail(ctx, EQP_ERR_CUDA, ex.what()); }
  catch (const std::exception& ex) { return fail(ctx, EQP_ERR_IO, ex.what()); }
}

int64_t eqp_annotation_bytes(eqp_ctx* ctx) { return ctx && ctx->annoLoaded ? (int64_t)ctx->D.bytes : 0; }

int eqp_count_chunk(eqp_ctx* ctx, const eqp_params* p, const uint8_t* t_sam, size_t t_len, const uint8_t* j_sam, size_t j_len,
                    const uint8_t* g_sam, size_t g_len, int on_device, eqp_counts* out) {
  if (!ctx || !p || !out) return fail(ctx, EQP_ERR_ARG, "eqp_count_chunk: null argument");
  if (!ctx->annoLoaded) return fail(ctx, EQP_ERR_ARG, "eqp_count_chunk: call eqp_load_annotation first");
  if ((t_len && !t_sam) || (j_len && !j_sam) || (g_len && !g_sam)) return fail(ctx, EQP_ERR_ARG, "eqp_count_chunk: null SAM buffer with non-zero length");
  if (p->stranded) return fail(ctx, EQP_ERR_UNSUPPORTED, derr_text(DERR_STRANDED));
  memset(out, 0, sizeof *out);
  try {
    auto wall = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    double w0 = wall();
    EQP_CUDA(cudaSetDevice(ctx->device));
    ctx->haveChunk = false; ctx->pl.reset(); ctx->be.reset();        // frees the previous chunk's device buffers (stream-ordered)
    double w1 = wall();
    ctx->be.reset(new CudaBackend(ctx->stream, ctx->nSm));
    RunParams P{};
    P.overlap[0] = p->gene_overlap; P.overlap[1] = p->exon_overlap; P.overlap[2] = p->junction_overlap;
    P.spliceConforming[0] = P.spliceConforming[1] = p->nonsplice_conforming ? 0 : 1; P.spliceConforming[2] = 1;   // CC:932-935,959-960,1034-1035
    P.containment[0] = p->containment ? 1 : 0; P.containment[1] = P.containment[2] = 0;
    P.unambiguous[0] = P.unambiguous[1] = p->unambiguous ? 1 : 0; P.unambiguous[2] = 0;
    P.unweighted = p->unweighted ? 1 : 0; P.modes = (uint8_t)(p->modes & 7); P.weightThreshold = p->read_weight_threshold; P.slack = p->distance_slack;
    const uint8_t* src[3] = {t_sam, j_sam, g_sam}; uint64_t len[3] = {t_len, j_len, g_len}; const char* text[3];
    ctx->retire(ctx->current, ctx->stream);
    if (on_device) { for (int f = 0; f < 3; f++) text[f] = (const char*)src[f]; }
    else {
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
      for (int f = 0; f < 3; f++) { text[f] = ctx->current.dev[f]; ctx->be->inputReady[f] = ctx->current.ready[f]; }
    }
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
    ctx->retire(ctx->current, ctx->stream);                         // the text is not needed once the results are on the host
    ctx->be->finish_timing(); ctx->be->report_stages();
    if (ctx->be->timing) fprintf(stderr, "[eqp host ] release %.3f ms, setup %.3f ms, run (events %.3f) %.3f ms wall\n", w1 - w0, w2 - w1, ms, wall() - w2);
    if (rc) {
      char buf[512]; snprintf(buf, sizeof buf, "%s [code %u, detail %u]", derr_text(ctx->R.err[0]), ctx->R.err[0], ctx->R.err[1]);
      return fail(ctx, derr_class(ctx->R.err[0]), buf);
    }
    const ChunkResult& R = ctx->R;
    out->n_gene = ctx->A.nFeat[0]; out->n_exon = ctx->A.nFeat[1]; out->n_junction = ctx->A.nFeat[2];
    out->gene = R.counts[0].data(); out->exon = R.counts[1].data(); out->junction = R.counts[2].data();
    out->gene_fixed = (const uint64_t*)R.fixed[0].data(); out->exon_fixed = (const uint64_t*)R.fixed[1].data(); out->junction_fixed = (const uint64_t*)R.fixed[2].data();
    out->junction_active = R.junctionActive.data();
    out->number_mapped_reads = (int64_t)R.stats[9]; out->number_expressed_reads = (int64_t)R.stats[1];
    for (int m = 0; m < 3; m++) { out->num_weight_objects[m] = (int64_t)R.stats[m * 3]; out->num_weight_objects_included[m] = (int64_t)R.stats[m * 3 + 1]; out->num_weight_objects_excluded[m] = (int64_t)R.stats[m * 3 + 2]; out->n_records[m] = R.nRecords[m]; }
    out->device_ms = ms; out->dominant_kernel_ms = ctx->be->decodeMs; out->kernel_launches = ctx->be->launches;
    ctx->haveChunk = true; return EQP_OK;
  } catch (const CudaError& ex) { return fail(ctx, EQP_ERR_CUDA, ex.what()); }
  catch (const std::exception& ex) { return fail(ctx, EQP_ERR_CUDA, ex.what()); }
}

int eqp_prefetch_chunk(eqp_ctx* ctx, const uint8_t* t_sam, size_t t_len, const uint8_t* j_sam, size_t j_len, const uint8_t* g_sam, size_t g_len) {
  if (!ctx) return fail(ctx, EQP_ERR_ARG, "eqp_prefetch_chunk: null context");
  if ((t_len && !t_sam) || (j_len && !j_sam) || (g_len && !g_sam)) return fail(ctx, EQP_ERR_ARG, "eqp_prefetch_chunk: null SAM buffer with non-zero length");
  if (ctx->prefetched.size() >= 2) return fail(ctx, EQP_ERR_ARG, "eqp_prefetch_chunk: two chunks are already in flight");
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
    if (ctx->vecLen) k_add_u64<<<ctx->nSm * 4, 256, 0, ctx->stream>>>((unsigned long long*)acc, ctx->vec, ctx->vecLen);
    EQP_CUDA(cudaGetLastError()); EQP_CUDA(cudaStreamSynchronize(ctx->stream)); return EQP_OK;
  } catch (const std::exception& ex) { return fail(ctx, EQP_ERR_CUDA, ex.what()); }
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
  else if (w == "weights:final") { src = pl.V.nFinal; bytes = N * 4; }
  else if (w == "a20") { src = pl.V.a20; bytes = N * 4; }
  else if (w == "selection") { src = pl.C.setC; bytes = N * sizeof(SetDesc); }
  else if (w == "blocks") { src = pl.C.blocks; bytes = (size_t)pl.nBlocks * sizeof(Block); }
  else if (w == "hits") { src = pl.C.hits; bytes = (size_t)pl.nHits * sizeof(Hit); }
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

const char* eqp_last_error(eqp_ctx* ctx) { return ctx ? ctx->err.c_str() : g_err.c_str(); }

void eqp_destroy(eqp_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  ctx->pl.reset(); ctx->be.reset();
  if (ctx->copyStream) cudaStreamSynchronize(ctx->copyStream);
  ctx->retire(ctx->current, ctx->stream); for (auto& c : ctx->prefetched) ctx->retire(c, ctx->stream);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  ctx->unpin();
  if (ctx->vec) cudaFree(ctx->vec);
  if (ctx->ev0) cudaEventDestroy(ctx->ev0);
  if (ctx->ev1) cudaEventDestroy(ctx->ev1);
  if (ctx->copyStream) cudaStreamDestroy(ctx->copyStream);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

}  // extern "C"
