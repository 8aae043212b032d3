// CUDA backend of the chunk pipeline (sm_100a): stream-ordered allocation, the generic per-item launcher,
// a three-phase exclusive scan, the work-class ordering and the launch of the one-pass SAM tokenizer (eqp_tokenize.cuh).
// The per-item logic lives in the eqp_*.h headers; this file owns how it is mapped to the machine.
#pragma once
#include <cuda_runtime.h>
#include <cuda_pipeline.h>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>
#include "eqp_pipeline.h"
#include "eqp_tokenize.cuh"

namespace eqpb {

struct CudaError : std::runtime_error { CudaError(const std::string& m) : std::runtime_error(m) {} };
#define EQP_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) throw CudaError(std::string(#x) + ": " + cudaGetErrorString(e_)); } while (0)

static const int EQP_NUM_SMS_DEFAULT = 148;

// ---------------------------------------------------------------- generic per-item kernel
// Tag: an empty struct named after the pipeline stage (eqp_pipeline.h), so that profiler listings read k_items<eqpb::st::blocks_fill, ...>
template <class Tag, class F>
__global__ void __launch_bounds__(256) k_items(uint64_t n, F f) {
  uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) f(i);
}

// same, compiled for at least MINB resident CTAs per SM: the latency-bound per-fragment stages trade registers for warps in flight
template <class Tag, int MINB, class F>
__global__ void __launch_bounds__(256, MINB) k_items_occ(uint64_t n, F f) {
  uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) f(i);
}

// items [0, *count) with the count read on the device: work lists built by an earlier kernel, no host round trip
template <class Tag, class F>
__global__ void __launch_bounds__(256) k_items_counted(const uint32_t* __restrict__ count, F f) {
  uint64_t n = *count, stride = (uint64_t)gridDim.x * blockDim.x;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) f(i);
}

// ---------------------------------------------------------------- scan (exclusive, in place, total at p[n])
static const int SCAN_THREADS = 256, SCAN_ITEMS = 8, SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* total) {   // 256 threads
  __shared__ uint32_t warpSum[8];
  uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5, x = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= (uint32_t)d) x += y; }
  if (lane == 31) warpSum[w] = x;
  __syncthreads();
  if (w == 0) {
    uint32_t s = lane < 8 ? warpSum[lane] : 0;
#pragma unroll
    for (int d = 1; d < 8; d <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, s, d); if (lane >= (uint32_t)d) s += y; }
    if (lane < 8) warpSum[lane] = s;
  }
  __syncthreads();
  uint32_t base = w ? warpSum[w - 1] : 0;
  *total = warpSum[7];
  __syncthreads();
  return base + x - v;
}

__global__ void __launch_bounds__(SCAN_THREADS) k_scan_tile_sums(const uint32_t* __restrict__ p, uint64_t n, uint32_t* __restrict__ sums) {
  uint64_t b = (uint64_t)blockIdx.x * SCAN_TILE; uint32_t s = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) { uint64_t i = b + (uint64_t)k * SCAN_THREADS + threadIdx.x; if (i < n) s += p[i]; }
  uint32_t tot; block_exclusive_scan(s, &tot);
  if (threadIdx.x == 0) sums[blockIdx.x] = tot;
}
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_single(uint32_t* p, uint64_t n) {       // one block walks the whole array
  uint32_t carry = 0;
  for (uint64_t b = 0; b < n; b += SCAN_THREADS) {
    uint64_t i = b + threadIdx.x; uint32_t v = i < n ? p[i] : 0, tot;
    uint32_t ex = block_exclusive_scan(v, &tot);
    if (i < n) p[i] = carry + ex;
    carry += tot;
  }
  if (threadIdx.x == 0) p[n] = carry;
}
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_apply(uint32_t* __restrict__ p, uint64_t n, const uint32_t* __restrict__ tileOff, uint64_t nTiles) {
  uint64_t b = (uint64_t)blockIdx.x * SCAN_TILE + (uint64_t)threadIdx.x * SCAN_ITEMS; uint32_t v[SCAN_ITEMS]; uint32_t s = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) { v[k] = b + k < n ? p[b + k] : 0; s += v[k]; }
  uint32_t tot; uint32_t ex = block_exclusive_scan(s, &tot) + tileOff[blockIdx.x];
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) { if (b + k < n) p[b + k] = ex; ex += v[k]; }
  if (blockIdx.x == nTiles - 1 && threadIdx.x == 0) p[n] = tileOff[nTiles];
}

// ---------------------------------------------------------------- work ordering
// Counting sort of the fragment slots by a small work key (descending, key 0 dropped), so that the lanes of a warp in the
// thread-per-fragment kernels get fragments of equal cost and the expensive fragments start first.
static const int ORD_BINS = 256, ORD_THREADS = 256, ORD_TILE = 4096;
__global__ void __launch_bounds__(ORD_THREADS) k_order_hist(const uint32_t* __restrict__ key, uint64_t n, uint32_t* __restrict__ blockHist, uint32_t nBlocks) {
  __shared__ uint32_t h[ORD_BINS];
  h[threadIdx.x] = 0; __syncthreads();
  uint64_t b = (uint64_t)blockIdx.x * ORD_TILE;
  for (int k = 0; k < ORD_TILE / ORD_THREADS; k++) {
    uint64_t i = b + (uint64_t)k * ORD_THREADS + threadIdx.x; bool ok = i < n; uint32_t v = ok ? key[i] : 0; if (v >= ORD_BINS) v = ORD_BINS - 1;
    uint32_t act = __ballot_sync(0xffffffffu, ok);
    if (ok) { uint32_t peers = __match_any_sync(act, v); if ((threadIdx.x & 31) == (uint32_t)(__ffs(peers) - 1)) atomicAdd(&h[v], __popc(peers)); }
  }
  __syncthreads();
  blockHist[(uint64_t)(ORD_BINS - 1 - threadIdx.x) * nBlocks + blockIdx.x] = h[threadIdx.x];      // bin-major, large keys first
}
__global__ void __launch_bounds__(ORD_THREADS) k_order_scatter(const uint32_t* __restrict__ key, uint64_t n, const uint32_t* __restrict__ blockBase, uint32_t nBlocks,
                                                               uint32_t* __restrict__ perm, uint32_t nKeep) {
  __shared__ uint32_t cur[ORD_BINS];
  cur[threadIdx.x] = blockBase[(uint64_t)(ORD_BINS - 1 - threadIdx.x) * nBlocks + blockIdx.x]; __syncthreads();
  uint64_t b = (uint64_t)blockIdx.x * ORD_TILE;
  for (int k = 0; k < ORD_TILE / ORD_THREADS; k++) {
    uint64_t i = b + (uint64_t)k * ORD_THREADS + threadIdx.x; bool ok = i < n; uint32_t v = ok ? key[i] : 0; if (v >= ORD_BINS) v = ORD_BINS - 1;
    uint32_t act = __ballot_sync(0xffffffffu, ok);
    if (ok) {
      uint32_t peers = __match_any_sync(act, v); uint32_t lane = threadIdx.x & 31, leader = __ffs(peers) - 1, base = 0;
      if (lane == leader) base = atomicAdd(&cur[v], __popc(peers));
      base = __shfl_sync(peers, base, leader);
      uint32_t pos = base + __popc(peers & ((1u << lane) - 1));
      if (pos < nKeep) perm[pos] = (uint32_t)i;                          // key-0 slots sort last and fall off the end
    }
  }
}

// ---------------------------------------------------------------- backend
struct CudaBackend {
  cudaStream_t stream; int nSm;
  cudaStream_t mainStream;
  std::vector<void*> allocs;
  float decodeMs = 0; long launches = 0; uint64_t bytesAllocated = 0;

  // No streams or other heavyweight driver objects are created per chunk: stream creation/destruction serialises on driver locks
  // (tens of ms when another process, e.g. an nvidia-smi poll, holds them).
  explicit CudaBackend(cudaStream_t s, int sms) : stream(s), nSm(sms > 0 ? sms : EQP_NUM_SMS_DEFAULT), mainStream(s) {}
  ~CudaBackend() { finish_timing(); release(); }
  // EQP_STAGE_TIMING=1: CUDA-event time between consecutive stage() marks on the main stream, printed to stderr per chunk
  static bool env_flag(const char* name) { const char* e = getenv(name); return e && e[0] == '1'; }
  bool timing = getenv("EQP_STAGE_TIMING") != nullptr; std::vector<std::pair<const char*, cudaEvent_t>> marks;
  void stage(const char* name) {
    if (!timing) return;
    cudaEvent_t e; EQP_CUDA(cudaEventCreate(&e)); EQP_CUDA(cudaEventRecord(e, mainStream)); marks.push_back({name, e});
  }
  void report_stages() {
    if (!timing || marks.empty()) return;
    EQP_CUDA(cudaEventSynchronize(marks.back().second)); float tot = 0;
    for (size_t i = 1; i < marks.size(); i++) { float ms = 0; cudaEventElapsedTime(&ms, marks[i - 1].second, marks[i].second); tot += ms; fprintf(stderr, "[eqp stage] %-20s %9.3f ms\n", marks[i].first, ms); }
    fprintf(stderr, "[eqp stage] %-20s %9.3f ms (decode kernels %.3f ms)\n", "total", tot, decodeMs);
    for (auto& m : marks) cudaEventDestroy(m.second);
    marks.clear();
  }
  void release() { for (void* p : allocs) cudaFreeAsync(p, stream); allocs.clear(); bytesAllocated = 0; }

  template <class T> T* alloc(uint64_t n) {
    void* p = nullptr; uint64_t bytes = (n ? n : 1) * sizeof(T) + 16;
    EQP_CUDA(cudaMallocAsync(&p, bytes, stream)); allocs.push_back(p); bytesAllocated += bytes; return (T*)p;
  }
  void memset(void* p, int v, uint64_t bytes) { if (bytes) EQP_CUDA(cudaMemsetAsync(p, v, bytes, stream)); }
  void to_host(void* dst, const void* src, uint64_t bytes) {
    EQP_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, stream)); EQP_CUDA(cudaStreamSynchronize(stream));
  }
  void to_device(void* dst, const void* src, uint64_t bytes) {
    EQP_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, stream)); EQP_CUDA(cudaStreamSynchronize(stream));
  }
  int grid_for(uint64_t n, int threads) const {
    uint64_t g = (n + threads - 1) / threads, cap = (uint64_t)nSm * 32;  // grid-stride beyond 32 CTAs per SM
    return (int)(g < cap ? (g ? g : 1) : cap);
  }
  template <class Tag = st::misc, class F> void launch(uint64_t n, F f) {
    if (!n) return;
    k_items<Tag><<<grid_for(n, 256), 256, 0, stream>>>(n, f); launches++;
    EQP_CUDA(cudaGetLastError());
  }
  template <class Tag = st::misc, class F> void launch_counted(const uint32_t* count, uint64_t maxN, F f) {
    if (!maxN) return;
    k_items_counted<Tag><<<nSm * 2, 256, 0, stream>>>(count, f); launches++;
    EQP_CUDA(cudaGetLastError());
  }
  template <class Tag, int MINB, class F> void launch_occ(uint64_t n, F f) {
    if (!n) return;
    k_items_occ<Tag, MINB><<<grid_for(n, 256), 256, 0, stream>>>(n, f); launches++;
    EQP_CUDA(cudaGetLastError());
  }
  void scan(uint32_t* p, uint64_t n) {
    if (n <= (uint64_t)SCAN_TILE * 4) { k_scan_single<<<1, SCAN_THREADS, 0, stream>>>(p, n); launches++; EQP_CUDA(cudaGetLastError()); return; }
    uint64_t nTiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    uint32_t* sums = alloc<uint32_t>(nTiles + 1);
    k_scan_tile_sums<<<(unsigned)nTiles, SCAN_THREADS, 0, stream>>>(p, n, sums); launches++;
    scan(sums, nTiles);
    k_scan_apply<<<(unsigned)nTiles, SCAN_THREADS, 0, stream>>>(p, n, sums, nTiles); launches++;
    EQP_CUDA(cudaGetLastError());
  }
  // perm[0..returned) = slots with key > 0, ordered by descending key bin
  uint32_t order_by_work(const uint32_t* key, uint64_t n, uint32_t* perm) {
    if (!n) return 0;
    uint32_t nBlocks = (uint32_t)((n + ORD_TILE - 1) / ORD_TILE); uint64_t m = (uint64_t)ORD_BINS * nBlocks;
    uint32_t* hist = alloc<uint32_t>(m + 1);
    k_order_hist<<<nBlocks, ORD_THREADS, 0, stream>>>(key, n, hist, nBlocks); launches++;
    scan(hist, m);
    uint32_t nKeep = 0; to_host(&nKeep, hist + (uint64_t)(ORD_BINS - 1) * nBlocks, 4);       // start of the key-0 bin = number of slots kept
    k_order_scatter<<<nBlocks, ORD_THREADS, 0, stream>>>(key, n, hist, nBlocks, perm, nKeep); launches++;
    EQP_CUDA(cudaGetLastError());
    return nKeep;
  }
  // a1/a2 in one pass (eqp_tokenize.cuh): allocates the line index + record SoA of `S`, launches the tokenizer, returns with S.nLines set.
  // The number of lines is only known afterwards, so the tables are sized for len/160 lines and the pass is repeated when that was too small.
  static const bool kFusedDecode = true;
  void alloc_records(SamView& S, uint8_t*& proc, uint64_t cap) {
    S.lineOff = alloc<uint64_t>(cap + 1);
    S.frag = alloc<uint32_t>(cap); S.ref = alloc<int32_t>(cap); S.pos = alloc<int32_t>(cap); S.rnext = alloc<int32_t>(cap); S.pnext = alloc<int32_t>(cap);
    S.flag = alloc<uint16_t>(cap); S.nm = alloc<int16_t>(cap); S.cigOff = alloc<uint16_t>(cap); S.cigLen = alloc<uint16_t>(cap); S.seqLen = alloc<uint16_t>(cap); S.cigM = alloc<uint16_t>(cap);
    S.rf = alloc<uint8_t>(cap); proc = alloc<uint8_t>(cap);
    S.hi = wantHi ? alloc<int32_t>(cap) : nullptr;
  }
  bool wantHi = false;                                                    // genomic-aligner input: keep the HI:i: hit indices
  // host inputs: the H2D copy of each SAM text runs on its own stream; the tokenizer of input f starts when copy f has landed,
  // so tokenizing T and J overlaps the copies of J and G
  cudaEvent_t inputReady[3] = {nullptr, nullptr, nullptr}; int nextInput = 0;
  void decode_fused(const AnnoView& A, SamView& S, uint8_t*& proc, const char* text, uint64_t len, uint32_t* err, uint32_t* fragMinMax) {
    S.text = text; S.len = len; S.nLines = 0;
    if (nextInput < 3 && inputReady[nextInput]) EQP_CUDA(cudaStreamWaitEvent(stream, inputReady[nextInput], 0));
    nextInput++;
    if (!len) { alloc_records(S, proc, 1); return; }
    static bool attr = false;
    if (!attr) { EQP_CUDA(cudaFuncSetAttribute(k_tokenize_sam, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(TokShared))); attr = true; }
    uint64_t vEnd = len + ((uintptr_t)text & 15); uint64_t nTiles = vEnd / TK_OWNED + 1;
    if (nTiles > 0xFFFFFFF0ull) throw CudaError("SAM input too large for the tile counter");
    unsigned long long* status = alloc<unsigned long long>(nTiles + 2);      // [nTiles] = newline total, [nTiles + 1] = tile ticket
    uint64_t cap = len / 160 + 1024;
    for (int attempt = 0; attempt < 2; attempt++) {
      if (cap > 0xFFFFFFF0ull) throw CudaError("more than 2^32 SAM lines in one input");
      alloc_records(S, proc, cap);
      memset(status, 0, (nTiles + 2) * 8);
      cudaEvent_t e0, e1; EQP_CUDA(cudaEventCreate(&e0)); EQP_CUDA(cudaEventCreate(&e1)); decEv.push_back({e0, e1});
      EQP_CUDA(cudaEventRecord(e0, stream));
      k_tokenize_sam<<<(unsigned)nTiles, TK_THREADS, sizeof(TokShared), stream>>>(A, S, (uint32_t)cap, (uint32_t)nTiles, status, (uint32_t*)(status + nTiles + 1), status + nTiles, err, fragMinMax); launches++;
      EQP_CUDA(cudaGetLastError());
      EQP_CUDA(cudaEventRecord(e1, stream));
      unsigned long long tot = 0; to_host(&tot, status + nTiles, 8);
      char last = 0; to_host(&last, text + len - 1, 1);
      uint64_t nl = tot - 1, lines = last == '\n' ? nl : nl + 1;             // tot counts the virtual newline in front of text[0]
      S.nLines = (uint32_t)lines;
      if (lines <= cap) return;
      cap = lines;                                                          // tables were too small (average line under 160 bytes): once more, exact
      for (int k = 0; k < 14 + (wantHi ? 1 : 0); k++) { EQP_CUDA(cudaFreeAsync(allocs.back(), stream)); allocs.pop_back(); }   // the first attempt's tables go back to the pool now
    }
    throw CudaError("line count changed between two passes over the same text");
  }
  void finish_timing() {
    for (auto& p : decEv) { float ms = 0; if (cudaEventElapsedTime(&ms, p.first, p.second) == cudaSuccess) decodeMs += ms; cudaEventDestroy(p.first); cudaEventDestroy(p.second); }
    decEv.clear();
  }
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> decEv;
};

}  // namespace eqpb
