// Kernels and host driver of the device inflate (eqp_inflate.h has the algorithm and the per-thread functions).
//   k_zf_search   thread per segment: first plausible dynamic-block start inside the segment
//   k_zf_decode   thread per listed segment: blocks -> 16-bit symbols in the segment's scratch region (Huffman tables in shared memory,
//                 interleaved over the CTA's threads)
//   k_zf_resolve_groups / k_zf_chain_windows  markers in every chain segment's last 32 K symbols, by groups of segments (see there)
//   k_zf_crc      CRC-32 of the finished text by 1 MB pieces (combined on the host) against the gzip trailer
//   k_zf_emit     CTA per (segment, slice): pass 0 writes the tails' bytes, pass 1 everything else (markers read the text in front of the segment)
#pragma once
#include <cuda_runtime.h>
#include <algorithm>
#include <string>
#include <vector>
#include "eqp_inflate.h"

namespace eqpb {
namespace zf {

// Occupancy is what these kernels live on (every lane follows its own bit stream: a step is a chain of dependent shared-memory and L2
// latencies): the code limits / bases stay in shared memory (128 bytes per thread), the symbol tables (640 bytes per thread) live in global
// memory, one contiguous table per thread — 1024 threads per SM instead of 256 with everything in shared memory.
#ifdef EQP_VAR_ZF_SMEM_SYMS                                          // A/B: everything in shared memory, 768 bytes per thread, 256 threads per SM
static const int ZF_THREADS = 128;
static const int ZF_CTAS = 2;
static const int ZF_TABLE_U16 = 16 + 16 + 16 + 16 + 288 + 32;
#else
static const int ZF_THREADS = 256;
static const int ZF_CTAS = 4;
static const int ZF_TABLE_U16 = 16 + 16 + 16 + 16;                   // limit / base of both codes, interleaved over the CTA's threads
#endif
static const size_t ZF_SMEM = (size_t)ZF_TABLE_U16 * ZF_THREADS * 2;
static const int ZF_SYMS = 320;                                      // 288 literal/length + 32 distance symbols per thread

__device__ __forceinline__ void zf_tables(uint16_t* sm, uint16_t* symTables, Huff& hl, Huff& hd) {
  const int t = threadIdx.x, S = ZF_THREADS;
  hl.limit = sm + t; hl.base = (int16_t*)(sm + 16 * S + t); hd.limit = sm + 32 * S + t; hd.base = (int16_t*)(sm + 48 * S + t);
  hl.stride = hd.stride = S;
#ifdef EQP_VAR_ZF_SMEM_SYMS
  (void)symTables; hl.sym = sm + 64 * S + t; hd.sym = sm + (64 + 288) * S + t; hl.symStride = hd.symStride = S;
#else
  hl.sym = symTables + ((size_t)blockIdx.x * ZF_THREADS + t) * ZF_SYMS; hd.sym = hl.sym + 288; hl.symStride = hd.symStride = 1;
#endif
}

// Every lane runs one bounded step of its state machine between two warp barriers, until all lanes of the warp are done: without the barrier
// the lanes drift apart at the first early exit and never meet again (measured: the plain per-thread loops ran at one active lane per warp).
__global__ void __launch_bounds__(ZF_THREADS, ZF_CTAS) k_zf_search(const uint32_t* w, uint64_t nWords, uint64_t totalBits, uint64_t dataBit, uint64_t segBits, uint32_t nSeg, uint64_t* segStart, uint16_t* symTables) {
  extern __shared__ __align__(16) uint16_t zf_sm[];
  Huff hl, hd; zf_tables(zf_sm, symTables, hl, hd);
  __shared__ uint8_t lut[512];
  for (uint32_t k = threadIdx.x; k < 512; k += ZF_THREADS) lut[k] = kraft3(k);
  __syncthreads();
  const uint32_t i = blockIdx.x * ZF_THREADS + threadIdx.x;
  Search s; search_init(s, 0, 0);
  if (i > 0 && i < nSeg) { uint64_t from = dataBit + (uint64_t)i * segBits, to = from + segBits; if (to > totalBits) to = totalBits; search_init(s, from, to); }
  while (true) {
    __syncwarp();
    const uint32_t scanning = __ballot_sync(0xffffffffu, !s.done && !s.pending), pending = __ballot_sync(0xffffffffu, !s.done && s.pending);
    if (!(scanning | pending)) break;
    if (__popc(pending) >= 8 || !scanning) { if (!s.done && s.pending) search_check(s, w, nWords, totalBits, hl, hd); }   // the long check, for several lanes at once
    else if (!s.done && !s.pending) search_scan(s, w, nWords, lut);
  }
  if (i < nSeg) segStart[i] = i == 0 ? dataBit : s.found;
}

// one thread per listed segment.  WRITE = false: the counting pass (block boundaries and output lengths, no memory traffic besides the
// compressed words); WRITE = true: 16-bit symbols into the segment's exact-size scratch region.
struct DecodeItem { uint64_t sym, cap; uint32_t seg, pad; };
template <bool WRITE>
__global__ void __launch_bounds__(ZF_THREADS, ZF_CTAS) k_zf_decode(const uint32_t* w, uint64_t nWords, uint64_t totalBits, const uint64_t* segStart, const uint8_t* bogus, uint32_t nSeg,
                                                                   const DecodeItem* list, uint32_t nList, uint16_t* scratch, SegOut* res, uint16_t* symTables) {
  extern __shared__ __align__(16) uint16_t zf_sm[];
  Huff hl, hd; zf_tables(zf_sm, symTables, hl, hd);
  const uint32_t k = blockIdx.x * ZF_THREADS + threadIdx.x;
  Dec d; d.state = ST_DONE;
  if (k < nList) {
    const DecodeItem it = list[k];
    Sink s; s.out = WRITE ? scratch + it.sym : nullptr; s.o = 0; s.cap = it.cap; s.window = it.seg != 0; s.status = 0;
    dec_init(d, w, nWords, segStart[it.seg], it.seg, nSeg, s);
  }
  while (true) {
    __syncwarp();
    if (!__any_sync(0xffffffffu, d.state != ST_DONE)) break;
    dec_step<WRITE>(d, totalBits, segStart, nSeg, bogus, hl, hd);      // every lane, finished or not: the step has a warp barrier inside
  }
  if (k < nList) res[k] = d.r;
}

struct ChainSeg { uint64_t sym; uint64_t abs; uint64_t len; };       // scratch offset (symbols), position in the stream's output, symbols

// ---- resolve: the markers in the last 32 K symbols of every chain segment.  A tail's markers point into the tail(s) in front of it, so the
// chain is sequential — but "window in front of segment s" is a FUNCTION of the window in front of an earlier segment, and functions compose:
//   k_zf_resolve_groups  the chain is cut into groups of consecutive segments, one CTA per group walks its segments in order with a
//                        window of 16-bit ENTRIES in shared memory (a byte, or 0x8000|k = "byte k of the window in front of the GROUP");
//                        afterwards every tail of the group and the group's final window F_g are expressed in the group's initial window
//   k_zf_chain_windows   one CTA, one cheap step per group: initial window of group g (bytes) -> through F_g -> initial window of g + 1
//   k_zf_emit pass 0     reads a tail's remaining markers from its group's initial window
// (The first version walked the whole chain in one CTA: 8-10 us per segment, 0.4-0.9 s per 1.5 GB of gzip, whatever the instruction count.)
static const int ZF_RES_THREADS = 512;
static const int ZF_RES_GROUPS = 9;                                    // 9 * 512 groups of 8 symbols >= 32768 / 8 + 1
struct TailRef { uint64_t a, lo, hi, absA; };                          // scratch index of group 0 (multiple of 8), valid scratch range [lo, hi), output position of a
__device__ __forceinline__ TailRef zf_tail_ref(const ChainSeg& cs) {
  const uint64_t T = cs.len < ZF_WINDOW ? cs.len : ZF_WINDOW; TailRef r; r.lo = cs.sym + (cs.len - T); r.hi = cs.sym + cs.len; r.a = r.lo & ~7ull;
  r.absA = cs.abs + (r.a - cs.sym);                                     // (may wrap below zero for the stream's first bytes: only used with valid elements)
  return r;
}
__device__ __forceinline__ void zf_tail_load(const uint16_t* scratch, const TailRef& r, uint32_t t, uint4 (&v)[ZF_RES_GROUPS]) {
#pragma unroll
  for (int i = 0; i < ZF_RES_GROUPS; i++) {
    const uint64_t g = r.a + 8ull * (t + (uint32_t)i * ZF_RES_THREADS);
    v[i] = g < r.hi ? *reinterpret_cast<const uint4*>(scratch + g) : make_uint4(0, 0, 0, 0);
  }
}
__global__ void __launch_bounds__(ZF_RES_THREADS) k_zf_resolve_groups(uint16_t* scratch, const ChainSeg* chain, uint32_t nChain, uint32_t perGroup, uint16_t* groupFinal) {
  extern __shared__ __align__(16) uint16_t win[];                      // win[x & 32767]: entry of output position x, for the 32 K positions in front of the current segment
  const uint32_t t = threadIdx.x, c0 = blockIdx.x * perGroup, c1 = min(nChain, c0 + perGroup);
  if (c0 >= c1) return;
  const ChainSeg none{0, 0, 0};
  ChainSeg cs = chain[c0], nx = c0 + 1 < c1 ? chain[c0 + 1] : none;
  const uint64_t absG = cs.abs;
  for (uint32_t k = t; k < ZF_WINDOW; k += ZF_RES_THREADS) win[(absG - ZF_WINDOW + k) & (ZF_WINDOW - 1)] = (uint16_t)(0x8000u | k);
  uint4 v[ZF_RES_GROUPS], vn[ZF_RES_GROUPS];
  zf_tail_load(scratch, zf_tail_ref(cs), t, v);
  __syncthreads();
  uint64_t absEnd = absG;
  for (uint32_t c = c0; c < c1; c++) {
    const ChainSeg nn = c + 2 < c1 ? chain[c + 2] : none;
    zf_tail_load(scratch, zf_tail_ref(nx), t, vn);                     // in flight during this step
    const TailRef r = zf_tail_ref(cs);
#pragma unroll
    for (int i = 0; i < ZF_RES_GROUPS; i++) {
      const uint64_t g = r.a + 8ull * (t + (uint32_t)i * ZF_RES_THREADS);
      uint32_t q[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
      if (g < r.hi && ((q[0] | q[1] | q[2] | q[3]) & 0x80008000u)) {
#pragma unroll
        for (int e = 0; e < 8; e++) {
          const uint32_t x = (q[e >> 1] >> (16 * (e & 1))) & 0xFFFFu;
          if ((x & 0x8000u) && g + e >= r.lo && g + e < r.hi) {         // entry w of the window in front of the segment
            const uint32_t b = win[(cs.abs - ZF_WINDOW + (x & 0x7FFFu)) & (ZF_WINDOW - 1)]; scratch[g + e] = (uint16_t)b;
            q[e >> 1] = (q[e >> 1] & ~(0xFFFFu << (16 * (e & 1)))) | (b << (16 * (e & 1)));
          }
        }
        v[i] = make_uint4(q[0], q[1], q[2], q[3]);
      }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < ZF_RES_GROUPS; i++) {
      const uint64_t g = r.a + 8ull * (t + (uint32_t)i * ZF_RES_THREADS);
      if (g < r.hi) {
        const uint32_t q[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
        const uint32_t at = (uint32_t)(r.absA + (g - r.a));
#pragma unroll
        for (int e = 0; e < 8; e++) if (g + e >= r.lo && g + e < r.hi) win[(at + e) & (ZF_WINDOW - 1)] = (uint16_t)(q[e >> 1] >> (16 * (e & 1)));
      }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < ZF_RES_GROUPS; i++) v[i] = vn[i];
    absEnd = cs.abs + cs.len; cs = nx; nx = nn;
  }
  uint16_t* F = groupFinal + (size_t)blockIdx.x * ZF_WINDOW;           // entry k = output position absEnd - 32768 + k
  for (uint32_t k = t; k < ZF_WINDOW; k += ZF_RES_THREADS) F[k] = win[(absEnd - ZF_WINDOW + k) & (ZF_WINDOW - 1)];
}
// winCarry: the 32 K bytes in front of the wave (byte k = output position start - 32768 + k).  groupInit[g] = window in front of group g.
__global__ void __launch_bounds__(1024) k_zf_chain_windows(const ChainSeg* chain, uint32_t nChain, uint32_t perGroup, uint32_t nGroups, const uint16_t* groupFinal, uint8_t* groupInit, uint8_t* winCarry, uint32_t* err) {
  extern __shared__ __align__(16) uint8_t zf_L[];                       // two windows
  uint8_t (*L)[ZF_WINDOW] = reinterpret_cast<uint8_t (*)[ZF_WINDOW]>(zf_L);
  const uint32_t t = threadIdx.x;
  for (uint32_t k = t; k < ZF_WINDOW; k += 1024) L[0][k] = winCarry[k];
  __syncthreads();
  int cur = 0;
  for (uint32_t g = 0; g < nGroups; g++) {
    const uint64_t absG = chain[g * perGroup].abs; const uint16_t* F = groupFinal + (size_t)g * ZF_WINDOW; uint8_t* init = groupInit + (size_t)g * ZF_WINDOW;
    for (uint32_t k = t; k < ZF_WINDOW; k += 1024) {
      init[k] = L[cur][k];
      const uint32_t e = F[k]; uint32_t b = e;
      if (e & 0x8000u) { const uint32_t w = e & 0x7FFFu; b = L[cur][w]; }
      L[cur ^ 1][k] = (uint8_t)b;
    }
    (void)absG;
    __syncthreads();
    cur ^= 1;
  }
  for (uint32_t k = t; k < ZF_WINDOW; k += 1024) winCarry[k] = L[cur][k];
  (void)err; (void)nChain;
}

// pass 0: the last min(len, 32 K) symbols of every chain segment (their markers now point into the group's initial window);
// pass 1: the symbols in front of them (markers read the text in front of the segment, which pass 0 has written)
static const uint32_t ZF_EMIT_SLICE = 1u << 16;
__global__ void __launch_bounds__(256) k_zf_emit(const uint16_t* scratch, const ChainSeg* chain, const uint32_t* sliceSeg, const uint32_t* sliceIdx, int pass, uint8_t* text,
                                                 uint32_t perGroup, const uint8_t* groupInit, uint32_t* err) {
  const uint32_t c = sliceSeg[blockIdx.x];
  const ChainSeg cs = chain[c];
  const uint64_t T = cs.len < ZF_WINDOW ? cs.len : ZF_WINDOW, head = cs.len - T;
  uint64_t k0 = (uint64_t)sliceIdx[blockIdx.x] * ZF_EMIT_SLICE, k1 = k0 + ZF_EMIT_SLICE;
  if (pass == 0) { k0 += head; k1 += head; if (k1 > cs.len) k1 = cs.len; } else if (k1 > head) k1 = head;
  const uint16_t* src = scratch + cs.sym; uint8_t* dst = text + cs.abs;
  const uint8_t* init = groupInit + (size_t)(c / perGroup) * ZF_WINDOW; const uint64_t absG = chain[(c / perGroup) * perGroup].abs;
  for (uint64_t k = k0 + threadIdx.x; k < k1; k += 256) {
    uint32_t v = src[k];
    if (v & 0x8000u) {
      const uint32_t w = v & 0x7FFFu;
      if (pass == 0) { if (absG < ZF_WINDOW && w < ZF_WINDOW - absG) atomicExch(err, 1u); v = init[w]; }          // (reaches in front of the stream: corrupt data)
      else { if (cs.abs < ZF_WINDOW && w < ZF_WINDOW - cs.abs) { atomicExch(err, 1u); v = 0; } else v = text[cs.abs - ZF_WINDOW + w]; }
    }
    dst[k] = (uint8_t)v;
  }
}

// CRC-32 of the text: CTA per 1 MB piece, thread per 4 KB sub-piece (16-byte loads, byte-wise table in shared memory), thread 0 combines
__global__ void __launch_bounds__(ZF_CRC_SUBS) k_zf_crc(const uint8_t* text, uint64_t n, uint32_t opSub, uint32_t* pieceCrc) {
  __shared__ uint32_t table[256]; __shared__ uint32_t sub[ZF_CRC_SUBS];
  const uint32_t t = threadIdx.x;
  table[t] = crc_table_entry(t);
  __syncthreads();
  const uint64_t base = (uint64_t)blockIdx.x * ZF_CRC_SUB * ZF_CRC_SUBS, plen = n - base < (uint64_t)ZF_CRC_SUB * ZF_CRC_SUBS ? n - base : (uint64_t)ZF_CRC_SUB * ZF_CRC_SUBS;
  const uint64_t lo = (uint64_t)t * ZF_CRC_SUB; uint32_t crc = 0;
  if (lo < plen) {
    const uint64_t len = plen - lo < ZF_CRC_SUB ? plen - lo : ZF_CRC_SUB; const uint8_t* p = text + base + lo;
    uint32_t c = 0xFFFFFFFFu; uint64_t i = 0;
    if ((((uintptr_t)p) & 15) == 0) for (; i + 16 <= len; i += 16) {
      const uint4 q = *reinterpret_cast<const uint4*>(p + i); const uint32_t wv[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
      for (int k = 0; k < 4; k++) { uint32_t x = wv[k];
#pragma unroll
        for (int b = 0; b < 4; b++) { c = table[(c ^ x) & 0xFFu] ^ (c >> 8); x >>= 8; } }
    }
    for (; i < len; i++) c = table[(c ^ p[i]) & 0xFFu] ^ (c >> 8);
    crc = ~c;
  }
  sub[t] = crc;
  __syncthreads();
  if (t == 0) pieceCrc[blockIdx.x] = crc_piece(sub, plen, opSub);
}

struct InflateInfo { uint64_t outBytes = 0; uint32_t nSeg = 0, nChain = 0, nBogus = 0, nWaves = 0, reruns = 0; float msSearch = 0, msDecode = 0, msResolve = 0, msEmit = 0; };
struct InflateError { std::string msg; bool unsupported; };          // unsupported: a valid gzip file this decoder does not take (several members)

#define ZF_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) throw InflateError{std::string("device inflate: ") + cudaGetErrorString(e_), false}; } while (0)

// Inflates ONE gzip member.  zHost: the whole file in host memory (header and trailer are read there); zDev: the same bytes in device
// memory, 4-byte aligned, followed by at least 8 zero bytes; copyDone: event after which zDev is complete (may be null).
// reserve(n) returns a device buffer of at least n bytes that keeps what was written so far (it may move).  Work runs on `st`; returns
// after the last kernel has finished.  segBytes / scratchBytes: segment size, scratch budget (ratio is unused: the counting pass sizes every segment exactly); maxGroups: CTAs of the
// resolve; testBogusEvery: test hook.
template <class Reserve>
InflateInfo device_inflate(const uint8_t* zHost, const uint8_t* zDev, uint64_t zLen, cudaStream_t st, Reserve reserve,
                           uint64_t segBytes = 65536, uint64_t ratio = 0, uint64_t scratchBytes = 24ull << 30, uint32_t maxGroups = 296, uint32_t testBogusEvery = 0, bool checkCrc = true) {
  InflateInfo info;
  if (zLen < 18 || zHost[0] != 0x1f || zHost[1] != 0x8b || zHost[2] != 8) throw InflateError{"not a gzip file (magic / method)", false};
  const uint8_t flg = zHost[3]; uint64_t off = 10;
  if (flg & 4) { if (off + 2 > zLen) throw InflateError{"truncated gzip header", false}; off += 2 + (zHost[off] | (zHost[off + 1] << 8)); }
  if (flg & 8) { while (off < zLen && zHost[off]) off++; off++; }
  if (flg & 16) { while (off < zLen && zHost[off]) off++; off++; }
  if (flg & 2) off += 2;
  if (off + 8 >= zLen) throw InflateError{"truncated gzip file", false};
  const uint64_t dataBit = off * 8, totalBits = (zLen - 8) * 8, nWords = (zLen + 3) / 4 + 1;
  const uint32_t isize = zHost[zLen - 4] | (zHost[zLen - 3] << 8) | (zHost[zLen - 2] << 16) | ((uint32_t)zHost[zLen - 1] << 24);
  const uint64_t nSeg64 = (zLen - 8 - off + segBytes - 1) / segBytes;
  if (nSeg64 > 0x7FFFFFF0ull) throw InflateError{"gzip file too large for the segment counter", false};
  const uint32_t nSeg = (uint32_t)nSeg64; info.nSeg = nSeg;
  const uint32_t* w = (const uint32_t*)zDev;
  (void)ratio;

  std::vector<void*> owned;
  auto dalloc = [&](size_t bytes) { void* p = nullptr; ZF_CUDA(cudaMallocAsync(&p, bytes ? bytes : 16, st)); owned.push_back(p); return p; };
  auto release = [&]() { for (void* p : owned) cudaFreeAsync(p, st); owned.clear(); };
  cudaEvent_t ev[2] = {nullptr, nullptr};
  try {
    ZF_CUDA(cudaEventCreate(&ev[0])); ZF_CUDA(cudaEventCreate(&ev[1]));
    auto timed = [&](float& acc, auto&& launch) { ZF_CUDA(cudaEventRecord(ev[0], st)); launch(); ZF_CUDA(cudaGetLastError()); ZF_CUDA(cudaEventRecord(ev[1], st)); ZF_CUDA(cudaEventSynchronize(ev[1])); float ms = 0; cudaEventElapsedTime(&ms, ev[0], ev[1]); acc += ms; };
    ZF_CUDA(cudaFuncSetAttribute(k_zf_search, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ZF_SMEM));
    ZF_CUDA(cudaFuncSetAttribute(k_zf_decode<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ZF_SMEM));
    ZF_CUDA(cudaFuncSetAttribute(k_zf_decode<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ZF_SMEM));
    uint64_t* dStart = (uint64_t*)dalloc((size_t)nSeg * 8); uint8_t* dBogus = (uint8_t*)dalloc(nSeg); SegOut* dRes = (SegOut*)dalloc((size_t)nSeg * sizeof(SegOut));
    DecodeItem* dList = (DecodeItem*)dalloc((size_t)nSeg * sizeof(DecodeItem));
    uint8_t* dWin = (uint8_t*)dalloc(ZF_WINDOW); uint32_t* dErr = (uint32_t*)dalloc(16);
    uint16_t* dSym = (uint16_t*)dalloc(((size_t)nSeg + ZF_THREADS) * ZF_SYMS * 2);      // one symbol table per launched thread
    ZF_CUDA(cudaMemsetAsync(dBogus, 0, nSeg, st)); ZF_CUDA(cudaMemsetAsync(dWin, 0, ZF_WINDOW, st)); ZF_CUDA(cudaMemsetAsync(dErr, 0, 16, st));
    timed(info.msSearch, [&] { k_zf_search<<<(nSeg + ZF_THREADS - 1) / ZF_THREADS, ZF_THREADS, ZF_SMEM, st>>>(w, nWords, totalBits, dataBit, segBytes * 8, nSeg, dStart, dSym); });
    std::vector<uint64_t> start(nSeg); ZF_CUDA(cudaMemcpyAsync(start.data(), dStart, (size_t)nSeg * 8, cudaMemcpyDeviceToHost, st)); ZF_CUDA(cudaStreamSynchronize(st));
    std::vector<uint8_t> bogus(nSeg, 0);
    if (testBogusEvery) {                                                 // test hook: every k-th found start is moved one bit on, i.e. made wrong
      for (uint32_t i = 1, k = 0; i < nSeg; i++) if (start[i] != ZF_NONE && ++k % testBogusEvery == 0) start[i] += 1;
      ZF_CUDA(cudaMemcpyAsync(dStart, start.data(), (size_t)nSeg * 8, cudaMemcpyHostToDevice, st));
    }
    // ---- counting pass over every segment with a start, then the chain walk (a chain member that passed a claimed start proves it bogus)
    std::vector<DecodeItem> items; std::vector<uint32_t> itemOf(nSeg, 0xFFFFFFFFu);
    for (uint32_t i = 0; i < nSeg; i++) if (start[i] != ZF_NONE) { itemOf[i] = (uint32_t)items.size(); items.push_back(DecodeItem{0, 0, i, 0}); }
    // A large member in which the search finds next to nothing (stored blocks, binary content: the search only recognises blocks that start with
    // text) would be decoded by a handful of threads for minutes: leave such a file to the host.
    if (nSeg >= 4096 && items.size() < nSeg / 64) throw InflateError{"no parallelism to be found in this member (" + std::to_string(items.size()) + " block starts in " + std::to_string(nSeg) + " segments: stored blocks or binary content)", true};
    std::vector<SegOut> res(items.size());
    ZF_CUDA(cudaMemcpyAsync(dList, items.data(), items.size() * sizeof(DecodeItem), cudaMemcpyHostToDevice, st));
    timed(info.msDecode, [&] { k_zf_decode<false><<<((uint32_t)items.size() + ZF_THREADS - 1) / ZF_THREADS, ZF_THREADS, ZF_SMEM, st>>>(w, nWords, totalBits, dStart, dBogus, nSeg, dList, (uint32_t)items.size(), nullptr, dRes, dSym); });
    ZF_CUDA(cudaMemcpyAsync(res.data(), dRes, items.size() * sizeof(SegOut), cudaMemcpyDeviceToHost, st)); ZF_CUDA(cudaStreamSynchronize(st));
    std::vector<ChainSeg> chain; std::vector<uint32_t> chainSeg; uint64_t abs = 0, endBit = 0; bool done = false;
    for (uint32_t cur = 0; !done;) {
      if (cur >= nSeg || itemOf[cur] == 0xFFFFFFFFu) throw InflateError{"deflate stream ends without a final block", false};
      const SegOut r = res[itemOf[cur]];
      if (r.status == ZF_RETRY) {
        if (r.endSeg >= nSeg || bogus[r.endSeg]) throw InflateError{"inconsistent block chain", false};
        bogus[r.endSeg] = 1; info.nBogus++; info.reruns++;
        ZF_CUDA(cudaMemcpyAsync(dBogus + r.endSeg, &bogus[r.endSeg], 1, cudaMemcpyHostToDevice, st));
        k_zf_decode<false><<<1, ZF_THREADS, ZF_SMEM, st>>>(w, nWords, totalBits, dStart, dBogus, nSeg, dList + itemOf[cur], 1u, nullptr, dRes + itemOf[cur], dSym);
        ZF_CUDA(cudaGetLastError());
        ZF_CUDA(cudaMemcpyAsync(&res[itemOf[cur]], dRes + itemOf[cur], sizeof(SegOut), cudaMemcpyDeviceToHost, st)); ZF_CUDA(cudaStreamSynchronize(st));
        continue;
      }
      if (r.status != ZF_OK && r.status != ZF_FINAL) throw InflateError{"invalid deflate data (segment " + std::to_string(cur) + ", code " + std::to_string(r.status) + ")", false};
      chain.push_back(ChainSeg{0, abs, r.outLen}); chainSeg.push_back(cur); abs += r.outLen;
      if (r.status == ZF_FINAL) { done = true; endBit = r.endBit; } else cur = r.endSeg;
    }
    info.nChain = (uint32_t)chain.size();
    if ((endBit + 7) / 8 + 8 != zLen) throw InflateError{"bytes after the first gzip member (several members / BGZF are not inflated on the device)", true};
    if ((uint32_t)abs != isize) throw InflateError{"gzip length check failed (ISIZE)", false};
    uint8_t* text = reserve(abs);
    // ---- writing pass in waves: as many chain segments as the scratch holds, tightly packed
    uint64_t scratchSyms = std::max<uint64_t>(std::min<uint64_t>(scratchBytes / 2, abs), 1);
    for (const ChainSeg& c : chain) scratchSyms = std::max(scratchSyms, c.len);          // (a single segment larger than the budget still has to fit)
    uint16_t* scratch = (uint16_t*)dalloc(scratchSyms * 2 + 64);     // (the resolve kernel reads whole 16-byte groups)
    ChainSeg* dChain = (ChainSeg*)dalloc(chain.size() * sizeof(ChainSeg) + 16); uint32_t* dSlice = nullptr; size_t sliceCap = 0;
    uint16_t* dGroupFinal = nullptr; uint8_t* dGroupInit = nullptr; uint32_t groupCap = 0;
    ZF_CUDA(cudaFuncSetAttribute(k_zf_resolve_groups, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(ZF_WINDOW * 2)));
    ZF_CUDA(cudaFuncSetAttribute(k_zf_chain_windows, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(ZF_WINDOW * 2)));
    const uint32_t nChain = (uint32_t)chain.size();
    for (uint32_t c0 = 0; c0 < nChain;) {
      uint32_t c1 = c0; uint64_t used = 0;
      while (c1 < nChain && (c1 == c0 || used + chain[c1].len <= scratchSyms)) { chain[c1].sym = used; used += chain[c1].len; c1++; }
      info.nWaves++;
      const uint32_t n = c1 - c0;
      std::vector<DecodeItem> wl(n); for (uint32_t c = c0; c < c1; c++) wl[c - c0] = DecodeItem{chain[c].sym, chain[c].len, chainSeg[c], 0};
      ZF_CUDA(cudaMemcpyAsync(dList, wl.data(), (size_t)n * sizeof(DecodeItem), cudaMemcpyHostToDevice, st));
      ZF_CUDA(cudaMemcpyAsync(dChain, chain.data() + c0, (size_t)n * sizeof(ChainSeg), cudaMemcpyHostToDevice, st));
      timed(info.msDecode, [&] { k_zf_decode<true><<<(n + ZF_THREADS - 1) / ZF_THREADS, ZF_THREADS, ZF_SMEM, st>>>(w, nWords, totalBits, dStart, dBogus, nSeg, dList, n, scratch, dRes, dSym); });
      std::vector<SegOut> wr(n); ZF_CUDA(cudaMemcpyAsync(wr.data(), dRes, (size_t)n * sizeof(SegOut), cudaMemcpyDeviceToHost, st)); ZF_CUDA(cudaStreamSynchronize(st));
      for (uint32_t k = 0; k < n; k++) {
        if (wr[k].status == ZF_BAD || wr[k].status == ZF_NOWINDOW) throw InflateError{"invalid deflate data (distance too far back, segment " + std::to_string(wl[k].seg) + ")", false};
        if (wr[k].outLen != wl[k].cap || (wr[k].status != ZF_OK && wr[k].status != ZF_FINAL)) throw InflateError{"the two decoding passes disagree (segment " + std::to_string(wl[k].seg) + ")", false};
      }
      std::vector<uint32_t> sl[2][2];                                      // per pass: slice -> (chain index within the wave, slice index)
      for (uint32_t c = 0; c < n; c++) {
        const uint64_t len = chain[c0 + c].len, T = len < ZF_WINDOW ? len : ZF_WINDOW, hd = len - T;
        for (uint64_t k = 0; k * ZF_EMIT_SLICE < T; k++) { sl[0][0].push_back(c); sl[0][1].push_back((uint32_t)k); }
        for (uint64_t k = 0; k * ZF_EMIT_SLICE < hd; k++) { sl[1][0].push_back(c); sl[1][1].push_back((uint32_t)k); }
      }
      const size_t nSl = 2 * (sl[0][0].size() + sl[1][0].size());
      if (nSl > sliceCap) { sliceCap = 2 * nSl + 16; dSlice = (uint32_t*)dalloc(sliceCap * 4); }
      uint32_t* dS[2][2]; size_t at = 0;
      for (int p = 0; p < 2; p++) for (int q = 0; q < 2; q++) { dS[p][q] = dSlice + at; if (!sl[p][q].empty()) ZF_CUDA(cudaMemcpyAsync(dS[p][q], sl[p][q].data(), sl[p][q].size() * 4, cudaMemcpyHostToDevice, st)); at += sl[p][q].size(); }
      const uint32_t nGroups = std::min<uint32_t>(n, std::max<uint32_t>(maxGroups, 1)), perGroup = (n + nGroups - 1) / nGroups, usedGroups = (n + perGroup - 1) / perGroup;
      if (usedGroups > groupCap) { groupCap = usedGroups; dGroupFinal = (uint16_t*)dalloc((size_t)groupCap * ZF_WINDOW * 2); dGroupInit = (uint8_t*)dalloc((size_t)groupCap * ZF_WINDOW); }
      timed(info.msResolve, [&] {
        k_zf_resolve_groups<<<usedGroups, ZF_RES_THREADS, ZF_WINDOW * 2, st>>>(scratch, dChain, n, perGroup, dGroupFinal);
        k_zf_chain_windows<<<1, 1024, ZF_WINDOW * 2, st>>>(dChain, n, perGroup, usedGroups, dGroupFinal, dGroupInit, dWin, dErr);
      });
      for (int p = 0; p < 2; p++) if (!sl[p][0].empty()) timed(info.msEmit, [&] { k_zf_emit<<<(uint32_t)sl[p][0].size(), 256, 0, st>>>(scratch, dChain, dS[p][0], dS[p][1], p, text, perGroup, dGroupInit, dErr); });
      ZF_CUDA(cudaStreamSynchronize(st));                                  // the host vectors above are reused by the next wave
      c0 = c1;
    }
    uint32_t herr = 0; ZF_CUDA(cudaMemcpyAsync(&herr, dErr, 4, cudaMemcpyDeviceToHost, st)); ZF_CUDA(cudaStreamSynchronize(st));
    if (herr) throw InflateError{"invalid deflate data (distance in front of the stream)", false};
    if (checkCrc && abs) {                                                 // the trailer's CRC-32 against the text as it now stands in device memory
      const uint64_t PIECE = (uint64_t)ZF_CRC_SUB * ZF_CRC_SUBS; const uint32_t nPieces = (uint32_t)((abs + PIECE - 1) / PIECE);
      uint32_t* dCrc = (uint32_t*)dalloc((size_t)nPieces * 4); std::vector<uint32_t> pc(nPieces);
      timed(info.msEmit, [&] { k_zf_crc<<<nPieces, ZF_CRC_SUBS, 0, st>>>(text, abs, crc_shift_op(ZF_CRC_SUB), dCrc); });
      ZF_CUDA(cudaMemcpyAsync(pc.data(), dCrc, (size_t)nPieces * 4, cudaMemcpyDeviceToHost, st)); ZF_CUDA(cudaStreamSynchronize(st));
      const uint32_t opPiece = crc_shift_op(PIECE); uint32_t acc = 0;
      for (uint32_t i = 0; i < nPieces; i++) { const uint64_t plen = abs - (uint64_t)i * PIECE < PIECE ? abs - (uint64_t)i * PIECE : PIECE; acc = crc_combine(acc, pc[i], plen == PIECE ? opPiece : crc_shift_op(plen)); }
      const uint32_t want = zHost[zLen - 8] | (zHost[zLen - 7] << 8) | (zHost[zLen - 6] << 16) | ((uint32_t)zHost[zLen - 5] << 24);
      if (acc != want) throw InflateError{"gzip CRC check failed", false};
    }
    info.outBytes = abs;
    release(); cudaEventDestroy(ev[0]); cudaEventDestroy(ev[1]);
    return info;
  } catch (...) { release(); if (ev[0]) cudaEventDestroy(ev[0]); if (ev[1]) cudaEventDestroy(ev[1]); throw; }
}

}  // namespace zf
}  // namespace eqpb
