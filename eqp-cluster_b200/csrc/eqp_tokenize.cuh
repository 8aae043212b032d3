// a1/a2 on the B200: ONE pass over the SAM text that builds the line index and decodes every record (sm_100a).
// Replaces SamReader.readSamFile's line loop + `new SamRecord(line)` (J/sam-source/SamReader.java:139-301, SamRecord.java:121-271).
//
// A CTA owns the lines that START in a 30,704-byte tile.  The tile plus a 16-byte lead-in and a 2 KB overhang (so the last
// owned line is complete) is brought into shared memory by one TMA bulk copy (cp.async.bulk, mbarrier completion).
//   phase 1a  byte-parallel: every 16-byte piece -> 16-bit masks of its '\n' and '\t' bytes (conflict-free LDS.128, exact SWAR
//             compare, one multiply gathers both masks of a word)
//   phase 1b  thread t owns 8 consecutive pieces: block scan of the newline counts, list of newline positions nlPos[]
//   look-back one thread publishes the tile's newline count and chains the exclusive prefix over earlier tiles (decoupled
//             look-back, tile ids from an atomic ticket so that every predecessor is already running) -> global line numbers
//   phase 2   thread per owned line: line = (nlPos[j], nlPos[j+1]); its tabs are read off the tab masks 64 text bytes per
//             word (StageSrc), names and tags are handled in 4-byte groups, SoA stores are coalesced over consecutive lines
// Lines longer than the overhang and tiles with more lines than nlPos holds take a per-line / per-tile slow path that
// parses from global memory with the generic decoder.
#pragma once
#include <cuda_runtime.h>
#include "eqp_decode.h"

namespace eqpb {

static const int TK_THREADS = 128;                                   // a tile holds ~90 lines: more threads would only idle in phase 2
static const int TK_MIN_CTAS = 5;                                    // per SM: 5 x 43 KB of shared memory
static const int TK_PIECES = 2048;                                   // 16-byte pieces staged per tile
static const int TK_STAGE = TK_PIECES * 16;                          // 32 KB
static const int TK_OV_PIECES = 128;                                 // 2 KB overhang
static const int TK_OWNED = (TK_PIECES - 1 - TK_OV_PIECES) * 16;     // 30,704 bytes of line starts per tile
static const int TK_PER_THREAD = TK_PIECES / TK_THREADS;             // 16 pieces = 256 bytes
static const int TK_NL_CAP = 1024;
static const int TK_FIRST_OV_THREAD = (TK_PIECES - TK_OV_PIECES) / TK_PER_THREAD;   // 120: first thread whose pieces are all overhang
static_assert((TK_PIECES - TK_OV_PIECES) % TK_PER_THREAD == 0 && TK_PER_THREAD % 8 == 0 && (TK_FIRST_OV_THREAD & 31) != 0, "tile geometry");

struct TokShared {
  alignas(128) char stage[TK_STAGE];
  alignas(16) uint16_t tabM[TK_PIECES + 8];                          // +8: StageSrc may read the 8-byte word after the last piece
  alignas(16) uint16_t nlM[TK_PIECES];
  uint16_t nlPos[TK_NL_CAP + 2];
  uint32_t warpTot[TK_THREADS / 32];
  alignas(8) unsigned long long bar, barBase;                        // tile staged / lineBase published
  unsigned long long lineBase;
  uint32_t tile, nNl, nOwned;
};

static const unsigned long long TK_FLAG_AGG = 1ULL << 62, TK_FLAG_PREFIX = 2ULL << 62, TK_VALUE_MASK = (1ULL << 62) - 1;

__device__ __forceinline__ uint32_t tk_smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// 4 text bytes -> top byte of the result: bits 24-27 = bytes equal to '\t', bits 28-31 = bytes equal to '\n' (exact compares).
// '\t' = 0x09 and '\n' = 0x0a are the bytes b with (b ^ 8) in {1, 2}: x = w ^ 0x08080808, a byte of x below 4 (and without its top bit)
// is one of 0x08..0x0b, bit 0 / bit 1 of x then tell tab / newline (0x0b has both and counts as neither).  (x | 0x80) - 4 cannot
// borrow across bytes.  The multiply moves the marker of byte k (bit 3 for tab, bit 7 for newline) to bit 24+k / 28+k: all 32
// partial-product positions are distinct, so nothing carries.
__device__ __forceinline__ uint32_t tk_marks4(uint32_t w) {
#ifdef EQP_VAR_OLD_MARKS
  uint32_t xn = w ^ 0x0a0a0a0au, xt = w ^ 0x09090909u;
  uint32_t zn0 = ~(((xn & 0x7f7f7f7fu) + 0x7f7f7f7fu) | xn | 0x7f7f7f7fu);
  uint32_t zt0 = ~(((xt & 0x7f7f7f7fu) + 0x7f7f7f7fu) | xt | 0x7f7f7f7fu);
  return (zn0 | (zt0 >> 4)) * 0x00204081u;
#endif
  uint32_t x = w ^ 0x08080808u;
  uint32_t t = (x | 0x80808080u) - 0x04040404u;                        // top bit of a byte clear <=> its low 7 bits are below 4
  uint32_t m = ~(t | x) & 0x80808080u;
  uint32_t a = x << 7, b = x << 6;                                      // bit 0 / bit 1 of every byte at its top bit
  uint32_t zt = m & a & ~b, zn = m & b & ~a;
  return (zn | (zt >> 4)) * 0x00204081u;
}
// four bytes (hi nibble = newline bits, lo nibble = tab bits, one byte per text word) -> 16-bit masks
__device__ __forceinline__ uint32_t tk_nibbles16(uint32_t t) {        // t: nibbles at bits 0-3, 8-11, 16-19, 24-27
  t = (t | (t >> 4)) & 0x00FF00FFu; return (t | (t >> 8)) & 0xFFFFu;
}

// Source for a line staged in shared memory: tabs from the tile's tab masks, 64 text bytes per 8-byte load; groups from two
// aligned words and a funnel shift.
struct StageSrc {
  const uint32_t* tabW;                                               // tab masks, one bit per staged byte, 32 bytes of text per word
  int s, lim, wi; uint32_t word;                                      // line start / end (staged positions), current mask word
  __device__ __forceinline__ StageSrc(const uint16_t* tabM, int s_, int n) : tabW(reinterpret_cast<const uint32_t*>(tabM)), s(s_), lim(s_ + n), wi(s_ >> 5) {
    word = tabW[wi] & (0xFFFFFFFFu << (s & 31));
  }
  __device__ __forceinline__ int next(const char*, int, int n) {
    while (!word && ((wi + 1) << 5) < lim) word = tabW[++wi];
    int res = n;
    if (word) { int q = (wi << 5) + (__ffs((int)word) - 1) - s; word &= word - 1; res = q < n ? q : n; }
    return res;
  }
  static __device__ __forceinline__ uint32_t load4(const char* q, int nvalid) {
    uint32_t a = (uint32_t)__cvta_generic_to_shared(q), lo, hi;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(lo) : "r"(a & ~3u));
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(hi) : "r"((a & ~3u) + 4u));          // past the stage buffer this is tabM: readable, masked below
    uint32_t v = __funnelshift_r(lo, hi, (a & 3u) * 8u);
    return nvalid >= 4 ? v : (nvalid <= 0 ? 0u : v & ((1u << (8 * nvalid)) - 1u));
  }
};

// Serial slow path of a whole tile (thread 0): same outputs as the fast path, parsed from global memory.
__device__ __noinline__ void tk_slow_tile(const AnnoView& A, const SamView& S, uint32_t capLines, long long vb, int a0, unsigned long long lineBase, bool firstTile, uint32_t* err, uint32_t* fragMinMax) {
  unsigned long long k = lineBase;
  for (long long v = vb - 1; v < vb + TK_OWNED - 1; v++) {
    long long g = v - a0;
    bool nl = (g == -1 && firstTile) || (g >= 0 && g < (long long)S.len && S.text[g] == '\n');
    if (!nl) continue;
    uint64_t st = (uint64_t)(g + 1);
    if (k <= capLines) S.lineOff[k] = st;
    if (st < S.len) {
      uint64_t e = st; while (e < S.len && S.text[e] != '\n') e++;
      if (k < capLines) { decode_record(A, S, (uint32_t)k, S.text + st, (int)(e - st), err); if (S.rf[k] & RF_REC) { atomicMin(fragMinMax, S.frag[k]); atomicMax(fragMinMax + 1, S.frag[k]); } }
      if (e == S.len && k + 1 <= capLines) S.lineOff[k + 1] = S.len + 1;
    }
    k++;
  }
}

__global__ void __launch_bounds__(TK_THREADS, TK_MIN_CTAS)
k_tokenize_sam(const __grid_constant__ AnnoView A, const __grid_constant__ SamView S, uint32_t capLines, uint32_t nTiles, unsigned long long* status, uint32_t* ticket, unsigned long long* nlTotal, uint32_t* err, uint32_t* fragMinMax) {
  extern __shared__ __align__(128) unsigned char tk_raw[];
  TokShared& sh = *reinterpret_cast<TokShared*>(tk_raw);
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int a0 = (int)((uintptr_t)S.text & 15);                      // v = text offset + a0: virtual coordinate with a 16-byte aligned origin
  const char* base = S.text - a0;
  const long long vEnd = (long long)S.len + a0, vEndAl = (vEnd + 15) & ~15LL;

  if (tid == 0) {
    sh.tile = atomicAdd(ticket, 1u);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(tk_smem_addr(&sh.bar)));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(tk_smem_addr(&sh.barBase)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const uint32_t tile = sh.tile;
  const long long vb = (long long)tile * TK_OWNED;                   // owned line starts: v in [vb, vb + TK_OWNED)
  const long long v0 = vb - 16;                                      // v of staged byte 0
  if (tid == 0) {
    long long c0 = v0 < 0 ? 0 : v0, c1 = v0 + TK_STAGE < vEndAl ? v0 + TK_STAGE : vEndAl;
    uint32_t bytes = (uint32_t)(c1 - c0);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tk_smem_addr(&sh.bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(tk_smem_addr(sh.stage + (c0 - v0))), "l"(base + c0), "r"(bytes), "r"(tk_smem_addr(&sh.bar)) : "memory");
  }
  { uint32_t done = 0, bar = tk_smem_addr(&sh.bar);                  // thread 0 waits for the copy, the others for thread 0 ...
    if (tid == 0) while (!done) asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(bar) : "memory");
    __syncthreads();
    done = 0;                                                          // ... and then observe the completed phase themselves (acquire)
    while (!done) asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(bar) : "memory"); }

  // ---- phase 1a: newline and tab masks per 16-byte piece
  const uint4* sv = reinterpret_cast<const uint4*>(sh.stage);
  const int vnPos = 15 + a0;                                         // staged position of the virtual newline in front of text[0] (tile 0)
#pragma unroll
  for (int k = 0; k < TK_PER_THREAD; k++) {
    uint32_t p = tid + k * TK_THREADS; uint4 q = sv[p];
    uint32_t r0 = tk_marks4(q.x), r1 = tk_marks4(q.y), r2 = tk_marks4(q.z), r3 = tk_marks4(q.w);
    uint32_t B = __byte_perm(__byte_perm(r0, r1, 0x0073), __byte_perm(r2, r3, 0x0073), 0x5410);   // top bytes of r0..r3
    uint32_t tb = tk_nibbles16(B & 0x0F0F0F0Fu), nl = tk_nibbles16((B >> 4) & 0x0F0F0F0Fu);
    long long g0 = v0 + 16LL * p - a0;                               // text offset of the piece's byte 0
    if (g0 < 0 || g0 + 16 > (long long)S.len) {                      // piece straddles the text's ends: keep valid bytes only
      long long lo = g0 < 0 ? -g0 : 0, hi = (long long)S.len - g0;
      if (lo > 16) lo = 16; if (hi > 16) hi = 16; if (hi < 0) hi = 0;
      uint32_t ok = (hi > lo) ? (((1u << hi) - 1u) & ~((1u << lo) - 1u)) : 0u;
      tb &= ok; nl &= ok;
    }
    if (p == 0) { nl &= 0x8000u; tb = 0; }                           // lead-in: only a newline right before the tile matters
    if (tile == 0 && p == (uint32_t)(vnPos >> 4)) nl |= 1u << (vnPos & 15);
    sh.tabM[p] = (uint16_t)tb; sh.nlM[p] = (uint16_t)nl;
  }
  if (tid < 8) sh.tabM[TK_PIECES + tid] = 0;
  __syncthreads();

  // ---- phase 1b: newline positions
  const int NW = TK_PER_THREAD / 2;                                  // two 16-bit piece masks per register
  uint32_t nlm[NW]; uint32_t cnt = 0;
#pragma unroll
  for (int q = 0; q < NW / 4; q++) { uint4 c = reinterpret_cast<const uint4*>(sh.nlM)[tid * (NW / 4) + q]; nlm[4 * q] = c.x; nlm[4 * q + 1] = c.y; nlm[4 * q + 2] = c.z; nlm[4 * q + 3] = c.w; }
#pragma unroll
  for (int h = 0; h < NW; h++) cnt += __popc(nlm[h]);
  uint32_t x = cnt;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= (uint32_t)d) x += y; }
  if (lane == 31) sh.warpTot[warp] = x;
  __syncthreads();
  uint32_t wbase = 0, totNl = 0;
#pragma unroll
  for (int w = 0; w < TK_THREADS / 32; w++) { uint32_t v = sh.warpTot[w]; if (w < (int)warp) wbase += v; totNl += v; }
  uint32_t l = wbase + x - cnt;                                      // newlines before this thread's bytes
  // the newline at the tile's last owned byte starts the NEXT tile's first line
  uint32_t lastBit = __shfl_sync(0xffffffffu, (nlm[NW - 1] >> 31) & 1u, (TK_FIRST_OV_THREAD - 1) & 31);
  const bool fits = totNl <= (uint32_t)TK_NL_CAP;
  unsigned long long* st = status + tile;
  if (tid == TK_FIRST_OV_THREAD) {
    uint32_t own = l - lastBit; sh.nOwned = own;
    *(volatile unsigned long long*)st = (tile == 0 ? TK_FLAG_PREFIX : TK_FLAG_AGG) | own;
  }
  if (fits) {
#pragma unroll
    for (int h = 0; h < NW; h++) {
      uint32_t m = nlm[h];
      while (m) { int bit = __ffs((int)m) - 1; m &= m - 1; sh.nlPos[l++] = (uint16_t)((tid * TK_PER_THREAD + 2 * h) * 16 + bit); }
    }
  }
  __syncthreads();                                                   // nlPos[], nOwned complete
  const uint32_t nOwned = sh.nOwned;
  // Decoupled look-back by the last warp, 32 predecessors per step; the other warps already parse their lines and only wait
  // (on an mbarrier) for the tile's first line number when they are about to store.
  if (warp == TK_THREADS / 32 - 1) {
    unsigned long long sum = 0;
    if (tile) {
      long long i = (long long)tile - 1;
      while (true) {
        long long idx = i - lane; unsigned long long v = TK_FLAG_PREFIX;                    // in front of tile 0: prefix 0
        if (idx >= 0) do { v = *(volatile unsigned long long*)(status + idx); } while (!(v >> 62));
        uint32_t pm = __ballot_sync(0xffffffffu, (v >> 62) == 2);
        uint32_t firstP = pm ? (uint32_t)(__ffs((int)pm) - 1) : 32u;
        uint32_t c = lane <= firstP ? (uint32_t)(v & TK_VALUE_MASK) : 0u;
        sum += __reduce_add_sync(0xffffffffu, c);
        if (pm) break;
        i -= 32;
      }
      if (lane == 0) *(volatile unsigned long long*)st = TK_FLAG_PREFIX | (sum + nOwned);
    }
    if (lane == 0) {
      sh.lineBase = sum;
      if (tile == nTiles - 1) *nlTotal = sum + nOwned;
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tk_smem_addr(&sh.barBase)) : "memory");
    }
  }
  auto line_base = [&]() -> unsigned long long {                     // blocks until the look-back has finished
    uint32_t done = 0, bar = tk_smem_addr(&sh.barBase);
    while (!done) asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(bar) : "memory");
    return *(volatile unsigned long long*)&sh.lineBase;
  };
  if (!fits) { if (tid == 0) tk_slow_tile(A, S, capLines, vb, a0, line_base(), tile == 0, err, fragMinMax); return; }

  // ---- phase 2: one thread per owned line
  const bool reachesEof = v0 + TK_STAGE >= vEnd;
  uint32_t fmin = 0xFFFFFFFFu, fmax = 0;                             // fragment id range of this thread's records
  // (ncu shows this loop peeled into two ~4500-instruction bodies and "no instruction" stalls of 1.3-1.7 warps per issue slot; forcing one
  // copy with `#pragma unroll 1` and real calls for the cold parser shrank the kernel from 23k to 8k instructions but raised it from 80 to
  // 96 registers and was SLOWER, 38.1 vs 36.2 ms per chunk, profiles/r03f: EQP_VAR_ONE_BODY keeps that variant for A/B runs.)
#ifdef EQP_VAR_ONE_BODY
#pragma unroll 1
#endif
  for (uint32_t j = tid; j < nOwned; j += TK_THREADS) {
    int s = (int)sh.nlPos[j] + 1;                                    // staged position of the line's first byte
    uint64_t g = (uint64_t)(v0 + s - a0);
    int n = -1; bool eofLine = false;
    if (j + 1 < totNl) n = (int)sh.nlPos[j + 1] - s;
    else if (reachesEof) { n = (int)(vEnd - v0) - s; eofLine = true; }
    DecodedRec rec{}; bool parsed = false;
    if (g < S.len && n >= 0) { StageSrc src(sh.tabM, s, n); rec = parse_record_t(A, sh.stage + s, n, src); parsed = true; }
    unsigned long long k = line_base() + j;
    if (k <= capLines) S.lineOff[k] = g;
    if (g >= S.len) continue;                                        // the text's final newline opens no line
    if (parsed) {
      if (k < capLines) store_record(S, (uint32_t)k, rec, err);
      if (rec.rf & RF_REC) { fmin = min(fmin, rec.frag); fmax = max(fmax, rec.frag); }
      if (eofLine && k + 1 <= capLines) S.lineOff[k + 1] = S.len + 1;
    } else {                                                         // line runs past the overhang: generic decoder on global memory
      uint64_t eg = g; while (eg < S.len && S.text[eg] != '\n') eg++;
      if (k < capLines) { decode_record(A, S, (uint32_t)k, S.text + g, (int)(eg - g), err); if (S.rf[k] & RF_REC) { fmin = min(fmin, S.frag[k]); fmax = max(fmax, S.frag[k]); } }
      if (eg == S.len && k + 1 <= capLines) S.lineOff[k + 1] = S.len + 1;
    }
  }
  fmin = __reduce_min_sync(0xffffffffu, fmin); fmax = __reduce_max_sync(0xffffffffu, fmax);
  if (lane == 0 && fmin <= fmax) { atomicMin(fragMinMax, fmin); atomicMax(fragMinMax + 1, fmax); }
}

}  // namespace eqpb
