// a1/a2 on the B200: ONE pass over the SAM text that builds the line index and decodes every record (sm_100a).
// Replaces SamReader.readSamFile's line loop + `new SamRecord(line)` (J/sam-source/SamReader.java:139-301, SamRecord.java:121-271).
//
// A CTA owns the lines that START in a 30,704-byte tile.  The tile plus a 16-byte lead-in and a 2 KB overhang (so the last
// owned line is complete) is brought into shared memory by one TMA bulk copy (cp.async.bulk, mbarrier completion).
//   phase 1a  byte-parallel: every 16-byte piece -> 16-bit mask of control characters (< 0x10), conflict-free LDS.128 + SWAR
//   phase 1b  thread t owns 8 consecutive pieces: classify the few marked bytes into '\t' / '\n', block scan of (events, newlines),
//             emit the tile's sorted event list ev[] and the event index of every newline lineEv[]
//   look-back one thread publishes the tile's newline count and chains the exclusive prefix over earlier tiles (decoupled
//             look-back, tile ids from an atomic ticket so that every predecessor is already running) -> global line numbers
//   phase 2   thread per owned line: the record's tabs are consecutive entries of ev[], so the field decode (decode_record_t)
//             never searches; SoA stores are coalesced over consecutive line numbers
// Lines longer than the overhang and tiles with more events than the tables hold take a per-line / per-tile slow path that
// parses from global memory with the generic decoder.
#pragma once
#include <cuda_runtime.h>
#include "eqp_decode.h"

namespace eqpb {

static const int TK_THREADS = 256;
static const int TK_PIECES = 2048;                                   // 16-byte pieces staged per tile
static const int TK_STAGE = TK_PIECES * 16;                          // 32 KB
static const int TK_OV_PIECES = 128;                                 // 2 KB overhang
static const int TK_OWNED = (TK_PIECES - 1 - TK_OV_PIECES) * 16;     // 30,704 bytes of line starts per tile
static const int TK_PER_THREAD = TK_PIECES / TK_THREADS;             // 8 pieces = 128 bytes
static const int TK_EV_CAP = 6144, TK_NL_CAP = 2048;
static const int TK_FIRST_OV_THREAD = (TK_PIECES - TK_OV_PIECES) / TK_PER_THREAD;   // 240

struct TokShared {
  alignas(128) char stage[TK_STAGE];
  alignas(16) uint16_t ctrl[TK_PIECES];
  uint16_t ev[TK_EV_CAP];
  uint16_t lineEv[TK_NL_CAP + 2];
  uint32_t warpTot[TK_THREADS / 32];
  alignas(8) unsigned long long bar;
  unsigned long long lineBase;
  uint32_t tile, nEv, nNl, nOwned;
};

static const unsigned long long TK_FLAG_AGG = 1ULL << 62, TK_FLAG_PREFIX = 2ULL << 62, TK_VALUE_MASK = (1ULL << 62) - 1;

__device__ __forceinline__ uint32_t tk_smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// 4 bytes -> 4-bit mask of bytes < 0x10 (high nibble zero), exact
__device__ __forceinline__ uint32_t tk_ctrl4(uint32_t w) {
  uint32_t t = (w & 0x70707070u) + 0x70707070u;                      // bit 7 of a byte set iff bits 4-6 are not all zero
  uint32_t z = ~(t | w) & 0x80808080u;                               // 0x80 iff the byte's high nibble is zero
  return (((z >> 7) * 0x00204081u) >> 21) & 0xFu;                    // gather bits 0,8,16,24 into a nibble
}

// Serial slow path of a whole tile (thread 0): same outputs as the fast path, parsed from global memory.
__device__ __noinline__ void tk_slow_tile(const AnnoView& A, const SamView& S, uint32_t capLines, long long vb, int a0, unsigned long long lineBase, bool firstTile, uint32_t* err) {
  unsigned long long k = lineBase;
  for (long long v = vb - 1; v < vb + TK_OWNED - 1; v++) {
    long long g = v - a0;
    bool nl = (g == -1 && firstTile) || (g >= 0 && g < (long long)S.len && S.text[g] == '\n');
    if (!nl) continue;
    uint64_t st = (uint64_t)(g + 1);
    if (k <= capLines) S.lineOff[k] = st;
    if (st < S.len) {
      uint64_t e = st; while (e < S.len && S.text[e] != '\n') e++;
      if (k < capLines) decode_record(A, S, (uint32_t)k, S.text + st, (int)(e - st), err);
      if (e == S.len && k + 1 <= capLines) S.lineOff[k + 1] = S.len + 1;
    }
    k++;
  }
}

__global__ void __launch_bounds__(TK_THREADS, 4)
k_tokenize_sam(AnnoView A, SamView S, uint32_t capLines, uint32_t nTiles, unsigned long long* status, uint32_t* ticket, unsigned long long* nlTotal, uint32_t* err) {
  extern __shared__ __align__(128) unsigned char tk_raw[];
  TokShared& sh = *reinterpret_cast<TokShared*>(tk_raw);
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int a0 = (int)((uintptr_t)S.text & 15);                      // v = text offset + a0: virtual coordinate with a 16-byte aligned origin
  const char* base = S.text - a0;
  const long long vEnd = (long long)S.len + a0, vEndAl = (vEnd + 15) & ~15LL;

  if (tid == 0) {
    sh.tile = atomicAdd(ticket, 1u);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(tk_smem_addr(&sh.bar)));
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
  { uint32_t done = 0, bar = tk_smem_addr(&sh.bar);
    while (!done) asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(bar) : "memory"); }

  // ---- phase 1a: control-character mask per 16-byte piece
  const uint4* sv = reinterpret_cast<const uint4*>(sh.stage);
#pragma unroll
  for (int k = 0; k < TK_PER_THREAD; k++) {
    uint32_t p = tid + k * TK_THREADS; uint4 q = sv[p];
    uint32_t m = tk_ctrl4(q.x) | (tk_ctrl4(q.y) << 4) | (tk_ctrl4(q.z) << 8) | (tk_ctrl4(q.w) << 12);
    long long g0 = v0 + 16LL * p - a0;                               // text offset of the piece's byte 0
    if (g0 < 0 || g0 + 16 > (long long)S.len) {                      // piece straddles the text's ends: keep valid bytes only
      long long lo = g0 < 0 ? -g0 : 0, hi = (long long)S.len - g0;
      if (lo > 16) lo = 16; if (hi > 16) hi = 16; if (hi < 0) hi = 0;
      m &= (hi > lo) ? (((1u << hi) - 1u) & ~((1u << lo) - 1u)) : 0u;
    }
    if (p == 0) m &= 0x8000u;                                        // lead-in: only the byte right before the tile matters
    sh.ctrl[p] = (uint16_t)m;
  }
  __syncthreads();
  const int vnPos = 15 + a0;                                         // staged position of the virtual newline before text[0] (tile 0)
  if (tile == 0 && tid == 0) sh.ctrl[vnPos >> 4] |= (uint16_t)(1u << (vnPos & 15));
  if (tile == 0) __syncthreads();

  // ---- phase 1b: classify, scan, emit
  uint32_t tabM[TK_PER_THREAD / 2], nlM[TK_PER_THREAD / 2];          // two 16-bit piece masks per register
  uint32_t nEv = 0, nNl = 0;
  { uint4 cm = reinterpret_cast<const uint4*>(sh.ctrl)[tid]; uint32_t c[4] = {cm.x, cm.y, cm.z, cm.w};
#pragma unroll
    for (int h = 0; h < 4; h++) {
      uint32_t m = c[h], tb = 0, nl = 0;
      while (m) {
        int bit = __ffs((int)m) - 1; m &= m - 1;
        int pos = (int)(tid * TK_PER_THREAD + 2 * h) * 16 + bit;     // bit 16.. belongs to the next piece: same formula
        char ch = sh.stage[pos]; bool forced = tile == 0 && pos == vnPos;
        if (forced || ch == '\n') nl |= 1u << bit; else if (ch == '\t') tb |= 1u << bit;
      }
      tabM[h] = tb; nlM[h] = nl; nEv += __popc(tb | nl); nNl += __popc(nl);
    } }
  uint32_t packed = nEv | (nNl << 16), x = packed;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= (uint32_t)d) x += y; }
  if (lane == 31) sh.warpTot[warp] = x;
  __syncthreads();
  uint32_t wbase = 0, tot = 0;
#pragma unroll
  for (int w = 0; w < TK_THREADS / 32; w++) { uint32_t v = sh.warpTot[w]; if (w < (int)warp) wbase += v; tot += v; }
  uint32_t ex = wbase + x - packed;                                  // exclusive prefix: events (low 16), newlines (high 16)
  uint32_t e = ex & 0xFFFFu, l = ex >> 16;
  const uint32_t totEv = tot & 0xFFFFu, totNl = tot >> 16;
  // the newline at the tile's last owned byte starts the NEXT tile's first line
  uint32_t lastBit = __shfl_sync(0xffffffffu, (nlM[3] >> 31) & 1u, (TK_FIRST_OV_THREAD - 1) & 31);
  const bool fits = totEv <= (uint32_t)TK_EV_CAP && totNl <= (uint32_t)TK_NL_CAP;
  unsigned long long* st = status + tile;
  if (tid == TK_FIRST_OV_THREAD) {
    uint32_t own = l - lastBit; sh.nOwned = own; sh.nEv = totEv; sh.nNl = totNl;
    *(volatile unsigned long long*)st = (tile == 0 ? TK_FLAG_PREFIX : TK_FLAG_AGG) | own;
  }
  if (fits) {
#pragma unroll
    for (int h = 0; h < 4; h++) {
      uint32_t m = tabM[h] | nlM[h];
      while (m) {
        int bit = __ffs((int)m) - 1; m &= m - 1;
        uint32_t pos = (tid * TK_PER_THREAD + 2 * h) * 16 + bit, isNl = (nlM[h] >> bit) & 1u;
        sh.ev[e] = (uint16_t)((pos << 1) | isNl);
        if (isNl) sh.lineEv[l++] = (uint16_t)e;
        e++;
      }
    }
  }
  if (tid == TK_FIRST_OV_THREAD) {                                   // decoupled look-back
    unsigned long long sum = 0;
    if (tile) {
      for (uint32_t i = tile; i-- > 0;) {
        unsigned long long v;
        do { v = *(volatile unsigned long long*)(status + i); } while (!(v >> 62));
        sum += v & TK_VALUE_MASK;
        if ((v >> 62) == 2) break;
      }
      *(volatile unsigned long long*)st = TK_FLAG_PREFIX | (sum + sh.nOwned);
    }
    sh.lineBase = sum;
    if (tile == nTiles - 1) *nlTotal = sum + sh.nOwned;
  }
  __syncthreads();
  const unsigned long long lineBase = sh.lineBase; const uint32_t nOwned = sh.nOwned;
  if (!fits) { if (tid == 0) tk_slow_tile(A, S, capLines, vb, a0, lineBase, tile == 0, err); return; }

  // ---- phase 2: one thread per owned line
  const bool reachesEof = v0 + TK_STAGE >= vEnd;
  for (uint32_t j = tid; j < nOwned; j += TK_THREADS) {
    uint32_t evi = sh.lineEv[j]; int s = (int)(sh.ev[evi] >> 1) + 1;  // staged position of the line's first byte
    unsigned long long k = lineBase + j; uint64_t g = (uint64_t)(v0 + s - a0);
    if (k <= capLines) S.lineOff[k] = g;
    if (g >= S.len) continue;                                        // the text's final newline opens no line
    int n; int evEnd;
    if (j + 1 < totNl) { evEnd = sh.lineEv[j + 1]; n = (int)(sh.ev[evEnd] >> 1) - s; }
    else if (reachesEof) { evEnd = (int)totEv; n = (int)(vEnd - v0) - s; if (k + 1 <= capLines) S.lineOff[k + 1] = S.len + 1; }
    else {                                                           // line runs past the overhang: generic decoder on global memory
      uint64_t eg = g; while (eg < S.len && S.text[eg] != '\n') eg++;
      if (k < capLines) decode_record(A, S, (uint32_t)k, S.text + g, (int)(eg - g), err);
      if (eg == S.len && k + 1 <= capLines) S.lineOff[k + 1] = S.len + 1;
      continue;
    }
    if (k < capLines) { EventTabs tabs{sh.ev, (int)evi + 1, evEnd, s}; decode_record_t(A, S, (uint32_t)k, sh.stage + s, n, tabs, err); }
  }
}

}  // namespace eqpb
