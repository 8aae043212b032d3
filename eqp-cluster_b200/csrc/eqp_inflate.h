// DEFLATE (RFC 1951) decoding on the device: the `zcat X.sam.gz | java ...` step of compute-intersection-bed-file.sh:644-662 (SURVEY.md §8f
// rank 1) without the host inflate.  A gzip member is ONE bit stream whose blocks can only be found by decoding, so the work is made
// parallel speculatively (the approach of pugz, Kerbiriou & Chikhi 2019, restated for a GPU):
//   search   the compressed bytes are cut into segments; a thread scans its segment bit by bit for a plausible dynamic-block header
//            (complete code-length code, complete literal/length code, end-of-block code present, first symbols are SAM text bytes)
//   decode   a thread decodes from its start up to the start of the next segment, one 16-bit symbol per output byte: 0..255 = the byte,
//            0x8000|w = "byte w of the 32 KB window in front of this segment" (a match that reaches back past the segment start).
//            A decoder that passes the next segment's claimed start without landing on it proves that start bogus; it is dropped and
//            the decoder re-run through it, so the accepted starts form one verified chain from the stream's first bit to its last block.
//   resolve  one CTA walks the chain in order and replaces the markers in the last 32 K symbols of every segment (window in shared memory)
//   emit     all segments in parallel: bytes of the tails first, then everything else, markers read from the text already written.
// These are the per-thread functions (also compiled on the host, tests/emu inflate); the kernels are in eqp_inflate.cuh.
#pragma once
#include "eqp_types.h"

namespace eqpb {
namespace zf {

enum { ZF_OK = 0, ZF_FINAL = 1, ZF_BAD = 2, ZF_OVERFLOW = 3, ZF_RETRY = 4, ZF_NOWINDOW = 5 };
static const uint64_t ZF_NONE = ~0ull;
static const uint32_t ZF_WINDOW = 32768;

// ---- bit reader: LSB-first, 32-bit words, at least 33 valid bits after bits_fill
// The word after the buffered ones is already loaded (nxt): a refill never waits for memory, the load it issues is used a few symbols later.
struct Bits { const uint32_t* w; uint64_t nWords, wi; uint64_t buf; uint32_t nxt; int cnt; };
EQP_HD uint32_t bits_word(const Bits& b, uint64_t i) { return i < b.nWords ? b.w[i] : 0u; }
EQP_HD void bits_seek(Bits& b, uint64_t bitpos) {
  b.wi = bitpos >> 5; const uint32_t sh = (uint32_t)(bitpos & 31); const uint32_t x = bits_word(b, b.wi);
  b.buf = (uint64_t)(x >> sh); b.cnt = 32 - (int)sh; b.wi++; b.nxt = bits_word(b, b.wi);
}
EQP_HD void bits_fill(Bits& b) { if (b.cnt <= 32) { b.buf |= (uint64_t)b.nxt << b.cnt; b.cnt += 32; b.wi++; b.nxt = bits_word(b, b.wi); } }
EQP_HD uint32_t bits_peek(const Bits& b, int n) { return (uint32_t)b.buf & ((1u << n) - 1u); }
EQP_HD void bits_drop(Bits& b, int n) { b.buf >>= n; b.cnt -= n; }
EQP_HD uint32_t bits_take(Bits& b, int n) { uint32_t v = bits_peek(b, n); bits_drop(b, n); return v; }
EQP_HD uint64_t bits_pos(const Bits& b) { return b.wi * 32 - (uint64_t)b.cnt; }
EQP_HD uint32_t rev15(uint32_t v) {                                    // the next 15 stream bits as a left-justified Huffman code
#if defined(__CUDA_ARCH__)
  return __brev(v) >> 17;
#else
  v = ((v >> 1) & 0x55555555u) | ((v & 0x55555555u) << 1); v = ((v >> 2) & 0x33333333u) | ((v & 0x33333333u) << 2);
  v = ((v >> 4) & 0x0F0F0F0Fu) | ((v & 0x0F0F0F0Fu) << 4); v = ((v >> 8) & 0x00FF00FFu) | ((v & 0x00FF00FFu) << 8);
  v = (v >> 16) | (v << 16); return v >> 17;
#endif
}

// ---- canonical Huffman code: limit[L] = left-justified (15-bit) end of the codes of length L, base[L] = index of the first symbol of
// length L minus its first code; limit[0] = shortest length in use.  Element k of an array sits at [k * stride] (shared memory is
// interleaved over the CTA's threads).
struct Huff { uint16_t* limit; int16_t* base; uint16_t* sym; int stride, symStride; };   // symStride: the symbol table may live elsewhere (global memory, one contiguous table per thread)
// returns false when over-subscribed; `complete`: Kraft sum is exactly one; `maxLen`: longest code
EQP_HD bool huff_build(const uint8_t* lens, int n, const Huff& h, bool& complete, int& maxLen) {
  uint16_t count[16]; for (int i = 0; i < 16; i++) count[i] = 0;
  for (int i = 0; i < n; i++) count[lens[i]]++;
  uint16_t offs[16]; uint32_t code = 0, idx = 0; int minLen = 0; maxLen = 0; uint32_t kraft = 0;
  for (int L = 1; L <= 15; L++) {
    if (code + count[L] > (1u << L)) return false;
    h.limit[L * h.stride] = (uint16_t)((code + count[L]) << (15 - L));
    h.base[L * h.stride] = (int16_t)((int)idx - (int)code);
    offs[L] = (uint16_t)idx; idx += count[L]; kraft += (uint32_t)count[L] << (15 - L);
    if (count[L]) { if (!minLen) minLen = L; maxLen = L; }
    code = (code + count[L]) << 1;
  }
  h.limit[0] = (uint16_t)(minLen ? minLen : 1);
  for (int i = 0; i < n; i++) if (lens[i]) h.sym[(offs[lens[i]]++) * h.symStride] = (uint16_t)i;
  complete = kraft == 32768u;
  return true;
}
// Fixed cost for every lane of a warp: the limits are non-decreasing in L, so the code's length is one more than the number of limits
// that are <= c — four probes of a binary search instead of a walk over the lengths (which costs the warp its slowest lane's walk).
EQP_HD int huff_decode(Bits& b, const Huff& h) {
  const uint32_t c = rev15((uint32_t)b.buf & 0x7FFFu);
  int L = c >= h.limit[8 * h.stride] ? 8 : 0;
  if (c >= h.limit[(L + 4) * h.stride]) L += 4;
  if (c >= h.limit[(L + 2) * h.stride]) L += 2;
  if (c >= h.limit[(L + 1) * h.stride]) L += 1;
  L += 1;
  if (L > 15) return -1;
  bits_drop(b, L);
  return h.sym[((int)h.base[L * h.stride] + (int)(c >> (15 - L))) * h.symStride];
}

// ---- block headers
// dynamic block (after the 3 header bits): code lengths -> the two codes.  zlib's rules (inftrees.c): an over-subscribed code is an error, an
// incomplete one too unless it has a single code of length 1; the end-of-block symbol must have a code.  strict (the search): every code complete.
EQP_HD bool read_dynamic(Bits& b, const Huff& hl, const Huff& hd, bool strict) {
  bits_fill(b);
  const int hlit = (int)bits_take(b, 5) + 257, hdist = (int)bits_take(b, 5) + 1, hclen = (int)bits_take(b, 4) + 4;
  if (hlit > 286 || hdist > 30) return false;
  uint8_t cl[19]; for (int i = 0; i < 19; i++) cl[i] = 0;
  const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
  uint32_t kraft = 0;
  for (int i = 0; i < hclen; i++) { bits_fill(b); uint8_t v = (uint8_t)bits_take(b, 3); cl[order[i]] = v; if (v) kraft += 128u >> v; }
  if (kraft != 128u) return false;                                     // inflate_table(CODES): incomplete or over-subscribed set
  uint16_t cLimit[16]; int16_t cBase[16]; uint16_t cSym[19]; Huff hc{cLimit, cBase, cSym, 1, 1};
  bool complete; int maxLen;
  if (!huff_build(cl, 19, hc, complete, maxLen)) return false;
  uint8_t lens[320]; int i = 0; const int total = hlit + hdist;
  uint32_t kl = 0;                                                      // running Kraft sum of the literal/length code: a wrong start over-subscribes it within a few lengths
  while (i < total) {
    bits_fill(b);
    int s = huff_decode(b, hc); if (s < 0) return false;
    if (s < 16) { if (s && i < hlit) { kl += 32768u >> s; if (kl > 32768u) return false; } lens[i++] = (uint8_t)s; continue; }
    int rep; uint8_t val = 0;
    if (s == 16) { if (!i) return false; val = lens[i - 1]; rep = 3 + (int)bits_take(b, 2); }
    else if (s == 17) rep = 3 + (int)bits_take(b, 3);
    else rep = 11 + (int)bits_take(b, 7);
    if (i + rep > total) return false;
    if (val) for (int r = 0; r < rep; r++) if (i + r < hlit) kl += 32768u >> val;
    if (kl > 32768u) return false;
    while (rep--) lens[i++] = val;
  }
  if (!lens[256]) return false;                                        // "invalid code -- missing end-of-block"
  if (!huff_build(lens, hlit, hl, complete, maxLen) || (!complete && (strict || maxLen != 1))) return false;
  if (!huff_build(lens + hlit, hdist, hd, complete, maxLen) || (!complete && (strict || maxLen > 1))) return false;
  return true;
}
EQP_HD void read_fixed(const Huff& hl, const Huff& hd) {
  uint8_t lens[288]; bool c; int m;
  for (int i = 0; i < 288; i++) lens[i] = (uint8_t)(i < 144 ? 8 : i < 256 ? 9 : i < 280 ? 7 : 8);
  huff_build(lens, 288, hl, c, m);
  for (int i = 0; i < 32; i++) lens[i] = 5;
  huff_build(lens, 32, hd, c, m);
}

// ---- symbol sink of one segment: 16-bit symbols in global memory, markers for matches that reach in front of the segment
struct Sink {
  uint16_t* out; uint64_t o, cap; bool window;                        // window: a window exists in front of the segment (false: first segment of the stream)
  int status;
};
// WRITE = false: the counting pass (lengths and block boundaries only; nothing is loaded or stored)
template <bool WRITE> EQP_HD void sink_lit(Sink& s, uint32_t v) {
  if (WRITE) { if (s.o < s.cap) s.out[s.o] = (uint16_t)v; else s.status = ZF_OVERFLOW; }
  s.o++;
}
// the search's probe stores nothing; its literals must be bytes of SAM text
EQP_HD bool sam_text_byte(uint32_t v) { return v == '\t' || v == '\n' || v == '\r' || (v >= 0x20 && v < 0x7F); }

EQP_HD uint32_t len_base(int sym, int& extra) {                        // length symbol 257..285
  if (sym < 265) { extra = 0; return (uint32_t)(sym - 254); }
  if (sym == 285) { extra = 0; return 258; }
  extra = (sym - 261) >> 2; return 3u + ((4u + (uint32_t)((sym - 265) & 3)) << extra);
}
EQP_HD uint32_t dist_base(int sym, int& extra) {                       // distance symbol 0..29
  if (sym < 4) { extra = 0; return (uint32_t)(sym + 1); }
  extra = (sym >> 1) - 1; return 1u + ((2u + (uint32_t)(sym & 1)) << extra);
}

// the first symbols of a candidate block: valid codes, literals are text, distances stay inside what could exist
EQP_HD bool probe_symbols(Bits& b, const Huff& hl, const Huff& hd, uint32_t maxSyms) {
  uint64_t o = 0;
  for (uint32_t k = 0; k < maxSyms; k++) {
    bits_fill(b);
    int sym = huff_decode(b, hl);
    if (sym < 0 || sym > 285) return false;
    if (sym < 256) { if (!sam_text_byte((uint32_t)sym)) return false; o++; continue; }
    if (sym == 256) return k > 0;                                      // an empty block is legal but not worth starting on
    int ex; uint32_t len = len_base(sym, ex); len += bits_take(b, ex);
    bits_fill(b);
    int ds = huff_decode(b, hd);
    if (ds < 0 || ds > 29) return false;
    uint32_t dist = dist_base(ds, ex); dist += bits_take(b, ex);
    if (dist > o + ZF_WINDOW) return false;
    o += len;
  }
  return true;
}

// ---- search: first bit position in [from, to) at which a non-final dynamic block plausibly starts (ZF_NONE: none).
// Three sieves: (1) the 13 header bits — BFINAL = 0, BTYPE = 2, HLIT <= 29, HDIST <= 29 — tested for 48 positions at once with
// bitwise operations on a 64-bit window (4 % of the positions of real gzip data pass); (2) the code-length code must be complete:
// Kraft sum of its (HCLEN + 4) three-bit lengths from a 512-entry table, three lengths per lookup (0.4 % of those pass);
// (3) the whole header, strict, and the first 96 symbols.  (1)+(2) are search_scan, (3) is search_check: the kernel lets lanes
// with a candidate wait until several have one, so that the long check runs for many lanes at once.
EQP_HD uint64_t bits64(const uint32_t* w, uint64_t nWords, uint64_t p) {
  const uint64_t i = p >> 5; const uint32_t sh = (uint32_t)(p & 31);
  const uint32_t a = i < nWords ? w[i] : 0u, b = i + 1 < nWords ? w[i + 1] : 0u, c = i + 2 < nWords ? w[i + 2] : 0u;
  const uint64_t lo = ((uint64_t)b << 32) | a;
  return sh ? (lo >> sh) | ((uint64_t)c << (64 - sh)) : lo;
}
EQP_HD uint8_t kraft3(uint32_t x) {                                   // table entry: three 3-bit code lengths -> sum of 2^(7 - len) over the non-zero ones
  uint32_t k = 0; for (int i = 0; i < 3; i++) { const uint32_t v = (x >> (3 * i)) & 7u; k += (128u >> v) & 127u; } return (uint8_t)k;
}
// One bounded sub-step per call (the kernel puts a warp barrier between calls): either load the next 48-position window and sieve it (1), or
// take ONE candidate of the current window through the Kraft test (2).  (ncu on the version that looped over a window's candidates inside
// the call, profiles/r05k: the lanes drifted apart inside the loop, its body ran 11.6 times per window with 7 of 32 lanes — 53 % of the kernel.)
struct Search { uint64_t p, to, cand, found, x, m; uint32_t span; bool pending, done; };
EQP_HD void search_init(Search& s, uint64_t from, uint64_t to) { s.p = from; s.to = to; s.cand = 0; s.found = ZF_NONE; s.x = s.m = 0; s.span = 0; s.pending = false; s.done = from >= to; }
EQP_HD void search_advance(Search& s) { s.p += s.span; s.span = 0; if (s.p >= s.to) s.done = true; }   // the current window is exhausted
EQP_HD void search_scan(Search& s, const uint32_t* w, uint64_t nWords, const uint8_t* lut) {
  if (!s.m) {
    const uint64_t x = bits64(w, nWords, s.p);
    uint64_t m = ~x & ~(x >> 1) & (x >> 2) & ~((x >> 4) & (x >> 5) & (x >> 6) & (x >> 7)) & ~((x >> 9) & (x >> 10) & (x >> 11) & (x >> 12));
    const uint64_t span = s.to - s.p < 48 ? s.to - s.p : 48;
    m &= (1ull << span) - 1ull;
    s.x = x; s.m = m; s.span = (uint32_t)span;
    if (!m) search_advance(s);
    return;
  }
#if defined(__CUDA_ARCH__)
  const int k = __ffsll((long long)s.m) - 1;
#else
  const int k = __builtin_ctzll(s.m);
#endif
  s.m &= s.m - 1;
  const int hclen = (int)((s.x >> (k + 13)) & 15u) + 4;
  const uint64_t y = bits64(w, nWords, s.p + k + 17) & ((1ull << (3 * hclen)) - 1ull);
  const uint32_t sum = (uint32_t)lut[y & 511u] + lut[(y >> 9) & 511u] + lut[(y >> 18) & 511u] + lut[(y >> 27) & 511u] + lut[(y >> 36) & 511u] + lut[(y >> 45) & 511u] + lut[(y >> 54) & 511u];
  if (sum == 128u) { s.pending = true; s.cand = s.p + k; return; }
  if (!s.m) search_advance(s);
}
EQP_HD void search_check(Search& s, const uint32_t* w, uint64_t nWords, uint64_t totalBits, const Huff& hl, const Huff& hd) {
  Bits c; c.w = w; c.nWords = nWords; bits_seek(c, s.cand + 3);
  if (read_dynamic(c, hl, hd, true) && probe_symbols(c, hl, hd, 96) && bits_pos(c) <= totalBits) { s.found = s.cand; s.done = true; return; }
  s.pending = false;
  if (!s.m) search_advance(s);
}
EQP_HD uint64_t find_block_start(const uint32_t* w, uint64_t nWords, uint64_t from, uint64_t to, uint64_t totalBits, const Huff& hl, const Huff& hd) {
  uint8_t lut[512]; for (uint32_t i = 0; i < 512; i++) lut[i] = kraft3(i);
  Search s; search_init(s, from, to);
  while (!s.done) { if (s.pending) search_check(s, w, nWords, totalBits, hl, hd); else search_scan(s, w, nWords, lut); }
  return s.found;
}

// ---- decode of one segment: blocks from `start` until a block boundary coincides with the start of a later segment (ZF_OK, endSeg = that
// segment), the final block ends (ZF_FINAL), or something is wrong.  segStart[j] == ZF_NONE: segment j has no start (nothing found);
// bogus[j]: its claimed start was proven wrong.  A boundary that passes a claimed start stops the decode with ZF_RETRY and endSeg = that
// segment: the host, walking the chain from the stream's first (genuine) segment, marks it bogus and re-runs this segment.  Only chain
// members may accuse: a bogus segment decodes garbage and would pass genuine starts.
// A state machine, one bounded step at a time (see Search): block boundary / header, up to two symbols, up to eight copied symbols,
// up to eight stored bytes.
struct SegOut { uint64_t endBit, outLen; uint32_t endSeg; int status; };
enum { ST_BOUNDARY = 0, ST_SYMBOL = 1, ST_COPY = 2, ST_STORED = 3, ST_DONE = 4, ST_DIST = 5 };
#if defined(__CUDA_ARCH__)
#define EQP_WARP_MEET() __syncwarp()
#else
#define EQP_WARP_MEET()
#endif
struct Dec { Bits b; Sink s; int state; uint32_t len, dist, self, j; bool final, first; SegOut r; };
EQP_HD void dec_init(Dec& d, const uint32_t* w, uint64_t nWords, uint64_t start, uint32_t self, uint32_t nSeg, const Sink& s) {
  d.b.w = w; d.b.nWords = nWords; bits_seek(d.b, start); d.s = s; d.state = ST_BOUNDARY; d.len = d.dist = 0; d.self = self; d.j = self + 1;
  d.final = false; d.first = true; d.r.status = ZF_BAD; d.r.endSeg = nSeg; d.r.endBit = start; d.r.outLen = 0;
}
EQP_HD void dec_finish(Dec& d, int status) { d.r.status = status; d.r.endBit = bits_pos(d.b); d.r.outLen = d.s.o; d.state = ST_DONE; }
EQP_HD void dec_block_end(Dec& d, uint64_t totalBits) {
  if (bits_pos(d.b) > totalBits) dec_finish(d, ZF_BAD);
  else if (d.final) dec_finish(d, ZF_FINAL);
  else d.state = ST_BOUNDARY;
}
// up to eight symbols of the pending match: markers for the part in front of the segment, else independent loads first, then their stores
// (a symbol-by-symbol copy would wait one L2 round trip per symbol)
EQP_HD void dec_copy_batch(Dec& d) {
  Sink& s = d.s; const uint64_t o = s.o; uint32_t n = d.len < 8u ? d.len : 8u;
  if (o < d.dist) {
    const uint64_t m64 = d.dist - o; if (m64 < n) n = (uint32_t)m64;
    for (uint32_t k = 0; k < n; k++) s.out[o + k] = (uint16_t)(0x8000u | (uint32_t)(ZF_WINDOW + o + k - d.dist));
  } else {
    if (d.dist < n) n = d.dist;
    const uint16_t* src = s.out + (o - d.dist); uint16_t v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) if ((uint32_t)k < n) v[k] = src[k];
#pragma unroll
    for (int k = 0; k < 8; k++) if ((uint32_t)k < n) s.out[o + k] = v[k];
  }
  s.o += n; d.len -= n;
}
// Straight-line: (A) ONE Huffman symbol — a literal/length symbol for the lanes in ST_SYMBOL, the distance symbol of a pending match for the
// lanes in ST_DIST, through the same instructions (only the table differs) — then, after the lanes have met again, (B) one batch of a pending
// copy for the lanes in ST_COPY, then the rare states.  (ncu on the first straight-line version, profiles/r05k: the distance path ran in
// almost every step with 3 active lanes, and the copy path twice per step because the lanes that skipped (A) reached it early.)
template <bool WRITE>
EQP_HD void dec_step(Dec& d, uint64_t totalBits, const uint64_t* segStart, uint32_t nSeg, const uint8_t* bogus, const Huff& hl, const Huff& hd) {
  Bits& b = d.b; Sink& s = d.s;
  if (d.state == ST_SYMBOL || d.state == ST_DIST) {
    const bool isDist = d.state == ST_DIST;
    Huff h; h.limit = isDist ? hd.limit : hl.limit; h.base = isDist ? hd.base : hl.base; h.sym = isDist ? hd.sym : hl.sym; h.stride = hl.stride; h.symStride = hl.symStride;
    bits_fill(b);
    const int sym = huff_decode(b, h);
    if (!isDist) {
      if (sym < 256) {
        if (sym < 0) dec_finish(d, ZF_BAD);
        else { sink_lit<WRITE>(s, (uint32_t)sym); if (WRITE && s.status) dec_finish(d, s.status); }
      } else if (sym == 256) dec_block_end(d, totalBits);
      else if (sym > 285) dec_finish(d, ZF_BAD);
      else { int ex; uint32_t len = len_base(sym, ex); d.len = len + bits_take(b, ex); d.state = ST_DIST; }
    } else {
      if (sym < 0 || sym > 29) dec_finish(d, ZF_BAD);
      else {
        int ex; uint32_t dist = dist_base(sym, ex); dist += bits_take(b, ex);
        if (!WRITE) { s.o += d.len; d.state = ST_SYMBOL; }
        else if (s.o + d.len > s.cap) dec_finish(d, ZF_OVERFLOW);
        else if (dist > s.o && (dist - s.o > ZF_WINDOW || !s.window)) dec_finish(d, s.window ? ZF_BAD : ZF_NOWINDOW);
        else { d.dist = dist; d.state = ST_COPY; }
      }
    }
  }
  if (WRITE) {
    EQP_WARP_MEET();
    if (d.state == ST_COPY) { dec_copy_batch(d); if (!d.len) d.state = ST_SYMBOL; }
  }
  if (d.state == ST_SYMBOL || d.state == ST_DIST || d.state == ST_COPY || d.state == ST_DONE) return;
  if (d.state == ST_STORED) {
    for (int k = 0; k < 8 && d.len; k++, d.len--) { bits_fill(b); sink_lit<WRITE>(s, bits_take(b, 8)); }
    if (WRITE && s.status) { dec_finish(d, s.status); return; }
    if (!d.len) dec_block_end(d, totalBits);
    return;
  }
  if (d.state != ST_BOUNDARY) return;
  const uint64_t pos = bits_pos(b);
  if (!d.first) {                                                       // a block boundary after at least one block
    while (d.j < nSeg && (segStart[d.j] == ZF_NONE || bogus[d.j])) d.j++;
    if (d.j < nSeg && segStart[d.j] == pos) { d.r.endSeg = d.j; dec_finish(d, ZF_OK); return; }
    if (d.j < nSeg && segStart[d.j] < pos) { d.r.endSeg = d.j; dec_finish(d, ZF_RETRY); return; }
  }
  d.first = false;
  if (pos + 3 > totalBits) { dec_finish(d, ZF_BAD); return; }
  bits_fill(b);
  const uint32_t hdr = bits_take(b, 3); const uint32_t type = hdr >> 1; d.final = hdr & 1u;
  if (type == 0) {                                                      // stored: skip to the byte boundary, LEN, NLEN, bytes
    const uint64_t p = (bits_pos(b) + 7) & ~7ull; bits_seek(b, p); bits_fill(b);
    const uint32_t len = bits_take(b, 16); bits_fill(b); const uint32_t nlen = bits_take(b, 16);
    if ((len ^ nlen) != 0xFFFFu || p + 32 + 8ull * len > totalBits) { dec_finish(d, ZF_BAD); return; }
    d.len = len;
    if (len) d.state = ST_STORED; else dec_block_end(d, totalBits);
  } else if (type == 1) { read_fixed(hl, hd); d.state = ST_SYMBOL; }
  else if (type == 2) { if (read_dynamic(b, hl, hd, false)) d.state = ST_SYMBOL; else dec_finish(d, ZF_BAD); }
  else dec_finish(d, ZF_BAD);
}
template <bool WRITE> EQP_HD SegOut inflate_segment(const uint32_t* w, uint64_t nWords, uint64_t totalBits, const uint64_t* segStart, uint32_t nSeg, uint32_t self,
                                                    const uint8_t* bogus, const Huff& hl, const Huff& hd, Sink& s) {
  Dec d; dec_init(d, w, nWords, segStart[self], self, nSeg, s);
  while (d.state != ST_DONE) dec_step<WRITE>(d, totalBits, segStart, nSeg, bogus, hl, hd);
  s = d.s;
  return d.r;
}

// ---- CRC-32 of the inflated text (the gzip trailer's check).  The text is cut into pieces whose CRCs are computed independently and
// combined: crc(A ‖ B) = crc(A) · x^(8|B|) mod P  xor  crc(B), polynomials over GF(2) in the reflected representation — the arithmetic of
// zlib's crc32_combine_gen / crc32_combine_op (zlib 1.2.12+, crc32.c: multmodp, x2nmodp), restated.
static const uint32_t ZF_CRC_POLY = 0xEDB88320u;
EQP_HD uint32_t crc_multmodp(uint32_t a, uint32_t b) {                  // a(x) * b(x) mod P(x)
  uint32_t m = 1u << 31, p = 0;
  for (;;) {
    if (a & m) { p ^= b; if ((a & (m - 1)) == 0) break; }
    m >>= 1;
    b = (b & 1u) ? (b >> 1) ^ ZF_CRC_POLY : b >> 1;
  }
  return p;
}
EQP_HD uint32_t crc_shift_op(uint64_t len) {                            // x^(8 * len) mod P: the operator that moves a CRC over `len` following bytes
  uint32_t sq = 1u << 30, p = 1u << 31;                                 // x^1, x^0
  sq = crc_multmodp(sq, sq); sq = crc_multmodp(sq, sq); sq = crc_multmodp(sq, sq);   // x^8 = x^(8 * 2^0)
  while (len) { if (len & 1) p = crc_multmodp(sq, p); len >>= 1; if (len) sq = crc_multmodp(sq, sq); }
  return p;
}
EQP_HD uint32_t crc_combine(uint32_t crc1, uint32_t crc2, uint32_t op) { return crc_multmodp(op, crc1) ^ crc2; }
EQP_HD uint32_t crc_table_entry(uint32_t i) { uint32_t c = i; for (int k = 0; k < 8; k++) c = (c & 1u) ? ZF_CRC_POLY ^ (c >> 1) : c >> 1; return c; }
EQP_HD uint32_t crc_bytes(const uint32_t* table, const uint8_t* p, uint64_t n) {   // the standard CRC-32 (pre- and post-conditioned) of n bytes
  uint32_t c = 0xFFFFFFFFu;
  for (uint64_t i = 0; i < n; i++) c = table[(c ^ p[i]) & 0xFFu] ^ (c >> 8);
  return ~c;
}
static const uint32_t ZF_CRC_SUB = 4096, ZF_CRC_SUBS = 256;            // a piece = 256 sub-pieces of 4 KB = 1 MB
// CRC of one piece from its sub-pieces' CRCs (sub-piece i covers bytes [i * 4 KB, ...) of the piece; the last ones may be short or empty)
EQP_HD uint32_t crc_piece(const uint32_t* sub, uint64_t pieceLen, uint32_t opSub) {
  uint32_t acc = 0;
  for (uint32_t i = 0; i < ZF_CRC_SUBS; i++) {
    const uint64_t lo = (uint64_t)i * ZF_CRC_SUB; if (lo >= pieceLen) break;
    const uint64_t len = pieceLen - lo < ZF_CRC_SUB ? pieceLen - lo : ZF_CRC_SUB;
    acc = crc_combine(acc, sub[i], len == ZF_CRC_SUB ? opSub : crc_shift_op(len));
  }
  return acc;
}

}  // namespace zf
}  // namespace eqpb
