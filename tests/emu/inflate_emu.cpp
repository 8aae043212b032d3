// Host emulation of the device inflate (eqp-cluster_b200/csrc/eqp_inflate.h): the same per-thread functions, the kernels' loops run one
// "thread" after the other, the host driver's chain walk and waves restated.  Test infrastructure (tests/test_inflate.py compares the
// output with zlib's); never linked into the library.
//   inflate_emu <in.gz> <out> [segBytes ratio segsPerWave bogusEvery maxGroups]      bogusEvery = k: every k-th found start is moved one bit on
//   (made wrong); maxGroups: CTAs of the group-parallel window resolve (default 296)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "../../eqp-cluster_b200/csrc/eqp_inflate.h"

using namespace eqpb::zf;

struct Tables { uint16_t lim[2][16]; int16_t base[2][16]; uint16_t sym[320]; Huff hl, hd;
  Tables() { hl = Huff{lim[0], base[0], sym, 1, 1}; hd = Huff{lim[1], base[1], sym + 288, 1, 1}; } };
struct Chain { uint64_t sym, abs, len; };

static int fail(const char* m) { fprintf(stderr, "ERROR: %s\n", m); return 1; }

int main(int argc, char** argv) {
  if (argc < 3) return fail("usage: inflate_emu <in.gz> <out> [segBytes ratio segsPerWave]");
  const uint64_t segBytes = argc > 3 ? strtoull(argv[3], 0, 10) : 32768, ratio = argc > 4 ? strtoull(argv[4], 0, 10) : 8;
  const uint32_t perWave = argc > 5 ? (uint32_t)strtoul(argv[5], 0, 10) : 1u << 30, bogusEvery = argc > 6 ? (uint32_t)strtoul(argv[6], 0, 10) : 0, maxGroups = argc > 7 ? (uint32_t)strtoul(argv[7], 0, 10) : 296;
  FILE* f = fopen(argv[1], "rb"); if (!f) return fail("cannot open input");
  std::vector<uint8_t> z; { uint8_t buf[1 << 16]; size_t n; while ((n = fread(buf, 1, sizeof buf, f)) > 0) z.insert(z.end(), buf, buf + n); } fclose(f);
  const uint64_t zLen = z.size();
  if (zLen < 18 || z[0] != 0x1f || z[1] != 0x8b || z[2] != 8) return fail("not a gzip file");
  const uint8_t flg = z[3]; uint64_t off = 10;
  if (flg & 4) off += 2 + (z[off] | (z[off + 1] << 8));
  if (flg & 8) { while (off < zLen && z[off]) off++; off++; }
  if (flg & 16) { while (off < zLen && z[off]) off++; off++; }
  if (flg & 2) off += 2;
  if (off + 8 >= zLen) return fail("truncated gzip file");
  const uint32_t isize = z[zLen - 4] | (z[zLen - 3] << 8) | (z[zLen - 2] << 16) | ((uint32_t)z[zLen - 1] << 24);
  z.resize(zLen + 16, 0);
  std::vector<uint32_t> words((zLen + 3) / 4 + 2); memcpy(words.data(), z.data(), zLen + 8);
  const uint32_t* w = words.data(); const uint64_t nWords = (zLen + 3) / 4 + 1, dataBit = off * 8, totalBits = (zLen - 8) * 8;
  const uint32_t nSeg = (uint32_t)((zLen - 8 - off + segBytes - 1) / segBytes);
  const uint64_t capSym = segBytes * ratio;
  Tables T;
  std::vector<uint64_t> start(nSeg, ZF_NONE); start[0] = dataBit;
  uint32_t found = 0;
  for (uint32_t i = 1; i < nSeg; i++) {                                  // k_zf_search
    uint64_t from = dataBit + (uint64_t)i * segBytes * 8, to = from + segBytes * 8; if (to > totalBits) to = totalBits;
    if (from < to) start[i] = find_block_start(w, nWords, from, to, totalBits, T.hl, T.hd);
    found += start[i] != ZF_NONE;
  }
  if (bogusEvery) for (uint32_t i = 1, k = 0; i < nSeg; i++) if (start[i] != ZF_NONE && ++k % bogusEvery == 0) start[i] += 1;
  std::vector<uint8_t> bogus(nSeg, 0), text; std::vector<SegOut> res(nSeg);
  std::vector<uint8_t> win(ZF_WINDOW, 0);
  uint64_t abs = 0, endBit = 0; uint32_t nBogus = 0, nWaves = 0;
  // counting pass (k_zf_decode<false> over every segment with a start), then the chain walk
  auto count = [&](uint32_t seg) { Sink s; s.out = nullptr; s.o = 0; s.cap = 0; s.window = seg != 0; s.status = 0;
    res[seg] = inflate_segment<false>(w, nWords, totalBits, start.data(), nSeg, seg, bogus.data(), T.hl, T.hd, s); };
  for (uint32_t i = 0; i < nSeg; i++) if (start[i] != ZF_NONE) count(i);
  std::vector<Chain> chain; std::vector<uint32_t> chainSeg; bool done = false;
  for (uint32_t cur = 0; !done;) {
    if (cur >= nSeg) return fail("deflate stream ends without a final block");
    const SegOut r = res[cur];
    if (r.status == ZF_RETRY) { if (r.endSeg >= nSeg || bogus[r.endSeg]) return fail("inconsistent block chain"); bogus[r.endSeg] = 1; nBogus++; count(cur); continue; }
    if (r.status != ZF_OK && r.status != ZF_FINAL) { fprintf(stderr, "segment %u status %d\n", cur, r.status); return fail("invalid deflate data"); }
    chain.push_back(Chain{0, abs, r.outLen}); chainSeg.push_back(cur); abs += r.outLen;
    if (r.status == ZF_FINAL) { done = true; endBit = r.endBit; } else cur = r.endSeg;
  }
  const uint32_t nChain = (uint32_t)chain.size();
  text.resize(abs);
  // writing pass in waves of at most scratchSyms symbols (perWave segments' worth here), tightly packed
  const uint64_t scratchSyms = std::max<uint64_t>((uint64_t)perWave * capSym, 1);
  for (uint32_t c0 = 0; c0 < nChain;) {
    uint32_t c1 = c0; uint64_t used = 0;
    while (c1 < nChain && (c1 == c0 || used + chain[c1].len <= scratchSyms)) { chain[c1].sym = used; used += chain[c1].len; c1++; }
    nWaves++;
    std::vector<uint16_t> scratch(used + 1);
    for (uint32_t c = c0; c < c1; c++) {                                 // k_zf_decode<true>
      Sink s; s.out = scratch.data() + chain[c].sym; s.o = 0; s.cap = chain[c].len; s.window = chainSeg[c] != 0; s.status = 0;
      SegOut r = inflate_segment<true>(w, nWords, totalBits, start.data(), nSeg, chainSeg[c], bogus.data(), T.hl, T.hd, s);
      if (r.status == ZF_NOWINDOW || r.status == ZF_BAD) return fail("invalid deflate data (distance)");
      if (r.status != res[chainSeg[c]].status || r.outLen != chain[c].len) return fail("the two passes disagree");
    }
    // k_zf_resolve_groups: per group of consecutive segments a window of 16-bit entries (a byte, or 0x8000|k = byte k of the window in front of the GROUP)
    const uint32_t n = c1 - c0, nGroups = std::min<uint32_t>(n, std::max<uint32_t>(maxGroups, 1)), perGroup = (n + nGroups - 1) / nGroups, usedGroups = (n + perGroup - 1) / perGroup;
    std::vector<std::vector<uint16_t>> F(usedGroups, std::vector<uint16_t>(ZF_WINDOW)); std::vector<std::vector<uint8_t>> init(usedGroups, std::vector<uint8_t>(ZF_WINDOW));
    for (uint32_t g = 0; g < usedGroups; g++) {
      const uint32_t g0 = c0 + g * perGroup, g1 = std::min(c1, g0 + perGroup); const uint64_t absG = chain[g0].abs; uint64_t absEnd = absG;
      std::vector<uint16_t> w16(ZF_WINDOW);
      for (uint32_t k = 0; k < ZF_WINDOW; k++) w16[(absG - ZF_WINDOW + k) & (ZF_WINDOW - 1)] = (uint16_t)(0x8000u | k);
      for (uint32_t c = g0; c < g1; c++) {
        const Chain& cs = chain[c];
        const uint32_t Tn = cs.len < ZF_WINDOW ? (uint32_t)cs.len : ZF_WINDOW; uint16_t* tail = scratch.data() + cs.sym + (cs.len - Tn); const uint64_t absTail = cs.abs + (cs.len - Tn);
        std::vector<uint16_t> r(Tn);
        for (uint32_t k = 0; k < Tn; k++) { uint16_t v = tail[k]; if (v & 0x8000u) { v = w16[(cs.abs - ZF_WINDOW + (v & 0x7FFFu)) & (ZF_WINDOW - 1)]; tail[k] = v; } r[k] = v; }
        for (uint32_t k = 0; k < Tn; k++) w16[(absTail + k) & (ZF_WINDOW - 1)] = r[k];
        absEnd = cs.abs + cs.len;
      }
      for (uint32_t k = 0; k < ZF_WINDOW; k++) F[g][k] = w16[(absEnd - ZF_WINDOW + k) & (ZF_WINDOW - 1)];
    }
    // k_zf_chain_windows: win = the 32 K bytes in front of the wave (byte k = position start - 32768 + k), one step per group
    for (uint32_t g = 0; g < usedGroups; g++) {
      init[g] = win; std::vector<uint8_t> nw(ZF_WINDOW);
      for (uint32_t k = 0; k < ZF_WINDOW; k++) { const uint16_t e = F[g][k]; nw[k] = (e & 0x8000u) ? win[e & 0x7FFFu] : (uint8_t)e; }
      win = nw;
    }
    for (int pass = 0; pass < 2; pass++) for (uint32_t c = c0; c < c1; c++) {  // k_zf_emit
      const Chain& cs = chain[c]; const uint32_t g = (c - c0) / perGroup; const uint64_t absG = chain[c0 + g * perGroup].abs;
      const uint64_t Tn = cs.len < ZF_WINDOW ? cs.len : ZF_WINDOW, hd = cs.len - Tn; const uint64_t k0 = pass == 0 ? hd : 0, k1 = pass == 0 ? cs.len : hd;
      for (uint64_t k = k0; k < k1; k++) {
        uint16_t v = scratch[cs.sym + k];
        if (v & 0x8000u) {
          const uint32_t wv = v & 0x7FFFu;
          if (pass == 0) { if (absG < ZF_WINDOW && wv < ZF_WINDOW - absG) return fail("distance in front of the stream"); v = init[g][wv]; }
          else { if (cs.abs < ZF_WINDOW && wv < ZF_WINDOW - cs.abs) return fail("distance in front of the stream"); v = text[cs.abs - ZF_WINDOW + wv]; }
        }
        text[cs.abs + k] = (uint8_t)v;
      }
    }
    c0 = c1;
  }
  if ((endBit + 7) / 8 + 8 != zLen) return fail("bytes after the first gzip member");
  if ((uint32_t)abs != isize) return fail("gzip length check failed (ISIZE)");
  { // k_zf_crc + the host's combine: sub-piece CRCs, piece CRCs, the text's CRC against the trailer
    uint32_t table[256]; for (uint32_t i = 0; i < 256; i++) table[i] = crc_table_entry(i);
    const uint64_t PIECE = (uint64_t)ZF_CRC_SUB * ZF_CRC_SUBS; const uint32_t opSub = crc_shift_op(ZF_CRC_SUB), opPiece = crc_shift_op(PIECE);
    uint32_t acc = 0;
    for (uint64_t base = 0; base < abs; base += PIECE) {
      const uint64_t plen = abs - base < PIECE ? abs - base : PIECE; uint32_t sub[ZF_CRC_SUBS];
      for (uint32_t t = 0; t < ZF_CRC_SUBS; t++) { const uint64_t lo = (uint64_t)t * ZF_CRC_SUB; sub[t] = lo < plen ? crc_bytes(table, text.data() + base + lo, plen - lo < ZF_CRC_SUB ? plen - lo : ZF_CRC_SUB) : 0; }
      acc = crc_combine(acc, crc_piece(sub, plen, opSub), plen == PIECE ? opPiece : crc_shift_op(plen));
    }
    const uint32_t want = z[zLen - 8] | (z[zLen - 7] << 8) | (z[zLen - 6] << 16) | ((uint32_t)z[zLen - 5] << 24);
    if (acc != want) { fprintf(stderr, "crc %08x, trailer %08x\n", acc, want); return fail("gzip CRC check failed"); }
    printf("crc %08x ", acc);
  }
  f = fopen(argv[2], "wb"); if (!f) return fail("cannot open output");
  if (!text.empty()) fwrite(text.data(), 1, text.size(), f);
  fclose(f);
  printf("segments %u starts_found %u chain %u bogus %u waves %u bytes %llu\n", nSeg, found, nChain, nBogus, nWaves, (unsigned long long)abs);
  return 0;
}
