// a1/a2: SAM line decode into structure-of-arrays records (per-line device function).
// Reference: J/sam-source/SamRecord.java:121-271 (field split, QNAME handling, flags),
// :617-640 (flag read index), :960-1018 (flag helpers), :48,1052-1068 (NM:i: search).
//
// decode_record_t is templated on a "source" that (a) yields the line's tab positions in increasing order and (b) loads
// unaligned 4-byte groups of the line.  ScanSrc searches the text (any memory, used by the host emulation and the kernel's
// slow paths); the tokenizer kernel passes a source backed by the tab masks it has already built in shared memory.
#pragma once
#include "eqp_annotation.hpp"

namespace eqpb {

EQP_HD void raise(uint32_t* err, uint32_t code, uint32_t detail) {
#if defined(__CUDA_ARCH__)
  if (atomicCAS(&err[0], 0u, code) == 0u) err[1] = detail;
#else
  if (err[0] == 0) { err[0] = code; err[1] = detail; }
#endif
}

EQP_HD int eqp_popc32(uint32_t m) {
#if defined(__CUDA_ARCH__)
  return __popc(m);
#else
  return __builtin_popcount(m);
#endif
}

// RNAME / RNEXT -> dictionary id.  Hash and comparison run over 4-byte groups (NameHash); the string pool keeps every name
// 4-byte aligned and zero-padded, so equal groups + equal length is string equality.
template <class Src>
EQP_HD int32_t lookup_ref(const AnnoView& A, const char* p, int n) {
  int32_t res = -1;
  if (!(n == 1 && p[0] == '*')) {
    NameHash H; int ng = (n + 3) >> 2;
    for (int k = 0; k < ng; k++) H.add(Src::load4(p + 4 * k, n - 4 * k));
    uint64_t h = H.finish(n), key = dict_key(h, n); uint32_t s = (uint32_t)h & A.dictMask;
    res = -2 - (int32_t)(h & 0x3fffffffULL);            // unknown name: identity by hash only (never joins, never whitelisted)
    bool go = n < 0x10000;
    while (go) {
      DictEntry e = A.dict[s];                                           // one 16-byte load: key (hash | length), id, string offset
      if (!e.key) go = false;
      else {
        if (e.key == key) {
          const uint32_t* q = reinterpret_cast<const uint32_t*>(A.strPool + e.off); uint32_t diff = 0;
          for (int k = 0; k < ng; k++) diff |= q[k] ^ Src::load4(p + 4 * k, n - 4 * k);   // no early exit: the loads are independent and overlap
          if (!diff) { res = e.val; go = false; }
        }
        s = (s + 1) & A.dictMask;
      }
    }
  }
  return res;
}

// Integer.parseInt on [p, p+n): optional sign, digits only.  Single exit so that the lanes of a warp reconverge after the call.
EQP_HD bool parse_int(const char* p, int n, int32_t& out) {
  bool ok = n > 0, neg = false; int i = 0;
  if (ok) { char c0 = p[0]; if (c0 == '-' || c0 == '+') { neg = c0 == '-'; i = 1; ok = n > 1; } }
  int64_t v = 0;
  for (; ok && i < n; i++) { uint32_t c = (uint32_t)(uint8_t)p[i] - (uint32_t)'0'; v = v * 10 + (int64_t)c; ok = c <= 9u && v <= 2147483648LL; }
  out = (int32_t)(neg ? -v : v);
  return ok;
}

// four ASCII digits in a little-endian group (first digit in the low byte) -> value, or -1
EQP_HD int32_t digits4(uint32_t x) {
  uint32_t t = x - 0x30303030u;
  bool ok = (((t + 0x76767676u) | t) & 0x80808080u) == 0;                // every byte 0..9 (a byte below '0' borrows, but is flagged itself)
  uint32_t u = t * 10u + (t >> 8);                                       // byte 0: d0*10+d1, byte 2: d2*10+d3
  return ok ? (int32_t)((u & 0xFFu) * 100u + ((u >> 16) & 0xFFu)) : -1;
}

// rf layout: bit0 record, bit1 mapped, bit2 hasMate, bits3-4 readIndex, bits5-6 /1 /2 name suffix
EQP_HD int flag_read_index(int flag, int suffix, int readIndex) {     // SamRecord.java:621-640
  if (flag & 1) { if (flag & 64) return 1; if (flag & 128) return 2; return 1; }
  if (suffix) return suffix;
  if (readIndex > 0) return readIndex;
  return 1;
}

struct alignas(16) Word4 { uint32_t x, y, z, w; };
EQP_HD uint32_t tab_marker(uint32_t w) { uint32_t x = w ^ 0x09090909u; return (x - 0x01010101u) & ~x & 0x80808080u; }   // lowest set marker is exact
EQP_HD int first_marker(uint32_t z) {
#if defined(__CUDA_ARCH__)
  return (__ffs((int)z) - 1) >> 3;
#else
  return __builtin_ctz(z) >> 3;
#endif
}
// first '\t' in p[from, n), or n.  Aligned 4-byte loads up to the next 16-byte boundary, then aligned 16-byte loads.
// Only aligned words / quads holding at least one byte of [from, n) are read, so no other page is ever touched.
EQP_HD int find_tab(const char* p, int from, int n) {
  // Single exit on purpose: with several `return`s inside the loops the compiler duplicated the caller's continuation per exit
  // and the lanes of a warp never reconverged (ncu: 1.3 active threads per instruction).
  int res = n;
  if (from < n) {
    const char* q = p + from; int mis = (int)((uintptr_t)q & 3);
    const uint32_t* w = reinterpret_cast<const uint32_t*>(q - mis);
    int idx = from - mis;                                               // index (relative to p) of byte 0 of *w
    uint32_t x = *w ^ 0x09090909u;
    x |= (mis ? (0xFFFFFFFFu >> (32 - 8 * mis)) : 0u);                   // bytes before `from` cannot match
    uint32_t z = (x - 0x01010101u) & ~x & 0x80808080u;
    while (!z) {                                                        // single words up to the next 16-byte boundary
      idx += 4; ++w;
      if (idx >= n || !((uintptr_t)w & 15)) break;
      z = tab_marker(*w);
    }
    if (!z && idx < n) {                                                // 16 bytes per step
      const Word4* v = reinterpret_cast<const Word4*>(w);
      uint32_t z0, z1, z2, z3;
      while (true) {
        Word4 c = *v; z0 = tab_marker(c.x); z1 = tab_marker(c.y); z2 = tab_marker(c.z); z3 = tab_marker(c.w);
        if ((z0 | z1 | z2 | z3) || idx + 16 >= n) break;
        idx += 16; ++v;
      }
      int add = 0; z = z0;
      if (!z) { z = z1; add = 4; }
      if (!z) { z = z2; add = 8; }
      if (!z) { z = z3; add = 12; }
      idx += add;
    }
    if (z) { int k = idx + first_marker(z); res = k < n ? k : n; }
  }
  return res;
}

// number of ':' in p[a, b): aligned 4-byte words, exact SWAR byte compare.  Only words holding at least one byte of [a, b) are read.
EQP_HD uint32_t colon_marker(uint32_t w) { uint32_t x = w ^ 0x3a3a3a3au; return ~(((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x | 0x7f7f7f7fu); }
EQP_HD int colon_count(const char* p, int a, int b) {
  int cnt = 0;
  if (a < b) {
    uintptr_t s = (uintptr_t)(p + a), e = (uintptr_t)(p + b), wa = s & ~(uintptr_t)3, we = (e - 1) & ~(uintptr_t)3;
    uint32_t m0 = 0xFFFFFFFFu << (8 * (int)(s - wa)), m1 = 0xFFFFFFFFu >> (8 * (int)(we + 4 - e));
    uint32_t z = colon_marker(*reinterpret_cast<const uint32_t*>(wa)) & m0;
    for (; wa < we; wa += 4) { cnt += eqp_popc32(z); z = colon_marker(*reinterpret_cast<const uint32_t*>(wa + 4)); }
    cnt += eqp_popc32(z & m1);
  }
  return cnt;
}

// General tag-block scan (any token shape): field count with the reference's rule, first [nN]M:i:<digits>, HI:i: at a token start
// (hiAt = offset of its value, the last such token wins as in the reference's loop).  Only reached for lines whose tags are not all of the
// form XX:T:value.
EQP_HD void tags_general(const char* p, int t0, int n, int& fields, int& nm, int& hiAt) {
  fields = 11; nm = -1; int k = t0; bool colon = false, any = k < n;
  for (; k < n; k++) {
    char c = p[k];
    if (c == '\t') { if (!colon) fields++; colon = false; any = k + 1 < n; }
    else if (c == ':') {
      colon = true;
      if (k + 2 < n && p[k + 1] == 'i' && p[k + 2] == ':' && k - 2 >= t0) {
        char c1 = p[k - 1], c0 = p[k - 2];
        if (nm < 0 && c1 == 'M' && (c0 == 'N' || c0 == 'n')) {
          int d = k + 3, v = 0; bool dig = false;
          while (d < n && p[d] >= '0' && p[d] <= '9' && v <= 30000) { v = v * 10 + (p[d] - '0'); d++; dig = true; }
          if (dig) nm = v;
        }
        if (c1 == 'I' && c0 == 'H' && (k - 2 == t0 || p[k - 3] == '\t')) hiAt = k + 3;
      }
    }
  }
  if (any && !colon) fields++;                                           // last token (a trailing empty token is not one)
}

// Source for text in any memory: tabs by search, groups byte by byte (never reads outside [p, p+n)).
struct ScanSrc {
  EQP_HD int next(const char* p, int from, int n) { return find_tab(p, from, n); }
  static EQP_HD uint32_t load4(const char* q, int nvalid) { return name_group_bytes(q, nvalid); }
};

// One decoded line, before it is stored: parsing does not need the line's number, so the tokenizer kernel parses while the
// line numbering (a chained scan over tiles) is still in flight.
struct DecodedRec {
  uint32_t code;                                                        // DERR_* of the first failed check, 0 = fine
  uint32_t frag; int32_t flag, pos, pnext, ref, rnext; int nm, cigOff, cigLen, seqLen; uint8_t rf; uint16_t cigM;
  int32_t hi; uint8_t hasHi;                                            // HI:i: tag (hit index), SamRecord.java:233-240
};

// One SAM line [p, p+n) (no '\n').  The text may live in shared or global memory.
template <class Src>
EQP_HD DecodedRec parse_record_t(const AnnoView& A, const char* p, int n, Src& src) {
  if (n > 0 && p[n - 1] == '\r') n--;
  uint32_t code = DERR_NONE;                                            // first failed check (single exit)
  uint32_t frag = 0xFFFFFFFFu; uint8_t rf = 0;
  int32_t flag = 0, pos = 0, pnext = 0, ref = -1, rnext = -1; int nm = -1, cigOff = 0, cigLen = 0, seqLen = 0; uint16_t cigM = 0xFFFFu;
  int32_t hiVal = -1; int hiAt = -1;
  if (n > 0 && p[0] != '@') {                                           // SamReader.java:173
    int nt = 0, from = 0;                                                // the first 11 tabs, as scalars (an indexed array would live in local memory)
#define EQP_TAB_STEP(i) int tab##i = n; if (nt == i) { int q = src.next(p, from, n); if (q < n) { tab##i = q; nt = i + 1; from = q + 1; } }
    EQP_TAB_STEP(0) EQP_TAB_STEP(1) EQP_TAB_STEP(2) EQP_TAB_STEP(3) EQP_TAB_STEP(4) EQP_TAB_STEP(5)
    EQP_TAB_STEP(6) EQP_TAB_STEP(7) EQP_TAB_STEP(8) EQP_TAB_STEP(9) EQP_TAB_STEP(10)
#undef EQP_TAB_STEP
    // field count with the reference's rule: the 11 mandatory tokens count, later tokens only when they hold no ':' (SamRecord.java:223-232).
    // The same pass over the tag block finds the first [nN]M:i:<digits> and rejects an HI:i: tag.
    int fields;
    if (nt < 10) fields = nt + (from < n ? 1 : 0);
    else if (nt == 10) fields = (from < n) ? 11 : 10;
    else {
      // tag block, one token per iteration.  bowtie2 tags are XX:T:value with no ':' in the value: then the token counts as a tag
      // (no field increment), the NM regex can only match at the token start and HI:i: is a plain prefix test.  "Every token has
      // ':' at offsets 2 and 4 and the block holds exactly two ':' per token" is that shape for all tokens at once.
      const uint32_t NMI = 'n' | ('M' << 8) | (':' << 16) | ((uint32_t)'i' << 24), HII = 'H' | ('I' << 8) | (':' << 16) | ((uint32_t)'i' << 24);
      fields = 11; bool odd = false; int t = from, nTok = 0;
      while (t < n && !odd) {
        int te = src.next(p, t, n), len = te - t;
        uint32_t w = Src::load4(p + t, len);
        if (len >= 5 && ((w >> 16) & 0xFFu) == ':' && p[t + 4] == ':') {
          nTok++;
          if (nm < 0 && (w | 0x20u) == NMI) {
            int d = t + 5, v = 0; bool dig = false;
            while (d < te && p[d] >= '0' && p[d] <= '9' && v <= 30000) { v = v * 10 + (p[d] - '0'); d++; dig = true; }
            if (dig) nm = v;
          }
          if (w == HII) hiAt = t + 5;
          t = te + 1;
        } else odd = true;
      }
      if (!odd && colon_count(p, from, n) != 2 * nTok) odd = true;
      if (odd) { hiAt = -1; tags_general(p, from, n, fields, nm, hiAt); }
      if (hiAt >= 0) {                                                    // new Integer(token.substring(5)): the rest of the token must be an integer
        int e = hiAt; while (e < n && p[e] != '\t') e++;
        if (!parse_int(p + hiAt, e - hiAt, hiVal)) code = DERR_INT;
      }
    }
    if (!code && (fields < 11 || fields > 12)) code = DERR_FIELDS;
    if (!code) {
      int e0 = tab0; int suffix = 0;
      // QNAME (SamRecord.java:155-176): F + nine digits, optionally "/1" "/2" or " comment"
      if (e0 == 10) {
        int32_t hi4 = digits4(Src::load4(p + 1, 4)), lo4 = digits4(Src::load4(p + 5, 4)); uint32_t d9 = (uint32_t)(uint8_t)p[9] - (uint32_t)'0';
        if (p[0] != 'F' || hi4 < 0 || lo4 < 0 || d9 > 9u) code = DERR_QNAME; else frag = (uint32_t)hi4 * 100000u + (uint32_t)lo4 * 10u + d9;
      } else {
        int qe = e0; bool sp = false;
        for (int i = 1; i < e0 && !sp; i++) if (p[i] == ' ') { qe = i; sp = true; }
        if (e0 >= 2 && p[e0 - 2] == '/' && (p[e0 - 1] == '1' || p[e0 - 1] == '2')) { suffix = p[e0 - 1] - '0'; if (qe == e0) qe = e0 - 2; }
        if (qe != 10 || p[0] != 'F') code = DERR_QNAME;
        else { frag = 0; for (int i = 1; i < 10; i++) { uint32_t c = (uint32_t)(uint8_t)p[i] - (uint32_t)'0'; if (c > 9u) code = DERR_QNAME; frag = frag * 10 + c; } }
      }
      int32_t mapq = 0; bool intsOk = true;
      intsOk &= parse_int(p + tab0 + 1, tab1 - tab0 - 1, flag);
      int readIndex = suffix;
      int fri = flag_read_index(flag, suffix, readIndex);
      bool riBad = intsOk && readIndex && readIndex != fri;
      if (!readIndex) readIndex = fri;
      { const char* q = p + tab2 + 1; int l = tab3 - tab2 - 1;
        if (l == 1 && q[0] == '*') pos = -1; else intsOk &= parse_int(q, l, pos); }
      intsOk &= parse_int(p + tab3 + 1, tab4 - tab3 - 1, mapq);
      { const char* q = p + tab6 + 1; int l = tab7 - tab6 - 1;
        if (l == 1 && q[0] == '*') pnext = -1; else intsOk &= parse_int(q, l, pnext); }
      if (!code && !intsOk) code = DERR_INT;
      if (!code && riBad) code = DERR_READ_INDEX;
      cigOff = tab4 + 1; cigLen = tab5 - tab4 - 1;
      { uint32_t v = 0; bool simple = cigLen >= 2 && cigLen <= 6 && p[cigOff + cigLen - 1] == 'M';       // "<n>M": the only block is [pos-1, pos-1+n)
        for (int i = 0; i + 1 < cigLen && simple; i++) { uint32_t c = (uint32_t)(uint8_t)p[cigOff + i] - (uint32_t)'0'; simple = c <= 9u; v = v * 10 + c; }
        cigM = (simple && v < 0xFFFFu) ? (uint16_t)v : (uint16_t)0xFFFFu; }
      seqLen = tab9 - tab8 - 1;
      int qualEnd = nt >= 11 ? tab10 : n, qualLen = qualEnd - tab9 - 1;
      if (!code && !(qualLen == 1 && p[tab9 + 1] == '*') && qualLen != seqLen) code = DERR_SEQ_QUAL;
      if (!code) {
        ref = lookup_ref<Src>(A, p + tab1 + 1, tab2 - tab1 - 1);
        const char* q = p + tab5 + 1; int l = tab6 - tab5 - 1;
        rnext = (l == 1 && q[0] == '=') ? ref : lookup_ref<Src>(A, q, l);   // getMateReferenceName (:888-894)
        bool mapped = pos >= 0 && (flag & 4) == 0;                          // :980-982
        bool hasMate = (flag & 1) && (flag & 2);                            // :964-966
        rf = (uint8_t)(RF_REC | (mapped ? RF_MAPPED : 0) | (hasMate ? RF_HASMATE : 0) | (readIndex << 3) | (suffix << 5));
      }
    }
  }
  if (code) { frag = 0xFFFFFFFFu; rf = 0; }
  return DecodedRec{code, frag, flag, pos, pnext, ref, rnext, nm, cigOff, cigLen, seqLen, rf, cigM, hiVal, (uint8_t)(hiAt >= 0 ? 1 : 0)};
}
EQP_HD void store_record(const SamView& S, uint32_t li, const DecodedRec& r, uint32_t* err) {
  if (r.code) raise(err, r.code, li);
  else if (r.hasHi && !S.hi) raise(err, DERR_FIELDS, li);                // HI:i: changes isMate (SamRecord.java:658-660): only accepted from a genomic aligner's SAM
  if (S.hi) S.hi[li] = r.rf ? r.hi : -1;
  S.frag[li] = r.frag; S.rf[li] = r.rf;
  if (r.rf) {
    S.flag[li] = (uint16_t)r.flag; S.ref[li] = r.ref; S.pos[li] = r.pos; S.rnext[li] = r.rnext; S.pnext[li] = r.pnext;
    S.nm[li] = (int16_t)r.nm; S.cigOff[li] = (uint16_t)r.cigOff; S.cigLen[li] = (uint16_t)r.cigLen; S.seqLen[li] = (uint16_t)r.seqLen; S.cigM[li] = r.cigM;
  }
}
template <class Src>
EQP_HD void decode_record_t(const AnnoView& A, const SamView& S, uint32_t li, const char* p, int n, Src& src, uint32_t* err) {
  store_record(S, li, parse_record_t(A, p, n, src), err);
}

#ifdef EQP_VAR_ONE_BODY
EQP_HD_CALL
#else
EQP_HD
#endif
void decode_record(const AnnoView& A, const SamView& S, uint32_t li, const char* p, int n, uint32_t* err) {
  ScanSrc src; decode_record_t(A, S, li, p, n, src, err);
}

EQP_HD void decode_line(const AnnoView& A, const SamView& S, uint32_t li, uint32_t* err) {
  uint64_t b = S.lineOff[li], e = S.lineOff[li + 1] - 1;              // [b,e): line without '\n'
  decode_record(A, S, li, S.text + b, (int)(e - b), err);
}

}  // namespace eqpb
