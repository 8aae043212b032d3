// a1/a2: SAM line decode into structure-of-arrays records (per-line device function).
// Reference: J/sam-source/SamRecord.java:121-271 (field split, QNAME handling, flags),
// :617-640 (flag read index), :960-1018 (flag helpers), :48,1052-1068 (NM:i: search).
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

EQP_HD int32_t lookup_ref(const AnnoView& A, const char* p, int n) {
  if (n == 1 && p[0] == '*') return -1;
  uint64_t h = fnv1a(p, n), key = dict_key(h, n); uint32_t s = (uint32_t)h & A.dictMask;
  if (n < 0x10000) while (true) {
    DictEntry e = A.dict[s];                                            // one 16-byte load: key (hash | length), id, string offset
    if (!e.key) break;
    if (e.key == key) {
      const char* q = A.strPool + e.off; int diff = 0;
      for (int i = 0; i < n; i++) diff |= q[i] ^ p[i];                   // no early exit: the loads are independent and overlap
      if (!diff) return e.val;
    }
    s = (s + 1) & A.dictMask;
  }
  return -2 - (int32_t)(h & 0x3fffffffULL);           // unknown name: identity by hash only (never joins, never whitelisted)
}

// Integer.parseInt on [p, p+n): optional sign, digits only
EQP_HD bool parse_int(const char* p, int n, int32_t& out) {
  if (n <= 0) return false;
  int i = 0; bool neg = false;
  if (p[0] == '-' || p[0] == '+') { neg = p[0] == '-'; i = 1; if (n == 1) return false; }
  int64_t v = 0;
  for (; i < n; i++) { char c = p[i]; if (c < '0' || c > '9') return false; v = v * 10 + (c - '0'); if (v > 2147483648LL) return false; }
  out = (int32_t)(neg ? -v : v); return true;
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

// General tag-block scan (any token shape): field count with the reference's rule, first [nN]M:i:<digits>, HI:i: at a token start.
// Returns false when an HI tag is present.  Only reached for lines whose tags are not all of the form XX:T:value.
EQP_HD bool tags_general(const char* p, int t0, int n, int& fields, int& nm) {
  fields = 11; nm = -1; int k = t0; bool colon = false, any = k < n;
  for (; k < n; k++) {
    char c = p[k];
    if (c == '\t') { if (!colon) fields++; colon = false; any = k + 1 < n; continue; }
    if (c == ':') {
      colon = true;
      if (k + 2 < n && p[k + 1] == 'i' && p[k + 2] == ':' && k - 2 >= t0) {
        char c1 = p[k - 1], c0 = p[k - 2];
        if (nm < 0 && c1 == 'M' && (c0 == 'N' || c0 == 'n')) {
          int d = k + 3, v = 0; bool dig = false;
          while (d < n && p[d] >= '0' && p[d] <= '9') { v = v * 10 + (p[d] - '0'); d++; dig = true; if (v > 30000) break; }
          if (dig) nm = v;
        }
        if (c1 == 'I' && c0 == 'H' && (k - 2 == t0 || p[k - 3] == '\t')) return false;
      }
    }
  }
  if (any && !colon) fields++;                                           // last token (a trailing empty token is not one)
  return true;
}

// number of ':' in p[a, b): aligned 4-byte words, exact SWAR byte compare.  Only words holding at least one byte of [a, b) are read.
EQP_HD int colon_count(const char* p, int a, int b) {
  if (a >= b) return 0;
  uintptr_t s = (uintptr_t)(p + a), e = (uintptr_t)(p + b), wa = s & ~(uintptr_t)3; int cnt = 0;
  for (; wa < e; wa += 4) {
    uint32_t x = *reinterpret_cast<const uint32_t*>(wa) ^ 0x3a3a3a3au;
    uint32_t z = ~(((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x | 0x7f7f7f7fu);   // 0x80 in every byte equal to ':'
    if (wa < s) z &= 0xFFFFFFFFu << (8 * (int)(s - wa));
    if (wa + 4 > e) z &= 0xFFFFFFFFu >> (8 * (int)(wa + 4 - e));
#if defined(__CUDA_ARCH__)
    cnt += __popc(z);
#else
    cnt += __builtin_popcount(z);
#endif
  }
  return cnt;
}

// Tab sources for decode_record_t: next(p, from, n) = first '\t' in p[from, n), or n.  Calls come with increasing `from`.
struct ScanTabs { EQP_HD int next(const char* p, int from, int n) { return find_tab(p, from, n); } };
// Tabs already located by the tokenizer kernel: ev[cur, end) are this line's tab events, (position in the staged tile << 1); s = line start.
struct EventTabs {
  const uint16_t* ev; int cur, end, s;
  EQP_HD int next(const char*, int, int n) { if (cur < end) { int q = (int)(ev[cur] >> 1) - s; cur++; return q < n ? q : n; } return n; }
};

// One SAM line [p, p+n) (no '\n').  The text may live in shared or global memory.
template <class Tabs>
EQP_HD void decode_record_t(const AnnoView& A, const SamView& S, uint32_t li, const char* p, int n, Tabs& tabs, uint32_t* err) {
  if (n > 0 && p[n - 1] == '\r') n--;
  S.frag[li] = 0xFFFFFFFFu; S.rf[li] = 0;
  if (n <= 0 || p[0] == '@') return;                                    // SamReader.java:173
  int tab[11]; int nt = 0;
  { int pos = 0; while (nt < 11) { int q = tabs.next(p, pos, n); if (q >= n) break; tab[nt++] = q; pos = q + 1; } }
  // field count with the reference's rule: the 11 mandatory tokens count, later tokens only when they hold no ':' (SamRecord.java:223-232).
  // The same pass over the tag block finds the first [nN]M:i:<digits> and rejects an HI:i: tag.
  int nm = -1; int fields;
  if (nt < 10) fields = nt + (((nt ? tab[nt - 1] + 1 : 0) < n) ? 1 : 0);
  else if (nt == 10) fields = (tab[9] + 1 < n) ? 11 : 10;
  else {
    // tag block, one token per iteration.  bowtie2 tags are XX:T:value with no ':' in the value: then the token counts as a tag
    // (no field increment), the NM regex can only match at the token start and HI:i: is a plain prefix test.  "Every token has
    // ':' at offsets 2 and 4 and the block holds exactly two ':' per token" is that shape for all tokens at once.
    fields = 11; bool odd = false, hi = false; int t = tab[10] + 1, nTok = 0;
    while (t < n) {
      int te = tabs.next(p, t, n), len = te - t;
      if (!(len >= 5 && p[t + 2] == ':' && p[t + 4] == ':')) { odd = true; break; }
      nTok++;
      bool isI = p[t + 3] == 'i';
      if (isI && p[t + 1] == 'M' && (p[t] == 'N' || p[t] == 'n') && nm < 0) {
        int d = t + 5, v = 0; bool dig = false;
        while (d < te && p[d] >= '0' && p[d] <= '9') { v = v * 10 + (p[d] - '0'); d++; dig = true; if (v > 30000) break; }
        if (dig) nm = v;
      }
      hi |= isI && p[t] == 'H' && p[t + 1] == 'I';
      t = te + 1;
    }
    if (!odd && colon_count(p, tab[10] + 1, n) != 2 * nTok) odd = true;
    if (odd) hi = !tags_general(p, tab[10] + 1, n, fields, nm);
    if (hi) { raise(err, DERR_FIELDS, li); return; }                      // HI:i: changes isMate (SamRecord.java:658-660): not produced by bowtie2
  }
  if (fields < 11 || fields > 12) { raise(err, DERR_FIELDS, li); return; }
  int e0 = tab[0];
  // QNAME (SamRecord.java:155-176)
  int qe = e0; { for (int i = 1; i < e0; i++) if (p[i] == ' ') { qe = i; break; } }
  int suffix = 0;
  if (e0 >= 2 && p[e0 - 2] == '/' && (p[e0 - 1] == '1' || p[e0 - 1] == '2')) { suffix = p[e0 - 1] - '0'; if (qe == e0) qe = e0 - 2; }
  if (qe != 10 || p[0] != 'F') { raise(err, DERR_QNAME, li); return; }
  uint32_t frag = 0;
  for (int i = 1; i < 10; i++) { char c = p[i]; if (c < '0' || c > '9') { raise(err, DERR_QNAME, li); return; } frag = frag * 10 + (uint32_t)(c - '0'); }
  int32_t flag, pos, mapq, pnext;
  if (!parse_int(p + tab[0] + 1, tab[1] - tab[0] - 1, flag)) { raise(err, DERR_INT, li); return; }
  int readIndex = suffix ? suffix : 0;
  int fri = flag_read_index(flag, suffix, readIndex);
  if (!readIndex) readIndex = fri; else if (readIndex != fri) { raise(err, DERR_READ_INDEX, li); return; }
  const char* rn = p + tab[1] + 1; int rnLen = tab[2] - tab[1] - 1;
  int32_t ref = lookup_ref(A, rn, rnLen);
  { const char* q = p + tab[2] + 1; int l = tab[3] - tab[2] - 1;
    if (l == 1 && q[0] == '*') pos = -1; else if (!parse_int(q, l, pos)) { raise(err, DERR_INT, li); return; } }
  if (!parse_int(p + tab[3] + 1, tab[4] - tab[3] - 1, mapq)) { raise(err, DERR_INT, li); return; }
  int cigOff = tab[4] + 1, cigLen = tab[5] - tab[4] - 1;
  int32_t rnext;
  { const char* q = p + tab[5] + 1; int l = tab[6] - tab[5] - 1;
    rnext = (l == 1 && q[0] == '=') ? ref : lookup_ref(A, q, l); }      // getMateReferenceName (:888-894)
  { const char* q = p + tab[6] + 1; int l = tab[7] - tab[6] - 1;
    if (l == 1 && q[0] == '*') pnext = -1; else if (!parse_int(q, l, pnext)) { raise(err, DERR_INT, li); return; } }
  int seqLen = tab[9] - tab[8] - 1;
  int qualEnd = nt >= 11 ? tab[10] : n, qualLen = qualEnd - tab[9] - 1;
  if (!(qualLen == 1 && p[tab[9] + 1] == '*') && qualLen != seqLen) { raise(err, DERR_SEQ_QUAL, li); return; }
  bool mapped = pos >= 0 && (flag & 4) == 0;                            // :980-982
  bool hasMate = (flag & 1) && (flag & 2);                              // :964-966
  S.frag[li] = frag; S.flag[li] = (uint16_t)flag; S.ref[li] = ref; S.pos[li] = pos; S.rnext[li] = rnext; S.pnext[li] = pnext;
  S.nm[li] = (int16_t)nm; S.cigOff[li] = (uint16_t)cigOff; S.cigLen[li] = (uint16_t)cigLen; S.seqLen[li] = (uint16_t)seqLen;
  S.rf[li] = (uint8_t)(RF_REC | (mapped ? RF_MAPPED : 0) | (hasMate ? RF_HASMATE : 0) | (readIndex << 3) | (suffix << 5));
}

EQP_HD void decode_record(const AnnoView& A, const SamView& S, uint32_t li, const char* p, int n, uint32_t* err) {
  ScanTabs tabs; decode_record_t(A, S, li, p, n, tabs, err);
}

EQP_HD void decode_line(const AnnoView& A, const SamView& S, uint32_t li, uint32_t* err) {
  uint64_t b = S.lineOff[li], e = S.lineOff[li + 1] - 1;              // [b,e): line without '\n'
  decode_record(A, S, li, S.text + b, (int)(e - b), err);
}

}  // namespace eqpb
