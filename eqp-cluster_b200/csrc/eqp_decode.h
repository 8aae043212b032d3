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

// first '\t' in p[from, n), or n.  Word-wise: aligned 4-byte loads, exact zero-byte test on (word ^ 0x09090909).
// Reads only aligned words that contain at least one byte of [from, n), so it never touches another page.
EQP_HD int find_tab(const char* p, int from, int n) {
  if (from >= n) return n;
  const char* q = p + from; int mis = (int)((uintptr_t)q & 3);
  const uint32_t* w = reinterpret_cast<const uint32_t*>(q - mis);
  int idx = from - mis;                                                 // index (relative to p) of byte 0 of *w
  uint32_t x = *w ^ 0x09090909u;
  x |= (mis ? (0xFFFFFFFFu >> (32 - 8 * mis)) : 0u);                     // bytes before `from` cannot match
  while (true) {
    uint32_t z = (x - 0x01010101u) & ~x & 0x80808080u;                   // lowest set marker is exact
    if (z) {
#if defined(__CUDA_ARCH__)
      int k = idx + ((__ffs((int)z) - 1) >> 3);
#else
      int k = idx + (__builtin_ctz(z) >> 3);
#endif
      return k < n ? k : n;
    }
    idx += 4; if (idx >= n) return n;
    x = *++w ^ 0x09090909u;
  }
}

// One SAM line [p, p+n) (no '\n').  The text may live in shared or global memory.
EQP_HD void decode_record(const AnnoView& A, const SamView& S, uint32_t li, const char* p, int n, uint32_t* err) {
  if (n > 0 && p[n - 1] == '\r') n--;
  S.frag[li] = 0xFFFFFFFFu; S.rf[li] = 0;
  if (n <= 0 || p[0] == '@') return;                                    // SamReader.java:173
  int tab[11]; int nt = 0;
  { int pos = 0; while (nt < 11) { int q = find_tab(p, pos, n); if (q >= n) break; tab[nt++] = q; pos = q + 1; } }
  // field count with the reference's rule: the 11 mandatory tokens count, later tokens only when they hold no ':' (SamRecord.java:223-232).
  // The same pass over the tag block finds the first [nN]M:i:<digits> and rejects an HI:i: tag.
  int nm = -1; int fields;
  if (nt < 10) fields = nt + (((nt ? tab[nt - 1] + 1 : 0) < n) ? 1 : 0);
  else if (nt == 10) fields = (tab[9] + 1 < n) ? 11 : 10;
  else {
    fields = 11; int k = tab[10] + 1; bool colon = false, any = k < n;
    for (; k < n; k++) {
      char c = p[k];
      if (c == '\t') { if (!colon) fields++; colon = false; any = k + 1 < n; continue; }
      if (c == ':') {
        colon = true;
        if (k + 2 < n && p[k + 1] == 'i' && p[k + 2] == ':' && p[k - 1] == 'M' && k - 2 > tab[10]) {
          char c0 = p[k - 2];
          if (nm < 0 && (c0 == 'N' || c0 == 'n')) {
            int d = k + 3, v = 0; bool dig = false;
            while (d < n && p[d] >= '0' && p[d] <= '9') { v = v * 10 + (p[d] - '0'); d++; dig = true; if (v > 30000) break; }
            if (dig) nm = v;
          }
        }
        if (k + 2 < n && p[k + 1] == 'i' && p[k + 2] == ':' && p[k - 1] == 'I' && k - 2 > tab[10] && p[k - 2] == 'H' && (k - 3 == tab[10] || p[k - 3] == '\t')) {
          raise(err, DERR_FIELDS, li); return;                           // HI:i: changes isMate (SamRecord.java:658-660): not produced by bowtie2
        }
      }
    }
    if (any && !colon) fields++;                                         // last token (a trailing empty token is not one)
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

EQP_HD void decode_line(const AnnoView& A, const SamView& S, uint32_t li, uint32_t* err) {
  uint64_t b = S.lineOff[li], e = S.lineOff[li + 1] - 1;              // [b,e): line without '\n'
  decode_record(A, S, li, S.text + b, (int)(e - b), err);
}

}  // namespace eqpb
