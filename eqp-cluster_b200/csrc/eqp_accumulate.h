// a15, a18-a25: alignment keys, in-fragment dedupe, multiplicity, final weight, conforming features, accumulation.
// Reference: WeightObjectAlignment.setGenomicAlignmentString (J/bed-source/WeightObjectAlignment.java:185-219),
// WeightObject.mergeWeightAlignments (:216-252) + combineExons (WeightObjectAlignment.java:312-374),
// WeightObject.getNumGenomicAlignments (:819-880), CombineReadWeightFiles default mode (:173-248),
// ComputeCounts main loop (:728-858), processWeightObject (:257-446), GenomicAlignment.getConformingCountObjects (:570-712),
// GenomicInterval.updateExonSet (GenomicInterval.java:479-523).
#pragma once
#include "eqp_align.h"

namespace eqpb {

#define EQP_FIXED_SHIFT_D 38
#ifndef EQP_FLAT_SCAN
#define EQP_FLAT_SCAN 64               // fragments with up to this many hits are regrouped (fragment_weight) and streamed by the counting runs; larger ones walk the duplicate chains
#endif
EQP_HD int eqp_popc(uint32_t m) {
#if defined(__CUDA_ARCH__)
  return __popc(m);
#else
  return __builtin_popcount(m);
#endif
}

// ---- canonical alignment keys (a18).  The reference compares STRINGS: WeightObjectAlignment.setGenomicAlignmentString (:185-219) over
// the per-half strings that GenomicAlignment.adjustGenomicAlignmentString (:873-887) rebuilds for every alignment: intervals as
// "chr/start/end/" joined by ":".  A tr- interval prints "" outside the gene run (GenomicInterval.java:711-713), and because the rebuild
// tests `alignmentString == ""` by identity, LEADING tr- intervals vanish together with their separator, later ones leave an empty token;
// a half of tr- intervals only is the literal "" and switches the alignment string to plain concatenation (:201-215).
// Here a half is its token list; keys are a hash of the same structure and equality is decided on the tokens themselves.
struct TokCur { uint32_t s0, s1; int nIv, r; bool gene, started; };
EQP_HD TokCur tok_begin(const Aln& a, int h, bool gene) {
  TokCur c; c.s0 = h == 1 ? a.hit0 : a.split; c.s1 = h == 1 ? a.split : a.hit1; c.nIv = h == 1 ? a.nIv1 : a.nIv2; c.r = -1; c.gene = gene; c.started = false; return c;
}
static const int32_t TOK_EMPTY = (int32_t)0x80000000;
EQP_HD bool tok_next(const HitX* hx, TokCur& c, int32_t& chr, int32_t& gs, int32_t& ge) {
  while (++c.r < c.nIv) {
    for (uint32_t i = c.s0; i < c.s1; i++) {
      const HitX& x = hx[i];
      if (!(x.flags & 1) || x.ivRank != c.r) continue;
      if (x.chr <= -2 && !c.gene) {
        if (!c.started) break;                                              // leading tr- interval: dropped
        chr = TOK_EMPTY; gs = ge = 0; return true;
      }
      c.started = true; chr = x.chr; gs = x.gs; ge = x.ge; return true;
    }
  }
  return false;
}
EQP_HD bool half_lit(const Aln& a, int h, bool gene) { return !gene && ((a.tr >> (3 * (h - 1))) & 2); }   // string is the literal ""
// 0 = "a-b" ordered by start (both halves, neither literal), 1 = "l-", 2 = "r-", 3 = concatenation
EQP_HD int key_form(const Aln& a, bool gene) {
  bool h1 = a.flags & 2, h2 = a.flags & 4;
  if (h1 && h2 && !half_lit(a, 1, gene) && !half_lit(a, 2, gene)) return 0;
  if (!h2) return 1;
  if (!h1) return 2;
  return 3;
}
EQP_HD uint64_t alignment_key_inl(const HitX* hx, const Aln& al, bool gene) {
  const uint8_t fl = al.flags; const int form = key_form(al, gene);
  uint64_t hsh = mix64(0x1234, (uint64_t)((fl & 1) ? 0 : (fl >> 3) & 3));    // alignmentBaseId: fragment id, or fragment id + /S1, /S2
  hsh = mix64(hsh, (uint64_t)form);
  const bool secondFirst = form == 0 && !(al.start1 <= al.start2);
  for (int part = 0; part < 2; part++) {
    int half = form == 0 ? (secondFirst ? 2 - part : 1 + part) : form == 1 ? 1 : form == 2 ? 2 : 1 + part;
    if ((form == 1 || form == 2) && part == 1) break;
    if (form == 0 && part == 1) hsh = mix64(hsh, 0xABCDEF00u);               // the "-" between the two halves
    TokCur c = tok_begin(al, half, gene); int32_t chr, gs, ge;
    while (tok_next(hx, c, chr, gs, ge)) hsh = mix64(mix64(mix64(hsh, (uint64_t)(uint32_t)chr), (uint64_t)gs), (uint64_t)ge);
  }
  return hsh;
}
// Real calls (the hit table is passed, not the view: a by-reference struct argument would force the caller to copy its kernel parameters to
// local memory).  Only the rare paths (a17, gene-run dedupe) use the call; inside alignment_build the call cost 19 ms per 25M-read chunk
// (measured, profiles/r03c), so the hot path inlines the key.
EQP_HD_CALL uint64_t alignment_key(const HitX* hx, const Aln& al, bool gene) { return alignment_key_inl(hx, al, gene); }
EQP_HD_CALL bool halves_equal(const HitX* hx, const Aln& a, int ha, const Aln& b, int hb, bool gene) {
  TokCur ca = tok_begin(a, ha, gene), cb = tok_begin(b, hb, gene);
  while (true) {
    int32_t c1, s1, e1, c2, s2, e2;
    bool na = tok_next(hx, ca, c1, s1, e1), nb = tok_next(hx, cb, c2, s2, e2);
    if (na != nb) return false;
    if (!na) return true;
    if (c1 != c2 || s1 != s2 || e1 != e2) return false;
  }
}
// string equality of two WeightObjectAlignment keys; on success `swapped` tells how b's halves map onto a's
EQP_HD_CALL bool keys_equal(const HitX* hx, const Aln& a, uint64_t ka, const Aln& b, uint64_t kb, bool gene, bool& swapped) {
  swapped = false;
  if (ka != kb) return false;
  if ((a.flags & 1) != (b.flags & 1)) return false;
  if (!(a.flags & 1) && ((a.flags >> 3) & 3) != ((b.flags >> 3) & 3)) return false;
  const int form = key_form(a, gene);
  if (form != key_form(b, gene)) return false;
  if (form == 0) {
    bool sa = !(a.start1 <= a.start2), sb = !(b.start1 <= b.start2);     // which half is printed first
    int a1 = sa ? 2 : 1, a2 = sa ? 1 : 2, b1 = sb ? 2 : 1, b2 = sb ? 1 : 2;
    if (!halves_equal(hx, a, a1, b, b1, gene) || !halves_equal(hx, a, a2, b, b2, gene)) return false;
    // combineExons (:312-337): straight pairing if ga1==ga1' and ga2==ga2', else the mate-swapped pairing
    swapped = !(halves_equal(hx, a, 1, b, 1, gene) && halves_equal(hx, a, 2, b, 2, gene));
    return true;
  }
  if (form == 1) return halves_equal(hx, a, 1, b, 1, gene);
  if (form == 2) return halves_equal(hx, a, 2, b, 2, gene);
  // concatenation: at most one half of each alignment has tokens
  bool straight = halves_equal(hx, a, 1, b, 1, gene) && halves_equal(hx, a, 2, b, 2, gene);
  if (straight) return true;
  if (halves_equal(hx, a, 1, b, 2, gene) && halves_equal(hx, a, 2, b, 1, gene)) { swapped = true; return true; }
  return false;
}

// ---- stage A (one alignment per item): sort the alignment's hits, build both halves, hash the canonical key
EQP_HD void alignment_build(const AnnoView& A, const ChunkView& C, const CountView& V, const RunParams& P, uint32_t a) {
  uint32_t b0 = C.alnBlk[a], b1 = C.alnBlk[a + 1];
  uint32_t k = C.hitCnt[b0], e = C.hitCnt[b1];                           // hits are grouped by block, blocks by alignment
  Aln al{}; al.repr = a; al.next = 0xFFFFFFFFu;                          // assembled in registers, stored once (48 bytes = three 16-byte stores)
  if (k == e) { V.alns[a] = al; return; }
  uint32_t frag = C.base + C.blocks[b0].slot;
  for (uint32_t i = k + 1; i < e; i++) {                                 // insertion sort by (readIdx, readStart, eRefStart, eRefEnd, strand, row); join_block emits every block's hits in that order
    Hit cur = C.hits[i]; RecKey ck = rec_key(A, C, cur); uint32_t j = i;
    while (j > k) { RecKey pk = rec_key(A, C, C.hits[j - 1]); int c = rec_cmp(pk, ck); if (c < 0 || (c == 0 && pk.row <= ck.row)) break; C.hits[j] = C.hits[j - 1]; j--; }
    if (j != i) C.hits[j] = cur;
  }
  uint32_t split = k; while (split < e && C.blocks[C.hits[split].block].readIdx == 1) split++;
  al.hit0 = k; al.split = split; al.hit1 = e;
  const Block fb = C.blocks[C.hits[k].block];
  uint8_t fl = fb.paired ? 1 : 0;
  for (uint32_t i = k; i < e; i++) if ((C.blocks[C.hits[i].block].paired ? 1 : 0) != (fl & 1)) raise(C.err, DERR_INCONSISTENT, frag);
  al.start1 = al.start2 = 0x7fffffff; al.nIv1 = al.nIv2 = 0; al.maxField1 = al.maxField2 = -1;
  uint8_t tr1 = 0, tr2 = 0;
  if (split > k) { build_half(A, C, V, P, k, split, 1, al.start1, al.nIv1, al.maxField1, tr1, frag); fl |= 2; }
  if (e > split) { build_half(A, C, V, P, split, e, 2, al.start2, al.nIv2, al.maxField2, tr2, frag); fl |= 4; }
  if (!(fl & 1)) fl |= (uint8_t)(((fl & 2) ? 1 : 2) << 3);
  if (!(fl & 1) && (fl & 2) && (fl & 4)) raise(C.err, DERR_INCONSISTENT, frag);     // single-read alignment with both read indices
  al.flags = fl; al.tr = (uint8_t)(tr1 | (tr2 << 3));
  al.key = alignment_key_inl(V.hx, al, false);
  V.alns[a] = al;
}

EQP_HD uint64_t hit_group_key(const HitX& x) { return ((uint64_t)((x.flags & 1) ? 0 : 1) << 40) | ((uint64_t)x.grp << 24) | ((uint64_t)x.half << 16) | x.ivRank; }

// ---- stage B: dedupe, a20, final weight
// a19 + a20 for one fragment: dedupe its alignments by key (first in file order is the representative, later ones are chained to it)
// and count the representatives.  GENE = false: keys of the weight pass and of the exon / junction runs, state in Aln; GENE = true: keys
// of the gene run, state in AlnG (only fragments with tr- exon hits need it: without a tr- interval the two keys are the same string).
// Cost (bisect r03l, 25M reads on the round-1 world, kernel fragment_weight): 22.98 ms as built; 6.71 ms with this loop skipped; 21.03 ms with
// round-1 style rank equality instead of the token lists; the hash pre-check in front of keys_equal: 24.1 vs 24.2 ms on the shaped world.  The
// loop is bound by the scattered 48-byte Aln reads of the fragment's alignments, not by the comparisons.
template <bool GENE>
EQP_HD uint32_t dedupe_fragment(const CountView& V, uint32_t a0, uint32_t a1) {
  int left = 0, right = 0; bool hasTr = false; uint32_t nA = a1 - a0, cap = 2 * nA; uint32_t* tab = V.dedupTab + 2 * (uint64_t)a0;
  if (GENE) { for (uint32_t a = a0; a < a1; a++) { AlnG& g = V.alnG[a]; g.repr = a; g.next = 0xFFFFFFFFu; g.flags = 0; g.key = V.alns[a].hit0 == V.alns[a].hit1 ? 0 : alignment_key(V.hx, V.alns[a], true); }
    if (nA > 4) for (uint32_t s = 0; s < cap; s++) tab[s] = 0; }            // the table was used by the first dedupe
  for (uint32_t a = a0; a < a1; a++) {
    Aln& al = V.alns[a]; if (al.hit0 == al.hit1) continue;
    const uint64_t key = GENE ? V.alnG[a].key : al.key;
    if (!GENE) al.repr = a;
    uint32_t found = 0xFFFFFFFFu; bool sw = false;
    if (nA <= 4) {                                                         // tiny fragment: compare with the earlier representatives directly
      for (uint32_t p = a0; p < a; p++) {
        const Aln& q = V.alns[p]; if (q.hit0 == q.hit1 || (GENE ? V.alnG[p].repr : q.repr) != p) continue;
        const uint64_t qk = GENE ? V.alnG[p].key : q.key;
        if (qk != key) continue;                                             // the hash decides almost every comparison without the call
        if (keys_equal(V.hx, q, qk, al, key, GENE, sw)) { found = p; break; }
      }
    } else {                                                               // per-fragment open-addressing table of representatives (2 slots per alignment, zeroed)
      uint32_t s = (uint32_t)(((key * 0x9E3779B97F4A7C15ULL) >> 32) % cap);
      while (true) {
        uint32_t e = tab[s];
        if (!e) { tab[s] = a - a0 + 1; break; }
        const uint32_t p = a0 + e - 1;
        const uint64_t pk = GENE ? V.alnG[p].key : V.alns[p].key;
        if (pk == key && keys_equal(V.hx, V.alns[p], pk, al, key, GENE, sw)) { found = p; break; }   // insertion order along the probe path = file order: first equal key wins
        s = s + 1 == cap ? 0 : s + 1;
      }
    }
    if (found != 0xFFFFFFFFu) {                                            // duplicate: append to the representative's chain (file order)
      if (GENE) { V.alnG[a].repr = found; if (sw) V.alnG[a].flags |= 32; V.alnG[V.alnTail[found]].next = a; }
      else { al.repr = found; if (sw) al.flags |= 32; V.alns[V.alnTail[found]].next = a; }
      V.alnTail[found] = a;                                              // (the tail as a 16-bit field inside Aln measured 48.6 vs 24.2 ms for this kernel, profiles/r03h)
    } else {
      V.alnTail[a] = a;
      // getNumGenomicAlignments (WeightObject.java:819-880): alignments of tr- intervals only are not counted but add one in total
      const bool trAln = (!(al.flags & 2) || (al.tr & 2)) && (!(al.flags & 4) || (al.tr & 16));
      if (trAln) hasTr = true; else if (key_form(al, GENE) == 2) right++; else left++;   // ":r-" forms (:859-865)
    }
  }
  return (uint32_t)(left > right ? left : right) + (hasTr ? 1u : 0u);
}

EQP_HD void fragment_weight(const AnnoView& A, const ChunkView& C, const CountView& V, const RunParams& P, uint32_t slot) {
  (void)A; (void)P;
  const SlotPack sp = V.pack[slot];
  if (sp.hit0 == sp.hit1) return;                                         // a20 / nFinal stay 0
  uint32_t a0 = sp.aln0, a1 = sp.aln1; const uint32_t nA = a1 - a0;
  if (nA > 0xFFFFu) { raise(C.err, DERR_CAPACITY, C.base + slot); return; }
  const uint32_t a20 = dedupe_fragment<false>(V, a0, a1);
  // tag every hit with its representative (index within the fragment) and with the representative's half it merges into
  // (combineExons pairs mate-swapped duplicates crosswise, WeightObjectAlignment.java:312-337): a22-a25 then read the
  // fragment's hits as one flat range instead of walking duplicate chains.  Fragments with tr- exon hits are left untagged: their
  // gene run has its own dedupe (fragment_weight_gene) and HitX::grp keeps the non-tr ordinal of the interval.
  const bool trFrag = (sp.g & SP_TR) != 0;
  if (!trFrag) for (uint32_t a = a0; a < a1; a++) {
    const Aln& al = V.alns[a]; if (al.hit0 == al.hit1) continue;
    bool sw = al.repr != a && (al.flags & 32); uint16_t g = (uint16_t)(al.repr - a0);
    for (uint32_t i = al.hit0; i < al.hit1; i++) { HitX& x = V.hx[i]; x.grp = g; if (sw) x.half = (uint8_t)(3 - x.half); }
  }
  // small fragments: regroup the hits by (representative, half, interval), stable, dropped duplicates last: the counting runs
  // then stream each fragment's hits once.  (hx is only indexed through Aln.hit0.. before this point; large fragments keep
  // their order and walk the duplicate chains.)
  if (!trFrag) { uint32_t h0 = sp.hit0, h1 = sp.hit1;
    bool sorted = true; uint64_t prev = 0;
    if (h1 - h0 <= EQP_FLAT_SCAN) for (uint32_t i = h0; i < h1; i++) { uint64_t k = hit_group_key(V.hx[i]); sorted &= k >= prev; prev = k; }   // independent loads
    if (h1 - h0 <= EQP_FLAT_SCAN && !sorted)
      for (uint32_t i = h0 + 1; i < h1; i++) {
        HitX cur = V.hx[i]; uint64_t ck = hit_group_key(cur); uint32_t j = i;
        while (j > h0 && hit_group_key(V.hx[j - 1]) > ck) { V.hx[j] = V.hx[j - 1]; j--; }
        if (j != i) V.hx[j] = cur;
      } }
  // final.wgt entry: BED weight (single-read, read index 1, n = a20) against the genome SAM weight with the T+J entry as class arbiter
  // (CombineReadWeightFiles.java:192-214); how the two combine does not depend on a20 and was worked out in slot order (slot_weight_mode)
  const uint32_t gN = sp.g & 0xFFFFFFu, how = (sp.g >> 24) & 3u;
  const uint32_t n = how == SP_A20 ? a20 : how == SP_GENOME ? gN : (a20 >= gN ? a20 : gN);
  V.pack[slot].a20 = a20; V.pack[slot].nFinal = n;
}
// the part of the final weight that does not depend on a20: numAlignments of the genome entry and how it combines (SlotPack::g)
EQP_HD uint32_t slot_weight_mode(const ChunkView& C, const RunParams& P, uint32_t slot) {
  if (P.genomic) { uint32_t n = (uint32_t)fe_n(C.st[FILE_T][slot]); return (n > 0xFFFFFFu ? 0xFFFFFFu : n) | ((uint32_t)SP_GENOME << 24); }   // ComputeCounts -w <ComputeReadWeightsSam output> (compute-genomic-counts.sh:1000-1010)
  const FragStat g = C.st[FILE_G][slot];
  if (!(g.flags & FS_EXISTS)) return (uint32_t)SP_A20 << 24;
  uint32_t ax = C.auxSlot[slot]; int cm = 0;
  if (ax != 0xFFFFFFFFu) cm = fe_cmp_mult(C.stTJ[ax], g);                   // aux.compareMultiplicity(entry2), CombineReadWeightFiles.java:192-195
  uint32_t n = (uint32_t)fe_n(g); if (n > 0xFFFFFFu) n = 0xFFFFFFu;
  return n | ((uint32_t)(cm == 1 ? SP_A20 : cm == -1 ? SP_GENOME : SP_MAX) << 24);
}

// second dedupe of a fragment with tr- exon hits, under the gene run's keys (a20 can differ between the runs)
EQP_HD void fragment_weight_gene(const ChunkView& C, const CountView& V, uint32_t slot) {
  (void)C;
  const SlotPack sp = V.pack[slot];
  V.pack[slot].a20G = sp.a20 == 0 ? 0u : dedupe_fragment<true>(V, sp.aln0, sp.aln1);
}

// ---- a17 for a fragment with a positive lengthDifference: WeightObject.adjustGenomicAlignments (:180-204) ->
// WeightObjectAlignment.adjustGenomicAlignmentString (:266-300) -> GenomicInterval.adjustChrStartEnd (GenomicInterval.java:284-421).
// A sequential sweep in file order, exactly as the reference runs it: every interval with a length difference is compared with every
// interval of the same half of every other alignment and may be snapped onto it (or snap it).  An interval lives in the records that
// carry its rank; its first record also owns the a17 state (IvAdj).  Keys are rebuilt afterwards.
struct IvRef { uint32_t first; uint32_t s0, s1; uint16_t rank; int32_t refId; };
EQP_HD bool iv_find(const CountView& V, uint32_t s0, uint32_t s1, int rank, uint32_t& first) {
  for (uint32_t i = s0; i < s1; i++) { const HitX& x = V.hx[i]; if ((x.flags & 1) && x.ivRank == rank) { first = i; return true; } }
  return false;
}
EQP_HD void iv_move(const CountView& V, const IvRef& x, int32_t gs, int32_t ge, int32_t ld) {
  for (uint32_t i = x.s0; i < x.s1; i++) { HitX& h = V.hx[i]; if ((h.flags & 1) && h.ivRank == x.rank) { h.gs = gs; h.ge = ge; } }
  V.ivAdj[x.first].ld = ld;
}
// GenomicInterval.adjustChrStartEnd(GenomicInterval) :284-358
EQP_HD bool adjust_pair(const CountView& V, const IvRef& x0, const IvRef& y0) {
  const IvRef* px = &x0; const IvRef* py = &y0;
  { const HitX& hx = V.hx[px->first]; const HitX& hy = V.hx[py->first];
    if (hx.chr != hy.chr) return false;
    if (px->refId == py->refId) return false;
    if (hx.gs == hy.gs && hx.ge == hy.ge) return false;
    const int32_t xld = V.ivAdj[px->first].ld, yld = V.ivAdj[py->first].ld;
    if (xld < yld || (xld == yld && hx.gs < hy.gs)) { const IvRef* t = px; px = py; py = t; } }   // the recursive call with swapped roles (:303-316)
  const IvRef& x = *px; const IvRef& y = *py;
  const HitX& hx = V.hx[x.first]; const HitX& hy = V.hx[y.first];
  const int32_t xld = V.ivAdj[x.first].ld, yld = V.ivAdj[y.first].ld;
  int32_t lx = hx.ge - hx.gs, ly = hy.ge - hy.gs; if (lx < 0) lx = -lx; if (ly < 0) ly = -ly;
  int32_t d = V.ivAdj[y.first].orig - V.ivAdj[x.first].orig; if (d < 0) d = -d;
  if (lx == ly && d <= xld && (xld > yld || (xld == yld && hx.gs > hy.gs))) { iv_move(V, x, hy.gs, hy.ge, yld); return true; }
  return false;
}
EQP_HD void fragment_adjust(const AnnoView& A, const ChunkView& C, const CountView& V, uint32_t slot) {
  uint32_t a0 = C.alnOff[slot], a1 = C.alnOff[slot + 1];
  for (uint32_t a = a0; a < a1; a++) {                                     // a17 state of every interval
    const Aln& al = V.alns[a];
    for (uint32_t i = al.hit0; i < al.hit1; i++) { const HitX& x = V.hx[i]; if (x.flags & 1) { V.ivAdj[i].ld = record_ld(A, C, C.hits[i]); V.ivAdj[i].orig = x.gs; } }
  }
  for (uint32_t a = a0; a < a1; a++) {
    const Aln& al = V.alns[a]; if (al.hit0 == al.hit1) continue;
    const int32_t refA = A.bktRefId[C.blocks[C.hits[al.hit0].block].bucket];
    for (int h = 1; h <= 2; h++) {
      if (!(al.flags & (h == 1 ? 2 : 4))) continue;
      IvRef x; x.s0 = h == 1 ? al.hit0 : al.split; x.s1 = h == 1 ? al.split : al.hit1; x.refId = refA;
      const int nIv = h == 1 ? al.nIv1 : al.nIv2;
      for (int r = 0; r < nIv; r++) {
        if (!iv_find(V, x.s0, x.s1, r, x.first)) continue;
        x.rank = (uint16_t)r;
        if (V.ivAdj[x.first].ld == 0) continue;
        for (uint32_t a2 = a0; a2 < a1; a2++) {
          // The reference skips alignments whose string equals the (partly rebuilt) string of this one; for those adjust_pair cannot
          // act anyway: it needs a different reference id and different coordinates of equal length within 3 bp, which two intervals
          // of one read alignment never have.  The alignment itself is excluded by the reference-id test.
          const Aln& bl = V.alns[a2]; if (bl.hit0 == bl.hit1 || !(bl.flags & (h == 1 ? 2 : 4))) continue;
          IvRef y; y.s0 = h == 1 ? bl.hit0 : bl.split; y.s1 = h == 1 ? bl.split : bl.hit1; y.refId = A.bktRefId[C.blocks[C.hits[bl.hit0].block].bucket];
          const int nIv2 = h == 1 ? bl.nIv1 : bl.nIv2;
          for (int r2 = 0; r2 < nIv2; r2++) { if (!iv_find(V, y.s0, y.s1, r2, y.first)) continue; y.rank = (uint16_t)r2; adjust_pair(V, x, y); }
        }
      }
    }
  }
  for (uint32_t a = a0; a < a1; a++) { Aln& al = V.alns[a]; if (al.hit0 != al.hit1) al.key = alignment_key(V.hx, al, false); }
}

// ---- stage C: conforming features + accumulation for one run (mode)
// Per-fragment tables live in local memory; their size decides how much of the L1/L2 a resident thread claims.  Almost every fragment
// needs a handful of entries, so the counting kernel is compiled twice: SMALL tables for all fragments, and the few that overflow them
// are handed (through a device work list) to a second launch with the LARGE tables, whose overflow is the DERR_CAPACITY error.
#define EQP_FS_CAP 160
#define EQP_FT_CAP 320
#ifndef EQP_FS_SMALL
#define EQP_FS_SMALL 24
#define EQP_FT_SMALL 32
#endif
template <int CAP> struct FeatSetT { static const int cap = CAP; uint32_t id[CAP]; uint32_t mask[CAP]; int n; };
template <class FS> EQP_HD bool fs_add(FS& s, uint32_t id, uint32_t bit) {
  for (int i = 0; i < s.n; i++) if (s.id[i] == id) { s.mask[i] |= bit; return true; }
  if (s.n >= FS::cap) return false;
  s.id[s.n] = id; s.mask[s.n] = bit; s.n++; return true;
}

// One genomic interval of a half: its exon set (first-seen order) -> updateExonSet -> features with the interval's field bit.
struct ExonSet { int32_t exon[32]; uint8_t gen[32]; int n; bool over; };
EQP_HD void exon_add(ExonSet& e, const HitX& x) {
  int ex = x.exon; bool have = false;
  for (int q = 0; q < e.n; q++) have |= e.exon[q] == ex;
  if (!have) { if (e.n >= 32) e.over = true; else { e.exon[e.n] = ex; e.gen[e.n] = (x.flags & 2) ? 1 : 0; e.n++; } }
}
// GenomicInterval.updateExonSet (GenomicInterval.java:479-523): among several exons keep the genomic ones and the non-genomic ones of maximal overlap
EQP_HD int exon_filter(const AnnoView& A, ExonSet& e, int32_t gs, int32_t ge) {
  int ne = e.n; int32_t* exon = e.exon; uint8_t* exGen = e.gen;
  if (ne > 1) {
    int maxOv = 0;
    for (int q = 0; q < ne; q++) if (!exGen[q]) { int en = A.ex[exon[q]].end < ge ? A.ex[exon[q]].end : ge; int ov = en - gs + 1; if (ov < 0) ov = 0; if (ov > maxOv) maxOv = ov; }
    if (maxOv > 0) { int w = 0; for (int q = 0; q < ne; q++) { int en = A.ex[exon[q]].end < ge ? A.ex[exon[q]].end : ge; int ov = en - gs + 1; if (ov < 0) ov = 0; if (exGen[q] || ov >= maxOv) { exon[w] = exon[q]; exGen[w] = exGen[q]; w++; } } ne = w; }
  }
  e.n = ne; return ne;
}
// features of exon `ex` in run `mode`: cnt entries; inl: the single entry is `first` itself, else they are feat[mode][first .. first + cnt)
EQP_HD void exon_features(const AnnoView& A, int mode, int ex, uint32_t& first, uint32_t& cnt, bool& inl) {
  const ExonFeat f = A.exf[ex];
  if (mode == MODE_JUNCTION) { first = f.jOff; cnt = f.n >> 16; inl = false; }
  else { cnt = mode == MODE_GENE ? (f.n & 0xFFu) : ((f.n >> 8) & 0xFFu); first = mode == MODE_GENE ? f.g : f.e; inl = cnt == 1; }
}
// Junction run, a half with exactly two intervals of one exon each (the ordinary spliced read): a junction conforms iff it is a
// candidate of both exons and the two fields are consecutive (:679-701), i.e. the intersection of the two exons' feature lists.
// The lists are sorted by id when the annotation is loaded, so this is one merge instead of a set with field masks.
// Returns false when the shape is different (the general path then handles the half).
template <class FS>
EQP_HD bool junction_pair(const AnnoView& A, const CountView& V, ExonSet& ea, int32_t gsa, int32_t gea, int fa, ExonSet& eb, int32_t gsb, int32_t geb, int fb, FS& out, bool& overflow) {
  if (ea.over || eb.over) return false;
  if (exon_filter(A, ea, gsa, gea) != 1 || exon_filter(A, eb, gsb, geb) != 1) return false;   // the filter is idempotent: the general path may run it again
  out.n = 0;
  if (fa - fb != 1 && fb - fa != 1) return true;
  const uint32_t* F = A.feat[MODE_JUNCTION];
  uint32_t i, i1, j, j1; bool inl;
  exon_features(A, MODE_JUNCTION, ea.exon[0], i, i1, inl); i1 += i; exon_features(A, MODE_JUNCTION, eb.exon[0], j, j1, inl); j1 += j;
  while (i < i1 && j < j1) {
    uint32_t va = F[i], vb = F[j], ia = va & 0x7fffffffu, ib = vb & 0x7fffffffu;
    if (ia < ib) i++; else if (ib < ia) j++;
    else {
      // a row survives the chunk's map filter (a27) or is in the transcript map
      bool pa = (va >> 31) || V.junctionActive[A.featAct ? A.featAct[i] : ia] != 0, pb = (vb >> 31) || V.junctionActive[A.featAct ? A.featAct[j] : ib] != 0;
      if (pa && pb) { if (out.n >= FS::cap) { overflow = true; return true; } out.id[out.n] = ia; out.mask[out.n] = 0; out.n++; }
      i++; j++;
    }
  }
  return true;
}
// returns false when a table overflowed
template <class FS>
EQP_HD bool interval_features(const AnnoView& A, const ChunkView& C, const CountView& V, const RunParams& P, int mode, ExonSet& e, int32_t gs, int32_t ge, int field,
                              FS& fs, uint32_t frag) {
  (void)C; (void)frag;
  if (e.over) return false;
  int ne = exon_filter(A, e, gs, ge); int32_t* exon = e.exon;
  if (mode == MODE_EXON && P.unambiguous[mode] && ne >= 2) ne = 0;      // -U with genomicIntervalMode (:603-608)
  const uint32_t bit = field >= 0 ? 1u << field : 0u;                   // a tr- record of the junction run has no field (:638-656)
  for (int q = 0; q < ne; q++) {
    int ex = exon[q]; uint32_t f0, cnt; bool inl, any = false;
    exon_features(A, mode, ex, f0, cnt, inl);
    for (uint32_t f = f0; f < f0 + cnt; f++) {
      uint32_t v = inl ? f0 : A.feat[mode][f]; uint32_t id = v & 0x7fffffffu;
      if (mode == MODE_JUNCTION && !(v & 0x80000000u) && !V.junctionActive[A.featAct ? A.featAct[f] : id]) continue;   // row removed by the per-chunk map filter (a27)
      any = true;
      if (!fs_add(fs, id, bit)) return false;
    }
    if (!any && !fs_add(fs, (uint32_t)A.nFeat[mode] + (uint32_t)ex, bit)) return false;   // unknown exon: own pseudo id
  }
  return true;
}
// field-coverage test of a half's candidate features (GenomicAlignment.java:570-712)
template <class FS>
EQP_HD void conforming_features(const FS& fs, int mode, int maxField, FS& out) {
  out.n = 0;
  for (int i = 0; i < fs.n; i++) {
    uint32_t m = fs.mask[i]; bool okf;
    if (mode == MODE_GENE) okf = eqp_popc(m) >= maxField + 1;
    else if (mode == MODE_EXON) okf = m != 0;
    else okf = (m & (m >> 1)) != 0;                                       // two consecutive field indices (:679-701)
    if (okf) { out.id[out.n] = fs.id[i]; out.mask[out.n] = 0; out.n++; }
  }
}

// features of half `h` of representative alignment R (index r) with all merged duplicates folded in: duplicate-chain walk
// (fragments with more than EQP_FLAT_SCAN hits, and fragments with tr- exon hits; the others are regrouped by fragment_weight and
// streamed, see fragment_count).  trFrag: the hits are untagged (HitX::half is the read index, HitX::grp the non-tr ordinal), the
// gene run walks the AlnG chains, and outside the gene run duplicates pair their non-tr intervals by ordinal while tr- intervals keep
// the representative's own exons only (GenomicAlignment.combineExons skips intervals that print "", :721-786).
// returns false when a table overflowed
template <class FS>
EQP_HD bool half_features(const AnnoView& A, const ChunkView& C, const CountView& V, const RunParams& P, int mode, uint32_t r, int h, bool trFrag, FS& out, uint32_t frag) {
  const Aln& R = V.alns[r]; out.n = 0;
  int nIv = h == 1 ? R.nIv1 : R.nIv2; int maxField = h == 1 ? R.maxField1 : R.maxField2;
  if (maxField >= 32) return false;
  if (mode == MODE_JUNCTION && nIv < 2) return true;                       // a junction needs two consecutive fields (:679-701)
  const bool useG = trFrag && mode == MODE_GENE;
  const uint32_t r0 = h == 1 ? R.hit0 : R.split, r1 = h == 1 ? R.split : R.hit1;
  FS fs; fs.n = 0;
  for (int iv = 0; iv < nIv; iv++) {
    int32_t gs = 0, ge = 0; int field = 0; bool haveIv = false, trIv = false; uint16_t ord = 0; ExonSet e; e.n = 0; e.over = false;
    if (trFrag) { uint32_t f; if (iv_find(V, r0, r1, iv, f)) { trIv = V.hx[f].chr <= -2; ord = V.hx[f].grp; } }
    for (uint32_t d = r; d != 0xFFFFFFFFu; d = useG ? V.alnG[d].next : V.alns[d].next) {          // R first, then later duplicates in file order
      const Aln& D = V.alns[d];
      const bool sw = trFrag && d != r && ((useG ? V.alnG[d].flags : D.flags) & 32);
      if (trFrag && d != r && trIv && !useG) break;
      for (uint32_t i = D.hit0; i < D.hit1; i++) {
        const HitX& x = V.hx[i];
        if (!(x.flags & 1)) continue;
        if ((sw ? 3 - x.half : x.half) != h) continue;
        if (trFrag && d != r && !useG) { if (x.grp != ord) continue; }
        else if (x.ivRank != iv) continue;
        if (!haveIv) { gs = x.gs; ge = x.ge; field = mode == MODE_JUNCTION ? x.ivFieldJ : x.ivField; haveIv = true; }   // R's own first record of the interval
        if (x.flags & (4 << mode)) exon_add(e, x);
      }
    }
    if (!interval_features(A, C, V, P, mode, e, gs, ge, field, fs, frag)) return false;
  }
  conforming_features(fs, mode, maxField, out);
  return true;
}

EQP_HD void atomic_add_u64(unsigned long long* p, unsigned long long v) {
#if defined(__CUDA_ARCH__)
  atomicAdd(p, v);
#else
  *p += v;
#endif
}

// returns 0 = the fragment is no weight object, 1 = it added to a count, 2 = it was excluded (or aborted on a capacity error)
// FS_CAP / FT_CAP: table sizes (see EQP_FS_SMALL); with small tables an overflow returns -1 and the caller queues the fragment for the large ones.
// wsum: the fragment's weightAdded (sum of its contributions, printed ids or not, ComputeCounts.java:772,849) in 2^-32 units
template <int FS_CAP, int FT_CAP>
EQP_HD int fragment_count(const AnnoView& A, const ChunkView& C, const CountView& V, const RunParams& P, int mode, uint32_t slot, uint64_t& wsum) {
  typedef FeatSetT<FS_CAP> FeatSet;
  const bool retry = FS_CAP < EQP_FS_CAP;                                  // an overflow is not final
  wsum = 0;
  const SlotPack sp = V.pack[slot];
  const bool trFrag = (sp.g & SP_TR) != 0;
  uint32_t a20 = trFrag && mode == MODE_GENE ? sp.a20G : sp.a20;
  if (sp.a20 == 0) return 0;
  uint32_t frag = C.base + slot; uint32_t a0 = sp.aln0, a1 = sp.aln1;
  const uint32_t hEnd = sp.hit1;                                          // end of the fragment's hit range
  double w = 1.0 / (double)sp.nFinal;
  if (w > 1.0 / (double)a20) w = 1.0 / (double)a20;                      // ComputeCounts.java:762-764
  if (!(w >= P.weightThreshold)) return 2;
  double addW = P.unweighted ? 1.0 : w;
  uint32_t fid[FT_CAP]; uint16_t k1[FT_CAP], k2[FT_CAP]; int nf = 0;
  FeatSet s1, s2; bool bad = false;
  // one representative alignment done: paired -> gene run intersects the halves, the others unite them (:325-335); then count per feature
  auto merge_alignment = [&](uint8_t flags) {
    if (flags & 1) {
      if (mode == MODE_GENE) { int wn = 0; for (int i = 0; i < s1.n; i++) { bool in2 = false; for (int j = 0; j < s2.n; j++) in2 |= s2.id[j] == s1.id[i]; if (in2) s1.id[wn++] = s1.id[i]; } s1.n = wn; }
      else for (int j = 0; j < s2.n; j++) { bool in1 = false; for (int i = 0; i < s1.n; i++) in1 |= s1.id[i] == s2.id[j]; if (!in1) { if (s1.n >= FS_CAP) { bad = true; return; } s1.id[s1.n++] = s2.id[j]; } }
      s2.n = 0;
    }
    for (int side = 0; side < 2; side++) {
      FeatSet& s = side ? s2 : s1;
      for (int i = 0; i < s.n; i++) {
        int q = 0; while (q < nf && fid[q] != s.id[i]) q++;
        if (q == nf) { if (nf >= FT_CAP) { bad = true; return; } fid[nf] = s.id[i]; k1[nf] = k2[nf] = 0; nf++; }
        if (side) k2[q]++; else k1[q]++;
      }
    }
  };
  const uint32_t hBeg = sp.hit0;
  if (hEnd - hBeg <= EQP_FLAT_SCAN && !trFrag) {
    // hits regrouped by fragment_weight: (representative, half, interval) runs, representatives in file order
    FeatSet fs; uint32_t i = hBeg;
    while (i < hEnd && (V.hx[i].flags & 1) && !bad) {
      const uint16_t g = V.hx[i].grp; const Aln& R = V.alns[a0 + g];
      s1.n = s2.n = 0;
      while (i < hEnd && (V.hx[i].flags & 1) && V.hx[i].grp == g && !bad) {
        const int h = V.hx[i].half; int maxField = h == 1 ? R.maxField1 : R.maxField2;
        if (maxField >= 32) { bad = true; break; }
        fs.n = 0;
        if (mode == MODE_JUNCTION && (h == 1 ? R.nIv1 : R.nIv2) < 2) {      // a junction needs two consecutive fields (:679-701): one interval cannot conform
          while (i < hEnd && (V.hx[i].flags & 1) && V.hx[i].grp == g && V.hx[i].half == h) i++;
          (h == 1 ? s1 : s2).n = 0; continue;
        }
        const bool pairShape = mode == MODE_JUNCTION && (h == 1 ? R.nIv1 : R.nIv2) == 2;   // candidate for junction_pair()
        ExonSet e2[2]; int32_t gs2[2], ge2[2]; int field2[2]; int nColl = 0;
        while (i < hEnd && !bad) {                                         // one interval per iteration
          HitX x = V.hx[i];
          if (!(x.flags & 1) || x.grp != g || x.half != h) break;
          const int iv = x.ivRank; ExonSet& e = e2[pairShape ? (nColl & 1) : 0]; e.n = 0; e.over = false;
          const int32_t gs = x.gs, ge = x.ge; const int field = mode == MODE_JUNCTION ? x.ivFieldJ : x.ivField;   // R's own first record of the interval
          while (true) {
            if (x.flags & (4 << mode)) exon_add(e, x);
            if (++i >= hEnd) break;
            x = V.hx[i];
            if (!(x.flags & 1) || x.grp != g || x.half != h || x.ivRank != iv) break;
          }
          if (pairShape && nColl < 2) { gs2[nColl] = gs; ge2[nColl] = ge; field2[nColl] = field; nColl++; }
          else if (!interval_features(A, C, V, P, mode, e, gs, ge, field, fs, frag)) bad = true;
        }
        if (pairShape) {
          bool done = nColl == 2 && junction_pair(A, V, e2[0], gs2[0], ge2[0], field2[0], e2[1], gs2[1], ge2[1], field2[1], h == 1 ? s1 : s2, bad);
          if (done) continue;
          for (int c = 0; c < nColl && !bad; c++) if (!interval_features(A, C, V, P, mode, e2[c], gs2[c], ge2[c], field2[c], fs, frag)) bad = true;
          if (bad) break;
        }
        conforming_features(fs, mode, maxField, h == 1 ? s1 : s2);
      }
      if (!bad) merge_alignment(R.flags);
    }
  } else {
    const bool useG = trFrag && mode == MODE_GENE;
    for (uint32_t r = a0; r < a1 && !bad; r++) {
      const Aln& R = V.alns[r]; if (R.hit0 == R.hit1 || (useG ? V.alnG[r].repr : R.repr) != r) continue;
      s1.n = s2.n = 0;
      if (R.flags & 2) bad |= !half_features(A, C, V, P, mode, r, 1, trFrag, s1, frag);
      if ((R.flags & 4) && !bad) bad |= !half_features(A, C, V, P, mode, r, 2, trFrag, s2, frag);
      if (!bad) merge_alignment(R.flags);
    }
  }
  if (bad) { if (retry) return -1; raise(C.err, DERR_CAPACITY, frag); return 2; }
  if (P.unambiguous[mode] && mode != MODE_EXON && nf >= 2) return 2;   // :371-382 returns 0
  bool added = false;
  for (int q = 0; q < nf; q++) {
    int k = k1[q] > k2[q] ? k1[q] : k2[q];
    double t = 0; for (int i = 0; i < k && t < 1.0; i++) t = t + addW;     // repeated addition (:100-111, :343); min(1, max(w1, w2)) (:427-428)
    double c = t < 1.0 ? t : 1.0;
    if (c > 0) added = true;
    wsum += (uint64_t)(c * 4294967296.0 + 0.5);
    if (fid[q] < (uint32_t)A.nFeat[mode]) atomic_add_u64(&V.acc[mode][fid[q]], (unsigned long long)(c * (double)(1ULL << EQP_FIXED_SHIFT_D) + 0.5));
  }
  return added ? 1 : 2;
}

}  // namespace eqpb
