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

// ---- stage A (one alignment per item): sort the alignment's hits, build both halves, hash the canonical key
EQP_HD void alignment_build(const AnnoView& A, const ChunkView& C, const CountView& V, const RunParams& P, uint32_t a) {
  uint32_t b0 = C.alnBlk[a], b1 = C.alnBlk[a + 1];
  uint32_t k = C.hitCnt[b0], e = C.hitCnt[b1];                           // hits are grouped by block, blocks by alignment
  Aln& al = V.alns[a]; al.hit0 = al.hit1 = al.split = 0; al.flags = 0; al.repr = a; al.key = 0; al.next = 0xFFFFFFFFu;
  if (k == e) return;
  uint32_t frag = C.base + C.blocks[b0].slot;
  for (uint32_t i = k + 1; i < e; i++) {                                 // insertion sort by (readIdx, readStart, eRefStart, eRefEnd, strand, row)
    Hit cur = C.hits[i]; RecKey ck = rec_key(A, C, cur); uint32_t j = i;
    while (j > k) { RecKey pk = rec_key(A, C, C.hits[j - 1]); int c = rec_cmp(pk, ck); if (c < 0 || (c == 0 && pk.row <= ck.row)) break; C.hits[j] = C.hits[j - 1]; j--; }
    C.hits[j] = cur;
  }
  uint32_t split = k; while (split < e && C.blocks[C.hits[split].block].readIdx == 1) split++;
  al.hit0 = k; al.split = split; al.hit1 = e;
  const Block& fb = C.blocks[C.hits[k].block];
  uint8_t fl = fb.paired ? 1 : 0;
  for (uint32_t i = k; i < e; i++) if ((C.blocks[C.hits[i].block].paired ? 1 : 0) != (fl & 1)) raise(C.err, DERR_INCONSISTENT, frag);
  al.start1 = al.start2 = 0x7fffffff; al.nIv1 = al.nIv2 = 0; al.maxField1 = al.maxField2 = -1;
  if (split > k) { build_half(A, C, V, P, k, split, 1, al.start1, al.nIv1, al.maxField1, frag); fl |= 2; }
  if (e > split) { build_half(A, C, V, P, split, e, 2, al.start2, al.nIv2, al.maxField2, frag); fl |= 4; }
  if (!(fl & 1)) fl |= (uint8_t)(((fl & 2) ? 1 : 2) << 3);
  if (!(fl & 1) && (fl & 2) && (fl & 4)) raise(C.err, DERR_INCONSISTENT, frag);     // single-read alignment with both read indices
  al.flags = fl;
  // canonical key (strand slot is empty in the reference's strings: GenomicInterval never sets originalStrand)
  bool both = (fl & 2) && (fl & 4); bool secondFirst = both && !(al.start1 <= al.start2);
  uint64_t hsh = mix64(0x1234, (uint64_t)((fl & 1) ? 0 : (fl >> 3) & 3));
  hsh = mix64(hsh, both ? 0 : (fl & 2) ? 1 : 2);
  for (int part = 0; part < 2; part++) {
    int half = both ? (secondFirst ? 2 - part : 1 + part) : ((fl & 2) ? 1 : 2);
    if (!both && part == 1) break;
    uint32_t s0 = half == 1 ? k : split, s1 = half == 1 ? split : e; int nIv = half == 1 ? al.nIv1 : al.nIv2;
    hsh = mix64(hsh, 0xABCDEF00u + (uint64_t)nIv);
    for (int r = 0; r < nIv; r++)
      for (uint32_t i = s0; i < s1; i++) { const HitX& x = V.hx[i]; if ((x.flags & 1) && x.ivRank == r) { hsh = mix64(mix64(mix64(hsh, (uint64_t)x.chr), (uint64_t)x.gs), (uint64_t)x.ge); break; } }
  }
  al.key = hsh;
}

// interval r of a half: first valid record with that rank
EQP_HD bool half_interval(const CountView& V, uint32_t s0, uint32_t s1, int r, int32_t& chr, int32_t& gs, int32_t& ge, int& field) {
  for (uint32_t i = s0; i < s1; i++) { const HitX& x = V.hx[i]; if ((x.flags & 1) && x.ivRank == r) { chr = x.chr; gs = x.gs; ge = x.ge; field = x.ivField; return true; } }
  return false;
}
EQP_HD bool halves_equal(const CountView& V, const Aln& a, int ha, const Aln& b, int hb) {
  int na = ha == 1 ? a.nIv1 : a.nIv2, nb = hb == 1 ? b.nIv1 : b.nIv2;
  if (na != nb) return false;
  uint32_t a0 = ha == 1 ? a.hit0 : a.split, a1 = ha == 1 ? a.split : a.hit1, b0 = hb == 1 ? b.hit0 : b.split, b1 = hb == 1 ? b.split : b.hit1;
  for (int r = 0; r < na; r++) {
    int32_t c1, s1, e1, c2, s2, e2; int f;
    half_interval(V, a0, a1, r, c1, s1, e1, f); half_interval(V, b0, b1, r, c2, s2, e2, f);
    if (c1 != c2 || s1 != s2 || e1 != e2) return false;
  }
  return true;
}
// string equality of two WeightObjectAlignment keys; on success `swapped` tells how b's halves map onto a's
EQP_HD bool keys_equal(const CountView& V, const Aln& a, const Aln& b, bool& swapped) {
  swapped = false;
  if (a.key != b.key) return false;
  if ((a.flags & 1) != (b.flags & 1)) return false;
  if (!(a.flags & 1) && ((a.flags >> 3) & 3) != ((b.flags >> 3) & 3)) return false;
  if ((a.flags & 6) != (b.flags & 6)) return false;
  if ((a.flags & 6) == 6) {
    bool sa = !(a.start1 <= a.start2), sb = !(b.start1 <= b.start2);     // which half is printed first
    int a1 = sa ? 2 : 1, a2 = sa ? 1 : 2, b1 = sb ? 2 : 1, b2 = sb ? 1 : 2;
    if (!halves_equal(V, a, a1, b, b1) || !halves_equal(V, a, a2, b, b2)) return false;
    // combineExons (:312-337): straight pairing if ga1==ga1' and ga2==ga2', else the mate-swapped pairing
    swapped = !(halves_equal(V, a, 1, b, 1) && halves_equal(V, a, 2, b, 2));
    return true;
  }
  int h = (a.flags & 2) ? 1 : 2;
  return halves_equal(V, a, h, b, h);
}

EQP_HD uint64_t hit_group_key(const HitX& x) { return ((uint64_t)((x.flags & 1) ? 0 : 1) << 40) | ((uint64_t)x.grp << 24) | ((uint64_t)x.half << 16) | x.ivRank; }

// ---- stage B: dedupe, a20, final weight
EQP_HD void fragment_weight(const AnnoView& A, const ChunkView& C, const CountView& V, const RunParams& P, uint32_t slot) {
  uint32_t b0 = C.blkOff[slot], b1 = C.blkOff[slot + 1];
  V.a20[slot] = 0; V.nFinal[slot] = 0;
  if (b0 == b1 || C.hitCnt[b0] == C.hitCnt[b1]) return;
  uint32_t a0 = C.alnOff[slot], a1 = C.alnOff[slot + 1];
  int left = 0, right = 0; uint32_t nA = a1 - a0, cap = 2 * nA; uint32_t* tab = V.dedupTab + 2 * (uint64_t)a0;
  for (uint32_t a = a0; a < a1; a++) {
    Aln& al = V.alns[a]; if (al.hit0 == al.hit1) continue;
    al.repr = a;
    uint32_t found = 0xFFFFFFFFu; bool sw = false;
    if (nA <= 4) {                                                         // tiny fragment: compare with the earlier representatives directly
      for (uint32_t p = a0; p < a; p++) {
        const Aln& q = V.alns[p]; if (q.hit0 == q.hit1 || q.repr != p) continue;
        if (keys_equal(V, q, al, sw)) { found = p; break; }
      }
    } else {                                                               // per-fragment open-addressing table of representatives (2 slots per alignment, zeroed)
      uint32_t s = (uint32_t)(((al.key * 0x9E3779B97F4A7C15ULL) >> 32) % cap);
      while (true) {
        uint32_t e = tab[s];
        if (!e) { tab[s] = a - a0 + 1; break; }
        if (keys_equal(V, V.alns[a0 + e - 1], al, sw)) { found = a0 + e - 1; break; }   // insertion order along the probe path = file order: first equal key wins
        s = s + 1 == cap ? 0 : s + 1;
      }
    }
    if (found != 0xFFFFFFFFu) {                                            // duplicate: append to the representative's chain (file order)
      al.repr = found; if (sw) al.flags |= 32;
      V.alns[V.alnTail[found]].next = a; V.alnTail[found] = a;
    } else V.alnTail[a] = a;
    if (al.repr == a) { if ((al.flags & 6) == 4) right++; else left++; }   // ":r-" forms (WeightObject.java:859-865)
  }
  // tag every hit with its representative (index within the fragment) and with the representative's half it merges into
  // (combineExons pairs mate-swapped duplicates crosswise, WeightObjectAlignment.java:312-337): a22-a25 then read the
  // fragment's hits as one flat range instead of walking duplicate chains
  if (nA > 0xFFFFu) { raise(C.err, DERR_CAPACITY, C.base + slot); return; }
  for (uint32_t a = a0; a < a1; a++) {
    const Aln& al = V.alns[a]; if (al.hit0 == al.hit1) continue;
    bool sw = al.repr != a && (al.flags & 32); uint16_t g = (uint16_t)(al.repr - a0);
    for (uint32_t i = al.hit0; i < al.hit1; i++) { HitX& x = V.hx[i]; x.grp = g; if (sw) x.half = (uint8_t)(3 - x.half); }
  }
  // small fragments: regroup the hits by (representative, half, interval), stable, dropped duplicates last: the counting runs
  // then stream each fragment's hits once.  (hx is only indexed through Aln.hit0.. before this point; large fragments keep
  // their order and walk the duplicate chains.)
  { uint32_t h0 = C.hitCnt[b0], h1 = C.hitCnt[b1];
    bool sorted = true; uint64_t prev = 0;
    if (h1 - h0 <= EQP_FLAT_SCAN) for (uint32_t i = h0; i < h1; i++) { uint64_t k = hit_group_key(V.hx[i]); sorted &= k >= prev; prev = k; }   // independent loads
    if (h1 - h0 <= EQP_FLAT_SCAN && !sorted)
      for (uint32_t i = h0 + 1; i < h1; i++) {
        HitX cur = V.hx[i]; uint64_t ck = hit_group_key(cur); uint32_t j = i;
        while (j > h0 && hit_group_key(V.hx[j - 1]) > ck) { V.hx[j] = V.hx[j - 1]; j--; }
        if (j != i) V.hx[j] = cur;
      } }
  uint32_t a20 = (uint32_t)(left > right ? left : right);
  V.a20[slot] = a20;
  // final.wgt entry: BED weight (single-read, read index 1, n = a20) vs genome SAM weight, T+J entry as class arbiter
  FragStat bed{}; bed.nSr1 = (uint16_t)(a20 > 65535 ? 65535 : a20); bed.flags = FS_EXISTS | (1 << 2);
  const FragStat& g = C.st[FILE_G][slot];
  uint32_t n = a20;
  if (g.flags & FS_EXISTS) {
    uint32_t ax = C.auxSlot[slot]; int cm = 0;
    if (ax != 0xFFFFFFFFu) cm = fe_cmp_mult(C.stTJ[ax], g);               // aux.compareMultiplicity(entry2), CombineReadWeightFiles.java:192-195
    if (cm == 1) n = a20; else if (cm == -1) n = (uint32_t)fe_n(g); else n = (int)a20 >= fe_n(g) ? a20 : (uint32_t)fe_n(g);
  }
  V.nFinal[slot] = n;
}

// ---- stage C: conforming features + accumulation for one run (mode)
#define EQP_FS_CAP 160
#define EQP_FT_CAP 320
struct FeatSet { uint32_t id[EQP_FS_CAP]; uint32_t mask[EQP_FS_CAP]; int n; };
EQP_HD bool fs_add(FeatSet& s, uint32_t id, uint32_t bit) {
  for (int i = 0; i < s.n; i++) if (s.id[i] == id) { s.mask[i] |= bit; return true; }
  if (s.n >= EQP_FS_CAP) return false;
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
    for (int q = 0; q < ne; q++) if (!exGen[q]) { int en = A.exEnd[exon[q]] < ge ? A.exEnd[exon[q]] : ge; int ov = en - gs + 1; if (ov < 0) ov = 0; if (ov > maxOv) maxOv = ov; }
    if (maxOv > 0) { int w = 0; for (int q = 0; q < ne; q++) { int en = A.exEnd[exon[q]] < ge ? A.exEnd[exon[q]] : ge; int ov = en - gs + 1; if (ov < 0) ov = 0; if (exGen[q] || ov >= maxOv) { exon[w] = exon[q]; exGen[w] = exGen[q]; w++; } } ne = w; }
  }
  e.n = ne; return ne;
}
// Junction run, a half with exactly two intervals of one exon each (the ordinary spliced read): a junction conforms iff it is a
// candidate of both exons and the two fields are consecutive (:679-701), i.e. the intersection of the two exons' feature lists.
// The lists are sorted by id when the annotation is loaded, so this is one merge instead of a set with field masks.
// Returns false when the shape is different (the general path then handles the half).
EQP_HD bool junction_pair(const AnnoView& A, const CountView& V, ExonSet& ea, int32_t gsa, int32_t gea, int fa, ExonSet& eb, int32_t gsb, int32_t geb, int fb, FeatSet& out, bool& overflow) {
  if (ea.over || eb.over) return false;
  if (exon_filter(A, ea, gsa, gea) != 1 || exon_filter(A, eb, gsb, geb) != 1) return false;   // the filter is idempotent: the general path may run it again
  out.n = 0;
  if (fa - fb != 1 && fb - fa != 1) return true;
  const uint32_t* F = A.feat[MODE_JUNCTION]; const uint32_t* O = A.featOff[MODE_JUNCTION];
  uint32_t i = O[ea.exon[0]], i1 = O[ea.exon[0] + 1], j = O[eb.exon[0]], j1 = O[eb.exon[0] + 1];
  while (i < i1 && j < j1) {
    uint32_t va = F[i], vb = F[j], ia = va & 0x7fffffffu, ib = vb & 0x7fffffffu;
    if (ia < ib) i++; else if (ib < ia) j++;
    else {
      bool act = ((va & vb) >> 31) != 0 || V.junctionActive[ia] != 0;    // a row survives the chunk's map filter (a27) or is in the transcript map
      bool pa = (va >> 31) || act, pb = (vb >> 31) || act;
      if (pa && pb) { if (out.n >= EQP_FS_CAP) { overflow = true; return true; } out.id[out.n] = ia; out.mask[out.n] = 0; out.n++; }
      i++; j++;
    }
  }
  return true;
}
EQP_HD bool interval_features(const AnnoView& A, const ChunkView& C, const CountView& V, const RunParams& P, int mode, ExonSet& e, int32_t gs, int32_t ge, int field,
                              FeatSet& fs, uint32_t frag) {
  if (e.over) { raise(C.err, DERR_CAPACITY, frag); return false; }
  int ne = exon_filter(A, e, gs, ge); int32_t* exon = e.exon;
  if (mode == MODE_EXON && P.unambiguous[mode] && ne >= 2) ne = 0;      // -U with genomicIntervalMode (:603-608)
  for (int q = 0; q < ne; q++) {
    int ex = exon[q]; uint32_t f0 = A.featOff[mode][ex], f1 = A.featOff[mode][ex + 1]; bool any = false;
    for (uint32_t f = f0; f < f1; f++) {
      uint32_t v = A.feat[mode][f]; uint32_t id = v & 0x7fffffffu;
      if (mode == MODE_JUNCTION && !(v & 0x80000000u) && !V.junctionActive[id]) continue;   // row removed by the per-chunk map filter (a27)
      any = true;
      if (!fs_add(fs, id, 1u << field)) { raise(C.err, DERR_CAPACITY, frag); return false; }
    }
    if (!any && !fs_add(fs, (uint32_t)A.nFeat[mode] + (uint32_t)ex, 1u << field)) { raise(C.err, DERR_CAPACITY, frag); return false; }   // unknown exon: own pseudo id
  }
  return true;
}
// field-coverage test of a half's candidate features (GenomicAlignment.java:570-712)
EQP_HD void conforming_features(const FeatSet& fs, int mode, int maxField, FeatSet& out) {
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
// (fragments with more than EQP_FLAT_SCAN hits; smaller ones are regrouped by fragment_weight and streamed, see fragment_count)
EQP_HD void half_features(const AnnoView& A, const ChunkView& C, const CountView& V, const RunParams& P, int mode, uint32_t r, int h, FeatSet& out, uint32_t frag) {
  const Aln& R = V.alns[r]; out.n = 0;
  int nIv = h == 1 ? R.nIv1 : R.nIv2; int maxField = h == 1 ? R.maxField1 : R.maxField2;
  if (maxField >= 32) { raise(C.err, DERR_CAPACITY, frag); return; }
  if (mode == MODE_JUNCTION && nIv < 2) return;                            // a junction needs two consecutive fields (:679-701)
  FeatSet fs; fs.n = 0;
  for (int iv = 0; iv < nIv; iv++) {
    int32_t gs = 0, ge = 0; int field = 0; bool haveIv = false; ExonSet e; e.n = 0; e.over = false;
    for (uint32_t d = r; d != 0xFFFFFFFFu; d = V.alns[d].next) {          // R first, then later duplicates in file order
      const Aln& D = V.alns[d];
      for (uint32_t i = D.hit0; i < D.hit1; i++) {
        const HitX& x = V.hx[i];
        if (x.half != h || !(x.flags & 1) || x.ivRank != iv) continue;
        if (!haveIv) { gs = x.gs; ge = x.ge; field = x.ivField; haveIv = true; }   // R's own first record of the interval
        if (x.flags & (4 << mode)) exon_add(e, x);
      }
    }
    if (!interval_features(A, C, V, P, mode, e, gs, ge, field, fs, frag)) return;
  }
  conforming_features(fs, mode, maxField, out);
}

EQP_HD void atomic_add_u64(unsigned long long* p, unsigned long long v) {
#if defined(__CUDA_ARCH__)
  atomicAdd(p, v);
#else
  *p += v;
#endif
}

// returns 0 = the fragment is no weight object, 1 = it added to a count, 2 = it was excluded (or aborted on a capacity error)
EQP_HD int fragment_count(const AnnoView& A, const ChunkView& C, const CountView& V, const RunParams& P, int mode, uint32_t slot) {
  uint32_t a20 = V.a20[slot];
  if (a20 == 0) return 0;
  uint32_t frag = C.base + slot; uint32_t a0 = C.alnOff[slot], a1 = C.alnOff[slot + 1];
  const uint32_t hEnd = C.hitCnt[C.blkOff[slot + 1]];                    // end of the fragment's hit range
  double w = 1.0 / (double)V.nFinal[slot];
  if (w > 1.0 / (double)a20) w = 1.0 / (double)a20;                      // ComputeCounts.java:762-764
  if (!(w >= P.weightThreshold)) return 2;
  double addW = P.unweighted ? 1.0 : w;
  uint32_t fid[EQP_FT_CAP]; uint16_t k1[EQP_FT_CAP], k2[EQP_FT_CAP]; int nf = 0;
  FeatSet s1, s2; bool bad = false;
  // one representative alignment done: paired -> gene run intersects the halves, the others unite them (:325-335); then count per feature
  auto merge_alignment = [&](uint8_t flags) {
    if (flags & 1) {
      if (mode == MODE_GENE) { int wn = 0; for (int i = 0; i < s1.n; i++) { bool in2 = false; for (int j = 0; j < s2.n; j++) in2 |= s2.id[j] == s1.id[i]; if (in2) s1.id[wn++] = s1.id[i]; } s1.n = wn; }
      else for (int j = 0; j < s2.n; j++) { bool in1 = false; for (int i = 0; i < s1.n; i++) in1 |= s1.id[i] == s2.id[j]; if (!in1) { if (s1.n >= EQP_FS_CAP) { bad = true; return; } s1.id[s1.n++] = s2.id[j]; } }
      s2.n = 0;
    }
    for (int side = 0; side < 2; side++) {
      FeatSet& s = side ? s2 : s1;
      for (int i = 0; i < s.n; i++) {
        int q = 0; while (q < nf && fid[q] != s.id[i]) q++;
        if (q == nf) { if (nf >= EQP_FT_CAP) { bad = true; return; } fid[nf] = s.id[i]; k1[nf] = k2[nf] = 0; nf++; }
        if (side) k2[q]++; else k1[q]++;
      }
    }
  };
  const uint32_t hBeg = C.hitCnt[C.blkOff[slot]];
  if (hEnd - hBeg <= EQP_FLAT_SCAN) {
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
          const int32_t gs = x.gs, ge = x.ge; const int field = x.ivField;   // R's own first record of the interval
          while (true) {
            if (x.flags & (4 << mode)) exon_add(e, x);
            if (++i >= hEnd) break;
            x = V.hx[i];
            if (!(x.flags & 1) || x.grp != g || x.half != h || x.ivRank != iv) break;
          }
          if (pairShape && nColl < 2) { gs2[nColl] = gs; ge2[nColl] = ge; field2[nColl] = field; nColl++; }
          else if (!interval_features(A, C, V, P, mode, e, gs, ge, field, fs, frag)) return 2;
        }
        if (pairShape) {
          bool done = nColl == 2 && junction_pair(A, V, e2[0], gs2[0], ge2[0], field2[0], e2[1], gs2[1], ge2[1], field2[1], h == 1 ? s1 : s2, bad);
          if (done) continue;
          for (int c = 0; c < nColl; c++) if (!interval_features(A, C, V, P, mode, e2[c], gs2[c], ge2[c], field2[c], fs, frag)) return 2;
        }
        conforming_features(fs, mode, maxField, h == 1 ? s1 : s2);
      }
      if (!bad) merge_alignment(R.flags);
    }
  } else {
    for (uint32_t r = a0; r < a1 && !bad; r++) {
      const Aln& R = V.alns[r]; if (R.hit0 == R.hit1 || R.repr != r) continue;
      s1.n = s2.n = 0;
      if (R.flags & 2) half_features(A, C, V, P, mode, r, 1, s1, frag);
      if (R.flags & 4) half_features(A, C, V, P, mode, r, 2, s2, frag);
      merge_alignment(R.flags);
    }
  }
  if (bad) { raise(C.err, DERR_CAPACITY, frag); return 2; }
  if (P.unambiguous[mode] && mode != MODE_EXON && nf >= 2) return 2;   // :371-382 returns 0
  bool added = false;
  for (int q = 0; q < nf; q++) {
    int k = k1[q] > k2[q] ? k1[q] : k2[q];
    double t = 0; for (int i = 0; i < k && t < 1.0; i++) t = t + addW;     // repeated addition (:100-111, :343); min(1, max(w1, w2)) (:427-428)
    double c = t < 1.0 ? t : 1.0;
    if (c > 0) added = true;
    if (fid[q] < (uint32_t)A.nFeat[mode]) atomic_add_u64(&V.acc[mode][fid[q]], (unsigned long long)(c * (double)(1ULL << EQP_FIXED_SHIFT_D) + 0.5));
  }
  return added ? 1 : 2;
}

}  // namespace eqpb
