// a12-a16, a18: intersection records -> genomic intervals, per-alignment canonical form (per-fragment device function).
// Reference: BedRecord ctor (J/bed-source/BedRecord.java:118-260), BedRecord.compareTo (:351-387),
// computeGenomicInterval (:404-546, :555-578), isConformingAlignment (:646-738), GenomicInterval ctor
// (GenomicInterval.java:119-157) and compareTo (:597-649), GenomicAlignment ctor (GenomicAlignment.java:92-454).
#pragma once
#include "eqp_blocks.h"

namespace eqpb {

struct RunParams {                 // eqp_params, device form
  int32_t overlap[3];              // gene, exon, junction overlap thresholds
  uint8_t spliceConforming[3], containment[3], unambiguous[3];
  uint8_t unweighted; uint8_t modes;
  uint8_t stranded;                // 0 none, 1 forward, 2 backward (compute-counts.sh:372-387)
  uint8_t genomic;                 // the one input is a genomic aligner's SAM (compute-genomic-counts.sh): weights = its own alignment counts
  double weightThreshold, slack;
};

struct IvAdj { int32_t ld, orig; };  // a17 scratch of an interval (kept at its first record): current lengthDifference, original start

struct CountView {
  HitX* hx; Aln* alns;             // hx parallel to hits; alns indexed by alnOff[slot] + alignment number
  AlnG* alnG;                      // gene-run dedupe of fragments with tr- exon hits (null when the annotation has no tr- exons)
  SlotPack* pack;                  // per slot: ranges, weight-stage inputs and outputs
  IvAdj* ivAdj;                    // parallel to hits; only touched for fragments with a positive length difference
  uint32_t* dedupTab; uint32_t* alnTail;   // a19 scratch: per-fragment representative table (2 per alignment, zeroed), chain tails
  unsigned long long* acc[3];      // fixed-point accumulators per mode (nFeat entries)
  const uint8_t* junctionActive;   // per junction feature
  unsigned long long* stats;       // [mode*3 + {objects, included, excluded}], [9] = mapped fragments
};

struct RecKey { int readIdx, readStart, eStart, eEnd, eStrand, row; };
EQP_HD RecKey rec_key(const AnnoView& A, const ChunkView& C, const Hit& h) {
  const Block b = C.blocks[h.block]; RecKey k;
  k.readIdx = b.readIdx; k.readStart = b.start + 1; k.row = h.row;
  const RowA r = A.row[h.row]; const ExonA e = A.ex[r.exon()];
  int es = r.start + 1, ee = r.end;
  int refDiff = ee - es + 1, genDiff = e.end - e.start + 1;
  if (es == 1 && refDiff < genDiff && !(r.flags() & 1)) es -= genDiff - refDiff;             // BedRecord.java:136-141
  k.eStart = es; k.eEnd = ee; k.eStrand = (int)((r.flags() >> 1) & 3);
  return k;
}
EQP_HD int rec_cmp(const RecKey& a, const RecKey& b) {                 // BedRecord.compareTo for records of one alignment
  if (a.readIdx != b.readIdx) return a.readIdx < b.readIdx ? -1 : 1;
  if (a.readStart != b.readStart) return a.readStart < b.readStart ? -1 : 1;
  if (a.eStart != b.eStart) return a.eStart < b.eStart ? -1 : 1;
  if (a.eEnd != b.eEnd) return a.eEnd < b.eEnd ? -1 : 1;
  if (a.eStrand != b.eStrand) return a.eStrand < b.eStrand ? -1 : 1;
  return 0;
}

// lengthDifference of a record: transcript exon length (after the 5' fix) minus genome exon length; genomic records have none
EQP_HD int record_ld(const AnnoView& A, const ChunkView& C, const Hit& h) {
  const RowA r = A.row[h.row];
  if (r.flags() & 1) return 0;
  RecKey rk = rec_key(A, C, h); const ExonA e = A.ex[r.exon()];
  return (rk.eEnd - rk.eStart + 1) - (e.end - e.start + 1);
}

EQP_HD uint64_t mix64(uint64_t h, uint64_t v) { h ^= v + 0x9E3779B97F4A7C15ULL + (h << 6) + (h >> 2); h *= 0xBF58476D1CE4E5B9ULL; return h ^ (h >> 29); }

// Build one half (read index) of an alignment: hits [h0,h1) are sorted by rec_cmp; duplicates are invalidated.
// trBits: bit0 the half has a tr- interval, bit1 all its intervals are tr-.
EQP_HD void build_half(const AnnoView& A, const ChunkView& C, const CountView& V, const RunParams& P, uint32_t h0, uint32_t h1, int half,
                       int32_t& gaStart, uint16_t& nIv, int16_t& maxField, uint8_t& trBits, uint32_t frag) {
  // first-wins dedupe (TreeSet<BedRecord>)
  uint32_t n = 0;
  for (uint32_t k = h0; k < h1; k++) {
    bool dup = false;
    if (k > h0) { RecKey a = rec_key(A, C, C.hits[k - 1]), b = rec_key(A, C, C.hits[k]); dup = rec_cmp(a, b) == 0; }
    V.hx[k].flags = dup ? 0 : 1; V.hx[k].half = (uint8_t)half; V.hx[k].grp = 0;
    if (!dup) n++;
  }
  const Block& b0 = C.blocks[C.hits[h0].block];
  bool genomic = A.row[C.hits[h0].row].flags() & 1; bool junction = A.bktFlags[b0.bucket] & 1, inv = A.bktFlags[b0.bucket] & 2;
  int i = -1, iJ = -1; int prevExon = -1; Block prevCopy{}; const Block* prevBlk = nullptr; bool anyTr = false, anyLd = false; int trSign = 1;
  for (uint32_t k = h0; k < h1; k++) {
    HitX& x = V.hx[k]; if (!(x.flags & 1)) continue;
    const Hit h = C.hits[k]; const Block b = C.blocks[h.block]; const RowA row = A.row[h.row]; const int ex = row.exon(); const ExonA exn = A.ex[ex];
    if (((row.flags() & 1) != 0) != genomic) raise(C.err, DERR_INCONSISTENT, frag);                 // "Change of BED record type"
    if (prevBlk && b.strand != prevBlk->strand) raise(C.err, DERR_INCONSISTENT, frag);               // differing strand
    const bool tr = exn.chr <= -2;                                       // tr- transcript exon (Exon.java:230-301): coordinates on the transcript
    int exS = exn.start, exE = exn.end; char ostrand = (char)exn.strand;
    int readStart = b.start + 1, readEnd = b.end;
    RecKey rk = rec_key(A, C, h);
    int gs, ge, ld = 0, fieldIdx, fieldJ; char ivStrand; bool first, last;
    if (genomic) {                                                       // :555-578
      if (prevBlk && prevBlk->maxSeg != b.maxSeg) raise(C.err, DERR_INCONSISTENT, frag);
      gs = readStart; ge = readEnd; ivStrand = (char)b.strand; fieldIdx = fieldJ = b.seg; first = b.seg == 0; last = b.seg == b.maxSeg;
    } else {                                                             // :404-546
      if (prevBlk && (prevBlk->start != b.start || prevBlk->end != b.end || prevBlk->L != b.L || prevBlk->I != b.I || prevBlk->D != b.D))
        raise(C.err, DERR_INCONSISTENT, frag);                           // "changed alignment" / differing L, I, D
      if (prevExon == ex) raise(C.err, DERR_INCONSISTENT, frag);         // "Unchanged exon"
      i++; if (!tr) iJ++;                                                // the junction run gives tr- records no index (GenomicAlignment.java:233-234)
      int aS = rk.eStart, aE = rk.eEnd; int diffT = aE - aS + 1, diffG = exE - exS + 1; ld = diffT - diffG;
      if (ld != 0) {
        anyLd = true;
        if (junction && i == 0) aS = aE - diffG + 1;
        else if (i == (int)n - 1) aE = aS + diffG - 1;
        else if (i == 0) aS = aE - diffG + 1;
        else { aS = aE - diffG / 2 + 1; aE = aS + (diffG - diffG / 2) - 1; }
      }
      if (ld > 0) C.slotLd[b.slot] = 1;                                  // a17 can move intervals of this fragment (GenomicInterval.java:330-334)
      int ovS = readStart > aS ? readStart : aS, ovE = readEnd < aE ? readEnd : aE;
      int sd = ovS - aS; if (sd < 0) sd = -sd; int ed = ovE - aE; if (ed < 0) ed = -ed;
      char exonStrand = '+';
      if (junction) ivStrand = (char)b.strand;
      else { exonStrand = inv ? (ostrand == '+' ? '-' : '+') : ostrand; ivStrand = exonStrand == (char)b.strand ? '+' : '-'; }
      if (tr) { anyTr = true; trSign = exonStrand == '+' ? 1 : -1; C.slotTr[b.slot] = 1; }
      int od = h.overlap - 1;
      if (sd <= ed) {
        if (exonStrand == '+') { gs = exS + sd; ge = exS + sd + od; } else { gs = exE - sd - od; if (gs < 1) gs = 1; ge = exE - sd; }
      } else {
        if (exonStrand == '+') { gs = exE - ed - od; if (gs < 1) gs = 1; ge = exE - ed; } else { gs = exS + ed; ge = exS + ed + od; }
      }
      fieldIdx = i; fieldJ = tr ? -1 : iJ; first = last = false;
    }
    if (fieldIdx > 126) { raise(C.err, DERR_CAPACITY, frag); fieldIdx = fieldJ = 126; }
    // isConformingAlignment per run (the gene/exon/junction runs differ in -C / -N)
    uint8_t conf = 0;
    for (int m = 0; m < 3; m++) {
      bool c = !(P.stranded && ostrand != ivStrand);                       // BedRecord.java:650-656
      if (c && genomic && (P.spliceConforming[m] || P.containment[m])) {
        if (exS != rk.eStart || exE != rk.eEnd) { raise(C.err, DERR_GENOME_COORD, frag); }
        if (b.L - b.I + b.D < h.overlap) { raise(C.err, DERR_GENOME_COORD, frag); }
        if (P.spliceConforming[m]) {
          if (b.L - b.I + b.D > h.overlap) c = false;
          else if (first && exS > readStart) c = false;
          else if (!first && exS != readStart) c = false;
          else if (last && exE < readEnd) c = false;
          else if (!last && exE != readEnd) c = false;
        } else c = exS <= readStart && readEnd <= exE;
      }
      int thr = P.overlap[m] < (exE - exS + 1) ? P.overlap[m] : (exE - exS + 1);
      if (c && h.overlap >= thr) conf |= (uint8_t)(4 << m);              // GenomicInterval.java:144
    }
    { HitX y; y.gs = gs; y.ge = ge; y.chr = exn.chr; y.ivField = (int8_t)fieldIdx; y.ivFieldJ = (int8_t)fieldJ; y.ivRank = 0; y.exon = ex;
      y.flags = (uint8_t)(1 | (genomic ? 2 : 0) | conf | (ivStrand == '-' ? 32 : 0)); y.half = (uint8_t)half; y.grp = 0;
      x = y; }                                                           // one 24-byte store
    prevCopy = b; prevBlk = &prevCopy; prevExon = ex;
  }
  // With a tr- record the junction run numbers the later records differently, and the adapted coordinates of a record with a length
  // difference depend on its number (:417-452): the two runs would then need different coordinates.  Not handled.
  if (anyTr && anyLd) raise(C.err, DERR_TR_LENDIFF, frag);
  // distinct intervals: rank in GenomicInterval.compareTo order (GenomicInterval.java:597-649): (chr,start,end,strand), or, as soon as a
  // tr- interval takes part in the comparison, the records' transcript order; the kept interval carries the field index of its first record
  gaStart = 0x7fffffff; nIv = 0; uint32_t nTrIv = 0;
  for (uint32_t k = h0; k < h1; k++) {
    HitX& x = V.hx[k]; if (!(x.flags & 1)) continue;
    if (x.gs < gaStart) gaStart = x.gs;
    const bool xTr = x.chr <= -2; const int xKey = anyTr ? trSign * rec_key(A, C, C.hits[k]).eStart : 0; const int sx = (x.flags >> 5) & 1;
    auto same = [&](uint32_t j, const HitX& y, bool yTr, int yKey) {
      (void)j; return (xTr || yTr) ? yKey == xKey : (y.chr == x.chr && y.gs == x.gs && y.ge == x.ge && ((y.flags >> 5) & 1) == sx); };
    int rank = 0, ord = 0; bool firstOfGroup = true; int8_t fld = x.ivField, fldJ = x.ivFieldJ;
    for (uint32_t j = h0; j < h1; j++) {
      const HitX& y = V.hx[j]; if (!(y.flags & 1) || j == k) continue;
      const bool yTr = y.chr <= -2; const int yKey = anyTr ? trSign * rec_key(A, C, C.hits[j]).eStart : 0; const int sy = (y.flags >> 5) & 1;
      bool eq = same(j, y, yTr, yKey);
      bool less = (xTr || yTr) ? yKey < xKey : (y.chr != x.chr ? y.chr < x.chr : y.gs != x.gs ? y.gs < x.gs : y.ge != x.ge ? y.ge < x.ge : sy < sx);
      if (eq) { if (j < k) { firstOfGroup = false; } }
      else if (less) {                                                   // count distinct smaller intervals: only their first record
        bool yFirst = true;
        for (uint32_t q = h0; q < j; q++) {
          const HitX& z = V.hx[q]; if (!(z.flags & 1)) continue;
          const bool zTr = z.chr <= -2;
          bool zy = (zTr || yTr) ? (trSign * rec_key(A, C, C.hits[q]).eStart == yKey) : (z.chr == y.chr && z.gs == y.gs && z.ge == y.ge && ((z.flags >> 5) & 1) == sy);
          if (zy) { yFirst = false; break; }
        }
        if (yFirst) { rank++; if (!yTr) ord++; }
      }
    }
    if (!firstOfGroup) for (uint32_t j = h0; j < k; j++) { const HitX& y = V.hx[j]; if (!(y.flags & 1)) continue; const bool yTr = y.chr <= -2;
      if (same(j, y, yTr, anyTr ? trSign * rec_key(A, C, C.hits[j]).eStart : 0)) { fld = y.ivField; fldJ = y.ivFieldJ; break; } }
    if (rank > 255) { raise(C.err, DERR_CAPACITY, frag); rank = 255; }
    x.ivRank = (uint16_t)rank; x.ivField = fld; x.ivFieldJ = fldJ;
    x.grp = xTr ? (uint16_t)0xFFFFu : (uint16_t)ord;                     // ordinal among the non-tr intervals (see HitX::grp)
    if (firstOfGroup) { nIv++; if (xTr) nTrIv++; }
  }
  trBits = (uint8_t)((nTrIv ? 1 : 0) | (nTrIv && nTrIv == nIv ? 2 : 0));
  maxField = genomic ? (int16_t)b0.maxSeg : (int16_t)((int)n - 1);
}

}  // namespace eqpb
