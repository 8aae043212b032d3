// ORACLE (test infrastructure) — GenomicInterval / GenomicAlignment (a13, a14, a16).
#pragma once
#include "count.hpp"

namespace eqp {

struct GAln; struct WOA;

// Java strings that are compared with == / != against the literal "" (SURVEY H2): keep identity.
struct JStr { std::string s; bool isNull = true; bool lit = false; };

struct GInterval {                                     // J/bed-source/GenomicInterval.java:119-157
  std::string chr; int start, end; char strand; int origStart, origEnd;
  int overlap, ld, L, I, D, fieldIndex; bool conforming; bool adjusted = false;
  std::vector<Exon> exonSet; GAln* ga = nullptr; BedRec* rec = nullptr;
  bool isTr() const { return is_tr(chr); }
  JStr str(const Params& p) const {                    // toString :709-721 (strand slot empty: originalStrand never set)
    if (isTr() && p.mode != MODE_GENE) return {"", false, true};
    return {chr + "/" + std::to_string(start) + "/" + std::to_string(end) + "/", false, false};
  }
  void addExons(const std::vector<Exon>& es) {         // HashSet.addAll: first object with a key stays
    for (auto& e : es) {
      bool have = false;
      for (auto& x : exonSet) if (x.chr == e.chr && x.start == e.start && x.end == e.end && x.ostrand == e.ostrand) { have = true; break; }
      if (!have) exonSet.push_back(e);
    }
  }
};

// BedRecord.isConformingAlignment (:646-738)
inline bool is_conforming(const BedRec& b, const Params& p) {
  if (p.stranded && b.exon.ostrand != b.intervalStrand) return false;
  if (!b.isGenomicRecord()) return true;
  if (!p.spliceConforming && !p.containment) return true;
  if (b.exon.start != b.eRefStart || b.exon.end != b.eRefEnd) throw Error("Alignment against genome and exon coodinates disagree in BedRecord: " + b.readIdFull);
  if (b.L - b.I + b.D < b.overlap) throw Error("Computed overlap and intersectBed overlap disagree in BedRecord: " + b.readIdFull);
  if (p.spliceConforming) { if (b.L - b.I + b.D > b.overlap) return false; }
  else return b.exon.start <= b.readStart && b.readEnd <= b.exon.end;
  if (b.first && b.exon.start > b.readStart) return false;
  if (!b.first && std::abs(b.exon.start - b.readStart) > p.exonSlack) return false;
  if (b.last && b.exon.end < b.readEnd) return false;
  if (!b.last && std::abs(b.exon.end - b.readEnd) > p.exonSlack) return false;
  return true;
}

inline std::unique_ptr<GInterval> make_interval(const std::string& chr, int s, int e, char strand, BedRec& b, int ld, int fieldIndex,
                                                GAln* ga, bool conforming, const Params& p) {
  auto g = std::make_unique<GInterval>();
  g->chr = chr; g->start = g->origStart = s; g->end = g->origEnd = e; g->strand = strand; g->overlap = b.overlap; g->ld = ld;
  g->L = b.L; g->I = b.I; g->D = b.D; g->fieldIndex = fieldIndex; g->ga = ga; g->rec = &b;
  g->conforming = conforming && (b.overlap >= std::min(p.overlapThreshold, b.exon.length()));   // :144
  if (g->conforming) g->exonSet.push_back(b.exon);
  return g;
}

// BedRecord.computeGenomicInterval(index, maxIndex, ga) :404-546 — transcript / junction records
inline std::unique_ptr<GInterval> interval_nongenomic(BedRec& b, int index, int maxIndex, GAln* ga, const Params& p) {
  int gs = b.exon.start, ge = b.exon.end;
  int diffT = b.adaptedEnd - b.adaptedStart + 1, diffG = ge - gs + 1, ld = diffT - diffG;
  if (ld != 0) {
    if (b.isJunction() && index == 0) b.adaptedStart = b.adaptedEnd - diffG + 1;
    else if (index == maxIndex - 1) b.adaptedEnd = b.adaptedStart + diffG - 1;
    else if (index == 0) b.adaptedStart = b.adaptedEnd - diffG + 1;
    else { b.adaptedStart = b.adaptedEnd - diffG / 2 + 1; b.adaptedEnd = b.adaptedStart + (diffG - diffG / 2) - 1; }
  }
  int ovS = std::max(b.readStart, b.adaptedStart), ovE = std::min(b.readEnd, b.adaptedEnd);
  b.startDiff = std::abs(ovS - b.adaptedStart); b.endDiff = std::abs(ovE - b.adaptedEnd);
  char exonStrand = '+';
  if (b.isGenomicRecord() || b.isJunction()) b.intervalStrand = b.alignStrand;
  else { exonStrand = b.exon.strand; b.intervalStrand = b.exon.strand == b.alignStrand ? '+' : '-'; }
  int od = b.overlap - 1;
  if (b.startDiff <= b.endDiff) {
    if (exonStrand == '+') { b.genomeStart = gs + b.startDiff; b.genomeEnd = gs + b.startDiff + od; }
    else { b.genomeStart = std::max(1, ge - b.startDiff - od); b.genomeEnd = ge - b.startDiff; }
  } else {
    if (exonStrand == '+') { b.genomeStart = std::max(1, ge - b.endDiff - od); b.genomeEnd = ge - b.endDiff; }
    else { b.genomeStart = gs + b.endDiff; b.genomeEnd = gs + b.endDiff + od; }
  }
  bool conf = is_conforming(b, p);
  return make_interval(b.exon.chr, b.genomeStart, b.genomeEnd, b.intervalStrand, b, ld, index, ga, conf, p);
}
// BedRecord.computeGenomicInterval(ga) :555-578 — genomic records
inline std::unique_ptr<GInterval> interval_genomic(BedRec& b, GAln* ga, const Params& p) {
  b.genomeStart = b.readStart; b.genomeEnd = b.readEnd; b.intervalStrand = b.alignStrand;
  bool conf = is_conforming(b, p);
  return make_interval(b.exon.chr, b.genomeStart, b.genomeEnd, b.alignStrand, b, 0, b.fieldIndex, ga, conf, p);
}

// GenomicInterval.compareTo :597-649
inline int gi_compare(const GInterval& a, const GInterval& b) {
  if (a.isTr() || b.isTr()) {
    if (a.rec->ref != b.rec->ref) throw Error("ERROR: tr-exon intervals of different references compared");
    if (a.rec->exon.strand != b.rec->exon.strand) throw Error("ERROR: tr-exon intervals with different exon strands compared");
    int c = (a.rec->exon.strand == '+' ? 1 : -1) * ((a.rec->eRefStart > b.rec->eRefStart) - (a.rec->eRefStart < b.rec->eRefStart));
    if (c == 0 && !(a.chr == b.chr && a.start == b.start && a.end == b.end)) throw Error("ERROR: equal transcript start for different genomic intervals");
    return c;
  }
  if (a.chr != b.chr) return a.chr.compare(b.chr);
  if (a.start != b.start) return a.start > b.start ? 1 : -1;
  if (a.end != b.end) return a.end > b.end ? 1 : -1;
  return (a.strand > b.strand) - (a.strand < b.strand);
}

// BedRecord.compareTo :351-387 (including its two slips: alignmentId vs the other's reference id; readAlignEnd vs itself)
inline int bedrec_compare(const BedRec& a, const BedRec& b) {
  if (a.alignmentId != b.alignmentId) return a.alignmentId.compare(b.ref);
  if (a.ref != b.ref) return a.ref.compare(b.ref);
  if (a.readIndex != b.readIndex) return a.readIndex > b.readIndex ? 1 : -1;
  if (a.readStart != b.readStart) return a.readStart > b.readStart ? 1 : -1;
  if (a.eRefStart != b.eRefStart) return a.eRefStart > b.eRefStart ? 1 : -1;
  if (a.eRefEnd != b.eRefEnd) return a.eRefEnd > b.eRefEnd ? 1 : -1;
  return (a.eRefStrand > b.eRefStrand) - (a.eRefStrand < b.eRefStrand);
}

struct GAln {                                          // J/bed-source/GenomicAlignment.java:92-454
  std::vector<std::unique_ptr<GInterval>> iv;          // TreeSet<GenomicInterval> order
  WOA* woa = nullptr; int start = INT32_MAX, maxFieldIndex = -1, readAlignStart = -1, ld = 0;
  bool genomic = false; JStr str;

  GInterval* findEqual(const GInterval& g) { for (auto& x : iv) if (gi_compare(*x, g) == 0) return x.get(); return nullptr; }
  void insert(std::unique_ptr<GInterval> g) {
    size_t k = 0; while (k < iv.size() && gi_compare(*iv[k], *g) < 0) k++;
    iv.insert(iv.begin() + k, std::move(g));
  }
  GAln(std::vector<BedRec*>& recs, WOA* w, const Params& p) : woa(w) {
    BedRec* old = nullptr; GInterval* oldG = nullptr; int i = -1, n = (int)recs.size();
    int sumL = 0, sumI = 0, sumD = 0; int I = -1, D = -1; char strand = 0;
    for (BedRec* bp : recs) {
      BedRec& b = *bp;
      if (!strand) strand = b.alignStrand;
      else if (strand != b.alignStrand) throw Error("BED record: " + b.readIdFull + " has a different strand than the old BED record");
      if (old && (genomic != b.isGenomicRecord() || (old->fieldIndex == b.fieldIndex && (I != b.I || D != b.D))))
        throw Error("Change of BED record type from " + old->readIdFull + " to " + b.readIdFull);
      genomic = b.isGenomicRecord(); D = b.D; I = b.I;
      if (readAlignStart == -1 || b.readStart < readAlignStart) readAlignStart = b.readStart;
      if (old && old->readAlignmentId != b.readAlignmentId) throw Error("Different read alignment ids: " + b.readAlignmentId + " vs " + old->readAlignmentId);
      GInterval* cur;
      if (genomic) {                                   // :166-218
        if (old && old->maxFieldIndex != b.maxFieldIndex) throw Error("Differing max field index for bed records of " + b.readIdFull);
        maxFieldIndex = b.maxFieldIndex;
        auto g = interval_genomic(b, this, p);
        GInterval* have = findEqual(*g);
        if (old && old->fieldIndex == b.fieldIndex && !(g->chr == oldG->chr && g->start == oldG->start && g->end == oldG->end))
          throw Error("Differing genomic intervals for " + old->readIdFull + " and " + b.readIdFull);
        if (!have) { cur = g.get(); insert(std::move(g)); } else { have->addExons(g->exonSet); cur = have; }
        // `oldGenomicInterval = genomicInterval` keeps the NEW object even when merged: coordinates are equal
      } else {                                         // :220-413
        if (i == -1) { maxFieldIndex = n - 1; b.setMaxFieldIndex(maxFieldIndex); }
        if (b.trExonAlignment && p.mode == MODE_JUNCTION) b.setFieldIndex(-1);
        else { i++; b.setFieldIndex(i); }
        if (sumL == 0) sumL = b.L; else if (sumL != b.L) throw Error("Aligned length: " + std::to_string(b.L) + " of BED record " + b.readIdFull + " differs from previous aligned length: " + std::to_string(sumL));
        if (sumI == 0) sumI = b.I; else if (sumI != b.I) throw Error("Number of insertions of BED record " + b.readIdFull + " differs from previous insertions");
        if (sumD == 0) sumD = b.D; else if (sumD != b.D) throw Error("Number of deletions of BED record " + b.readIdFull + " differs from previous deletions");
        if (old) {
          if (!(b.ref == old->ref && b.readStart == old->readStart && b.readEnd == old->readEnd)) throw Error("BedRecord: " + b.readIdFull + " -  changed alignment");
          if (b.exon.key() == old->exon.key()) throw Error("Unchanged exon: " + b.exon.key() + " in " + b.readIdFull);
        }
        auto g = interval_nongenomic(b, i, n, this, p);
        ld += g->ld;
        GInterval* have = findEqual(*g);
        if (!have) { cur = g.get(); insert(std::move(g)); } else { have->addExons(g->exonSet); cur = have; }
      }
      if (cur->start < start) start = cur->start;
      old = bp; oldG = cur;
    }
  }
  JStr buildString(const Params& p) const {            // setGenomicAlignmentString :896-912 (null-check variant)
    JStr r;
    for (auto& g : iv) { JStr t = g->str(p); if (r.isNull) r = t; else { r.s += ":" + t.s; r.lit = false; } }
    return r;
  }
  const JStr& alignmentString(const Params& p) { if (str.isNull) str = buildString(p); return str; }   // toGenomicAlignmentString :921-929
  bool isTrAlignment() const { for (auto& g : iv) if (!g->isTr()) return false; return true; }        // :794-810
};

}  // namespace eqp
