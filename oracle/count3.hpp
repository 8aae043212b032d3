// ORACLE (test infrastructure) — WeightObjectAlignment / WeightObject (a15, a17-a20).
#pragma once
#include "count2.hpp"

namespace eqp {

struct WOA {                                           // J/bed-source/WeightObjectAlignment.java
  std::string alignmentId, baseId, fragmentId, referenceId; bool haveRef = false, paired = false, fromGenomic = false;
  std::vector<BedRec*> recs1, recs2;                   // TreeSet<BedRecord> by bedrec_compare
  std::unique_ptr<GAln> ga1, ga2; JStr str;

  void addBedRecord(BedRec* b) {                       // :107-155
    if ((!recs1.empty() || !recs2.empty()) && b->paired != paired) throw Error("BedRecord: " + b->readIdFull + " differs in paired-end status from previous records of the same alignment.");
    if (!haveRef) { referenceId = b->ref; haveRef = true; }
    else if (referenceId != b->ref) throw Error("BedRecord: " + b->readIdFull + " has a different reference sequence id than the weightObjectAlignment: " + referenceId);
    paired = b->paired;
    if (b->readIndex != 1 && b->readIndex != 2) throw Error("Unknown read index for bed record: " + b->readIdFull);
    auto& v = b->readIndex == 1 ? recs1 : recs2;
    size_t k = 0;
    for (; k < v.size(); k++) { int c = bedrec_compare(*b, *v[k]); if (c == 0) return; if (c < 0) break; }   // TreeSet: equal -> first kept
    v.insert(v.begin() + k, b);
  }
  void setString(const Params& p) {                    // setGenomicAlignmentString :185-219 (reference compares with != "")
    if (!ga1 && !ga2) return;
    JStr s1{"", false, true}, s2{"", false, true};
    if (ga1) s1 = ga1->alignmentString(p);
    if (ga2) s2 = ga2->alignmentString(p);
    std::string g;
    if (!s1.lit && !s2.lit) g = ga1->start <= ga2->start ? s1.s + "-" + s2.s : s2.s + "-" + s1.s;
    else if (!ga2) g = "l-" + s1.s;
    else if (!ga1) g = "r-" + s2.s;
    else g = s1.s + s2.s;
    str = {baseId + ":" + g, false, false};
  }
  void initialize(const Params& p) {                   // initializeGenomicAlignmentString :230-248
    if (!recs1.empty()) { ga1 = std::make_unique<GAln>(recs1, this, p); fromGenomic = ga1->genomic; }
    if (!recs2.empty()) { ga2 = std::make_unique<GAln>(recs2, this, p); fromGenomic = ga2->genomic; }
    setString(p);
  }
  std::string getString() const { return str.isNull ? alignmentId + ": No genomic alignment string" : str.s; }   // :166-174
  bool isTrAlignment() const { return (!ga1 || ga1->isTrAlignment()) && (!ga2 || ga2->isTrAlignment()); }        // :514-528
};

// test instrumentation: how often the shapes of SURVEY.md 8d were actually exercised (printed into stats.txt)
struct ShapeCounters { long snaps = 0, trRecords = 0, posLdRecords = 0; };
inline ShapeCounters& shape_counters() { static ShapeCounters c; return c; }

// GenomicInterval.adjustChrStartEnd(GenomicInterval) :284-358
inline bool adjust_pair(GInterval& x, GInterval& y, const Params& p) {
  if (x.chr != y.chr) return false;
  if (x.ga->woa->referenceId == y.ga->woa->referenceId) return false;
  if (x.start == y.start && x.end == y.end) return false;
  if (x.ld < y.ld || (x.ld == y.ld && x.start < y.start)) {
    if (adjust_pair(y, x, p)) { y.ga->str = y.ga->buildString(p); y.ga->woa->setString(p); return true; }
    return false;
  }
  if (std::abs(x.end - x.start) == std::abs(y.end - y.start) && std::abs(y.origStart - x.origStart) <= x.ld) {
    if (x.ld > y.ld || (x.ld == y.ld && x.start > y.start)) { x.start = y.start; x.end = y.end; x.ld = y.ld; shape_counters().snaps++; return true; }
  }
  return false;
}

struct WeightObject {                                  // J/bed-source/WeightObject.java
  std::string fragmentId;
  std::vector<std::unique_ptr<BedRec>> store;
  std::vector<std::unique_ptr<WOA>> alignments;        // weightObjectAlignments (Vector, file order)
  std::vector<WOA*> merged;                            // weightObjectAlignmentSet (first-seen order; the reference iterates a HashSet)
  bool done = false;

  void add(std::unique_ptr<BedRec> b) {                // addBedRecord :148-169 / addWeightObjectAlignment :110-133
    BedRec* bp = b.get(); store.push_back(std::move(b));
    if (alignments.empty() || !(bp->alignmentId == alignments.back()->alignmentId && bp->alignmentBaseId == alignments.back()->baseId)) {
      auto w = std::make_unique<WOA>(); w->alignmentId = bp->alignmentId; w->baseId = bp->alignmentBaseId; w->fragmentId = fragmentId;
      alignments.push_back(std::move(w));
    }
    alignments.back()->addBedRecord(bp);
  }

  void adjustAndMerge(const Params& p) {               // adjustGenomicAlignments :180-204 + mergeWeightAlignments :216-252
    if (done) return;                                  // second call sees an emptied Vector: no-op
    done = true;
    for (auto& w : alignments) w->initialize(p);
    bool anyPositive = p.noAdjustShortcut;
    for (auto& w : alignments) for (GAln* g : {w->ga1.get(), w->ga2.get()}) if (g) for (auto& x : g->iv) if (x->ld > 0) anyPositive = true;
    for (auto& w : alignments) {                       // WeightObjectAlignment.adjustGenomicAlignmentString :266-300
      int idx = 0;
      for (GAln* g : {w->ga1.get(), w->ga2.get()}) {
        idx++;
        if (!g) continue;
        // GenomicAlignment.adjustGenomicAlignmentString :873-887 (`== ""` variant).  The string is rebuilt in place, so
        // GenomicAlignment.equals() inside the sweep sees the partial string of the alignment being adjusted.
        JStr& s = g->str; s = JStr{"", false, true};
        for (auto& x : g->iv) {
          // GenomicInterval.adjustChrStartEnd(Vector, index) :367-421.  It can only move an interval whose lengthDifference
          // is > 0 (test at :330-334), so without any positive value in the fragment the O(n^2) sweep is skipped.
          if (x->ld != 0 && !x->adjusted) {
            x->adjusted = true;
            if (anyPositive)
              for (auto& w2 : alignments) {
                GAln* g2 = idx == 1 ? w2->ga1.get() : w2->ga2.get();
                if (g2 && g->alignmentString(p).s != g2->alignmentString(p).s) for (auto& y : g2->iv) adjust_pair(*x, *y, p);
              }
          }
          JStr t = x->str(p);
          if (s.lit) s = t; else { s.s += ":" + t.s; s.lit = false; }
        }
      }
      w->setString(p);
    }
    std::unordered_map<std::string, WOA*> table;
    for (auto& w : alignments) {
      std::string key = w->getString();
      if (key.empty()) continue;
      auto it = table.find(key);
      if (it == table.end()) { table[key] = w.get(); merged.push_back(w.get()); }
      else combineExons(*it->second, *w, p);
    }
  }

  static bool gaEq(GAln* a, GAln* b, const Params& p) { return a->alignmentString(p).s == b->alignmentString(p).s; }
  static void combineGa(GAln& a, GAln& b, const Params& p) {                // GenomicAlignment.combineExons :721-786
    size_t j = 0; GInterval* other = nullptr;
    for (auto& x : a.iv) {
      if (x->str(p).s.empty()) continue;
      if (j < b.iv.size()) { other = b.iv[j++].get(); while (j < b.iv.size() && other->str(p).s.empty()) other = b.iv[j++].get(); }
      else throw Error("ERROR: too few genomic intervals when merging alignments of " + a.woa->fragmentId);
      if (x->str(p).s != other->str(p).s) throw Error("ERROR: genomic interval mismatch when merging alignments of " + a.woa->fragmentId);
      x->addExons(other->exonSet);
    }
  }
  static void combineExons(WOA& a, WOA& b, const Params& p) {               // WeightObjectAlignment.combineExons :312-374
    GAln *a1 = a.ga1.get(), *a2 = a.ga2.get(), *b1 = b.ga1.get(), *b2 = b.ga2.get();
    if (a1 && a2 && b1 && b2) {
      if (gaEq(a1, b1, p) && gaEq(a2, b2, p)) { combineGa(*a1, *b1, p); combineGa(*a2, *b2, p); return; }
      if (gaEq(a1, b2, p) && gaEq(a2, b1, p)) { combineGa(*a1, *b2, p); combineGa(*a2, *b1, p); return; }
    }
    if (a1 && b1 && gaEq(a1, b1, p)) { combineGa(*a1, *b1, p); return; }
    if (a2 && b2 && gaEq(a2, b2, p)) { combineGa(*a2, *b2, p); return; }
    if (a1 && b2 && gaEq(a1, b2, p)) { combineGa(*a1, *b2, p); return; }
    if (a2 && b1 && gaEq(a2, b1, p)) { combineGa(*a2, *b1, p); return; }
    throw Error("Inconsistent weightObjectAlignments when combining the exons of " + a.alignmentId + " with: " + b.alignmentId);
  }

  int numGenomicAlignments(const Params& p) {          // getNumGenomicAlignments :819-880
    adjustAndMerge(p);
    int left = 0, right = 0; bool hasTr = false;
    for (WOA* w : merged) {
      if (!w->isTrAlignment()) { if (w->getString().find(":r-") != std::string::npos && w->getString().find(":r-") > 0) right++; else left++; }
      else hasTr = true;
    }
    return std::max(left, right) + (hasTr ? 1 : 0);
  }
};

// ---- a21: ComputeReadWeights.main (:59-198) — mode "none", output "Fid\t<a20>"
template <class F>
inline void for_each_weight_object(sv intersection, F&& fn) {               // shared fragment grouping (ComputeCounts.java:728-858)
  Lines L(intersection); sv line; std::unique_ptr<WeightObject> wo;
  while (L.next(line)) {
    auto b = std::make_unique<BedRec>(parse_bed_record(line));
    if (b->overlap <= 0) continue;
    if (!wo || b->fragmentId != wo->fragmentId) {
      if (wo) fn(*wo);
      wo = std::make_unique<WeightObject>(); wo->fragmentId = b->fragmentId;
    }
    wo->add(std::move(b));
  }
  if (wo) fn(*wo);
}
inline std::string compute_read_weights(sv intersection) {
  Params p; p.mode = MODE_NONE; std::string out;
  for_each_weight_object(intersection, [&](WeightObject& wo) { out += wo.fragmentId + "\t" + std::to_string(wo.numGenomicAlignments(p)) + "\n"; });
  return out;
}

}  // namespace eqp
