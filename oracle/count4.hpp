// ORACLE (test infrastructure) — a22-a26: ComputeCounts (J/bed-source/ComputeCounts.java).
#pragma once
#include "count3.hpp"

namespace eqp {

// GenomicInterval.updateExonSet :479-523 (note getOverlap(Exon) uses its own start twice, :462-464)
inline void update_exon_set(GInterval& g) {
  if (g.exonSet.size() <= 1) return;
  auto ov = [&](const Exon& e) { return std::max(0, std::min(e.end, g.end) - std::max(g.start, g.start) + 1); };
  int maxOv = 0;
  for (auto& e : g.exonSet) if (!e.isGenomic) maxOv = std::max(maxOv, ov(e));
  if (maxOv > 0) {
    g.overlap = std::max(g.overlap, maxOv);
    std::vector<Exon> keep;
    for (auto& e : g.exonSet) if (e.isGenomic || ov(e) >= maxOv) keep.push_back(e);
    g.exonSet.swap(keep);
  }
}

// GenomicAlignment.getConformingCountObjects :570-712
inline std::vector<std::string> conforming_features(GAln& ga, int threshold, FeatureTable& ft, bool consecutive, bool excludeAmbiguousIntervals, const Params& p) {
  if (p.mode == MODE_GENE || threshold < 0) threshold = ga.maxFieldIndex + 1;
  std::vector<std::string> order; std::unordered_map<std::string, std::vector<int>> bits;
  for (auto& gp : ga.iv) {
    GInterval& g = *gp;
    std::unordered_set<std::string> cur;
    update_exon_set(g);
    if (excludeAmbiguousIntervals && g.exonSet.size() >= 2) g.exonSet.clear();
    for (auto& e : g.exonSet) {
      auto it = ft.byExon.find(e.key());
      if (it == ft.byExon.end()) it = ft.byExon.emplace(e.key(), std::vector<std::string>{e.key()}).first;   // :622-626
      for (const std::string& f : it->second) {
        if (!cur.insert(f).second) continue;
        auto b = bits.find(f);
        if (b == bits.end()) { b = bits.emplace(f, std::vector<int>(std::max(0, ga.maxFieldIndex + 1), 0)).first; order.push_back(f); }
        if (g.fieldIndex >= 0) {
          if (g.fieldIndex >= (int)b->second.size()) throw Error("ERROR: field index out of range in alignment of " + ga.woa->fragmentId);
          b->second[g.fieldIndex] = 1;
        }
      }
    }
  }
  std::vector<std::string> out;
  for (const std::string& f : order) {
    const auto& a = bits[f]; int sum = 0, maxC = -1;
    for (int v : a) { if (consecutive && v == 0 && sum > 0) { maxC = std::max(maxC, sum); sum = 0; } sum += v; }
    if (std::max(sum, maxC) >= threshold) out.push_back(f);
  }
  return out;
}

struct CountStats { long numWeightObjects = 0, included = 0, excluded = 0; double sumObjectWeight = 0; };
struct CountResult {
  std::unordered_map<std::string, double> table; CountStats stats;
};

// ComputeCounts.processWeightObject :257-446
inline double process_weight_object(WeightObject& wo, double w, int minExonNum, bool consecutive, bool pairedEndOnly,
                                    std::unordered_map<std::string, double>& countTable, FeatureTable& ft, const Params& p) {
  bool genomicIntervalMode = p.mode == MODE_EXON;
  bool excludeIntervals = p.excludeAmbiguous && genomicIntervalMode;
  wo.adjustAndMerge(p);
  std::vector<std::string> order; std::unordered_map<std::string, double> t1, t2;
  auto add = [&](std::unordered_map<std::string, double>& t, const std::string& f) {   // addToTableEntry :100-111 (repeated addition)
    if (!t1.count(f) && !t2.count(f)) order.push_back(f);
    auto it = t.find(f);
    if (it == t.end()) t[f] = 0 + w; else it->second = it->second + w;
  };
  for (WOA* a : wo.merged) {
    std::vector<std::string> ids1, ids2;
    if (a->ga1) ids1 = conforming_features(*a->ga1, minExonNum, ft, consecutive, excludeIntervals, p);
    if (a->ga2) ids2 = conforming_features(*a->ga2, minExonNum, ft, consecutive, excludeIntervals, p);
    if (a->paired) {                                                        // :325-335
      if (pairedEndOnly) { std::vector<std::string> k; for (auto& f : ids1) if (std::find(ids2.begin(), ids2.end(), f) != ids2.end()) k.push_back(f); ids1.swap(k); }
      else for (auto& f : ids2) if (std::find(ids1.begin(), ids1.end(), f) == ids1.end()) ids1.push_back(f);
      ids2.clear();
    }
    for (auto& f : ids1) add(t1, f);
    for (auto& f : ids2) add(t2, f);
  }
  const std::vector<std::string>& ids = order;                              // union of both tables' keys
  if (p.excludeAmbiguous && !genomicIntervalMode && ids.size() >= 2) return 0;   // :371-382
  double sumAdded = 0;
  for (auto& f : ids) {                                                     // :411-434
    double w1 = t1.count(f) ? t1[f] : 0, w2 = t2.count(f) ? t2[f] : 0;
    double c = std::min(1.0, std::max(w1, w2));
    auto it = countTable.find(f);
    if (it == countTable.end()) countTable[f] = 0 + c; else it->second = it->second + c;
    sumAdded += c;
  }
  return sumAdded;
}

// lock-step weight lookup, ComputeCounts.getReadWeight :124-237 (cols 1-2 of final.wgt)
struct WeightScan {
  Lines L; explicit WeightScan(sv t) : L(t) {}
  double get(const std::string& fragmentId) {
    sv line;
    while (L.next(line)) {
      auto t = tokenize(line, '\t');
      if (t.size() < 2) throw Error("No weight for read " + std::string(t.empty() ? sv() : t[0]) + " found.");
      if (t[0] == fragmentId) return 1.0 / strtod(std::string(t[1]).c_str(), nullptr);
    }
    throw Error("Fragment id: " + fragmentId + " not found in read weight file.");
  }
};

// ComputeCounts.main loop :728-858
inline CountResult compute_counts(sv intersection, sv weights, FeatureTable& ft, const Params& p) {
  CountResult r; WeightScan ws(weights);
  int minExonNum = p.mode == MODE_JUNCTION ? 2 : 1; bool consecutive = p.mode == MODE_JUNCTION, pairedEndOnly = p.mode == MODE_GENE;
  for_each_weight_object(intersection, [&](WeightObject& wo) {
    r.stats.numWeightObjects++;
    int n = wo.numGenomicAlignments(p);
    double w = ws.get(wo.fragmentId);
    if (w > 1.0 / n) w = 1.0 / n;                                           // :762-764
    if (w >= p.readWeightThreshold) {
      double added = process_weight_object(wo, p.useReadWeights ? w : 1.0, minExonNum, consecutive, pairedEndOnly, r.table, ft, p);
      r.stats.sumObjectWeight += added;
      if (added > 0) r.stats.included++; else r.stats.excluded++;
    } else r.stats.excluded++;
  });
  return r;
}

// Output :875-950
inline std::string format_counts(const CountResult& r, const FeatureTable& ft, const Params& p) {
  std::string out;
  if (ft.columns == 2) {
    for (const std::string& id : ft.ids) {                                  // TreeSet<String> order
      auto it = r.table.find(id);
      if (it != r.table.end()) { if (p.outputZeroes || it->second != 0) out += id + "\t" + java_double(it->second) + "\n"; }
      else if (p.outputZeroes) out += id + "\t0\n";
    }
  } else {
    for (auto& row : ft.rows3) {
      auto it = r.table.find(row.second);
      if (it != r.table.end()) { if (p.outputZeroes || it->second != 0) out += row.first + "\t" + java_double(it->second) + "\n"; }
      else if (p.outputZeroes) out += row.first + "\t0.0\n";
    }
  }
  return out;
}

}  // namespace eqp
