// ORACLE (test infrastructure) — weight-file merge, best-of-two SAM selection, SAM -> BED6 blocks.
// Follows J/util-source/CombineReadWeightFiles.java, J/sam-source/{CombineSamFiles,ConvertSamBed,
// SamProcessorBed,BedEntry}.java and SamRecord.computeBedEntries/printBedRecords.
#pragma once
#include "sam_side.hpp"

namespace eqp {

// ---- a6: CombineReadWeightFiles.main (:291-485) + FragmentEntryFactory (:36-113)
struct EntryFactory {
  Lines L; bool has; sv line; bool lineOk;
  bool cur = false; FragEntry entry; std::string id; bool set = false;     // fragmentEntry / fragmentEntryId
  explicit EntryFactory(const std::string* text) : L(text ? sv(*text) : sv()), has(text != nullptr) {
    lineOk = has && L.next(line);
    cur = has;                                                              // new FragmentEntry() placeholder (:54)
  }
  bool next(FragEntry& out) {                                               // getNextFragmentEntry() :60-73
    cur = false;
    if (lineOk) { entry = parse_frag_entry(line); cur = true; lineOk = L.next(line); }
    set = true;
    if (cur) out = entry;
    return cur;
  }
  // getNextFragmentEntry(query) :75-106 — first entry with id >= query id, NO equality check
  const FragEntry* seek(const FragEntry* q) {
    if (!q || !has) return nullptr;
    FragEntry tmp;
    while (cur && id.compare(q->name) < 0) { if (next(tmp)) id = entry.name; }
    return cur ? &entry : nullptr;
  }
};

inline std::string combine_read_weight_files(const std::string& w1, const std::string& w2, const std::string* aux,
                                             bool useEditDistance, double slack = 0.25) {
  EntryFactory f1(&w1), f2(&w2), fa(aux);
  FragEntry e1, e2; bool h1 = f1.next(e1), h2 = f2.next(e2);
  const FragEntry* ax = fa.seek(h1 ? &e1 : nullptr);
  std::string out;
  while (h1 || h2) {
    int comp;                                                               // getCompValue :141-168
    if (!h1) comp = 1; else if (!h2) comp = -1; else comp = e1.name.compare(e2.name);
    const FragEntry* r;
    if (comp < 0) r = &e1; else if (comp > 0) r = &e2;
    else {                                                                  // getRelevantFragmentEntry :173-248
      int cm = ax ? ax->compareMultiplicity(e2) : 0;
      if (cm == 1) r = &e1; else if (cm == -1) r = &e2;
      else if (!useEditDistance) r = e1.numAlignments() >= e2.numAlignments() ? &e1 : &e2;
      else {
        if (e1.editDistance() < 0 || e2.editDistance() < 0) throw Error("ERROR: No edit distance for " + e1.name);
        r = e1.compareEditDistance(e2, slack) >= 0 ? &e1 : &e2;
      }
    }
    out += r->print(); out += '\n';
    if (comp <= 0) {
      std::string old = e1.name; h1 = f1.next(e1); ax = fa.seek(h1 ? &e1 : nullptr);
      if (h1 && old.compare(e1.name) >= 0) throw Error("oldFramentId1: " + old + " is not smaller than the new fragmentId1: " + e1.name);
    }
    if (comp >= 0) {
      std::string old = e2.name; h2 = f2.next(e2);
      if (h2 && old.compare(e2.name) >= 0) throw Error("oldFramentId2: " + old + " is not smaller than the new fragmentId2: " + e2.name);
    }
  }
  return out;
}

// ---- a7: CombineSamFiles.main (:558-990)
struct SamGroupReader {                                                     // getSamRecords :104-199
  Lines L; bool started = false, pendingOk = false; SamRec pending;
  explicit SamGroupReader(sv text) : L(text) {}
  bool next(std::vector<SamRec>& out, const std::vector<SamRec>* old) {
    sv line;
    if (!started) {
      started = true;
      if (!L.next(line)) return false;
      while (starts_with(line, "@")) if (!L.next(line)) return false;
      pending = parse_sam(line); pendingOk = true;
    }
    if (!pendingOk) return false;
    out.clear();
    std::string name = pending.frag;
    while (pendingOk && pending.frag == name) {
      out.push_back(pending);
      if (L.next(line)) pending = parse_sam(line); else pendingOk = false;
    }
    if (old && !old->empty() && out[0].frag.compare((*old)[0].frag) < 0)
      throw Error("ERROR: SAM records out of order " + out[0].frag + " and " + (*old)[0].frag);
    return true;
  }
};
struct WeightReader {                                                       // getFragmentEntry :366-395
  Lines L; explicit WeightReader(sv t) : L(t) {}
  bool get(const std::string& name, FragEntry& fe) {
    sv line;
    if (!L.next(line)) return false;
    fe = parse_frag_entry(line);
    while (name.compare(fe.name) > 0) { if (!L.next(line)) break; fe = parse_frag_entry(line); }
    if (name != fe.name) throw Error("ERROR: No fragment entry corresponding to fragment " + name + " found - fragment name found: " + fe.name);
    return true;
  }
};

inline std::string combine_sam_files(sv sam1, sv sam2, sv wgt1, sv wgt2, double slack = 0.25) {
  std::string out;
  auto printAll = [&](const std::vector<SamRec>& rs) { for (auto& r : rs) { out.append(r.line); out += '\n'; } };
  auto printIdx = [&](const std::vector<SamRec>& rs, int idx) { for (auto& r : rs) if (r.readIndex == idx) { out.append(r.line); out += '\n'; } };
  auto isMapped = [](const std::vector<SamRec>& rs) { for (auto& r : rs) if (r.mapped()) return true; return false; };
  auto idxSet = [](const std::vector<SamRec>& rs) {                         // getReadIndexSet :404-432
    int m = 0; for (auto& r : rs) if (r.mapped()) { if (r.readIndex == 0) return 3; m |= r.readIndex == 1 ? 1 : r.readIndex == 2 ? 2 : 0; } return m; };
  auto printTwo = [&](const std::vector<SamRec>& a, const std::vector<SamRec>& b, const FragEntry& fa, const FragEntry& fb) {   // :290-354
    if (fa.hasMate) { printAll(a); return; }
    if (fb.hasMate) { printAll(b); return; }
    int s = idxSet(a);
    if (s == 3) { printAll(a); return; }
    if (s & 1) { printIdx(a, 1); printIdx(b, 2); return; }
    if (s & 2) { printIdx(b, 1); printIdx(a, 2); return; }
    throw Error("ERROR: unmapped SAM records for " + a[0].frag);
  };
  SamGroupReader r1(sam1), r2(sam2); WeightReader w1(wgt1), w2(wgt2);
  std::vector<SamRec> s1, s2, tmp; bool h1 = r1.next(s1, nullptr), h2 = r2.next(s2, nullptr);
  std::string name, oldName;
  while (h1 && h2) {
    const std::string &n1 = s1[0].frag, &n2 = s2[0].frag;
    name = n1.compare(n2) < 0 ? n1 : n2;
    if (oldName == name) throw Error("ERROR: Duplicate fragment name: old " + oldName + " vs new " + name);
    bool m1 = isMapped(s1) && name == n1, m2 = isMapped(s2) && name == n2;
    FragEntry fe1, fe2;
    if (!m1 && !m2) { if (name == n1) printAll(s1); else if (name == n2) printAll(s2); }
    else if (!m2) { w1.get(name, fe1); printAll(s1); }
    else if (!m1) { w2.get(name, fe2); printAll(s2); }
    else {
      bool g1 = w1.get(name, fe1), g2 = w2.get(name, fe2);
      if (!g2) printAll(s1);
      else if (!g1) printAll(s2);
      else {
        int cm = fe1.compareMultiplicity(fe2);
        if (cm == 1) printAll(s1);
        else if (cm == -1) printAll(s2);
        else {
          if (fe1.editDistance() < 0 || fe2.editDistance() < 0) throw Error("ERROR: No edit distance for " + name);
          if (fe1.compareEditDistance(fe2, slack) == 1) printTwo(s1, s2, fe1, fe2); else printTwo(s2, s1, fe2, fe1);
        }
      }
    }
    oldName = name;
    tmp = s1; h1 = r1.next(s1, &tmp);                                       // both advance (:839-843)
    tmp = s2; h2 = r2.next(s2, &tmp);
    while (h1 && h2 && s1[0].frag != s2[0].frag) {                          // realign (:861-912)
      oldName = name;
      FragEntry fe;
      if (s1[0].frag.compare(s2[0].frag) > 0) {
        name = s2[0].frag; if (isMapped(s2)) w2.get(name, fe); printAll(s2);
        tmp = s2; h2 = r2.next(s2, &tmp);
      } else {
        name = s1[0].frag; if (isMapped(s1)) w1.get(name, fe); printAll(s1);
        tmp = s1; h1 = r1.next(s1, &tmp);
      }
    }
  }
  // tail (:922-954): when file 2 still has groups, `samRecords = samRecords1` (null) drops them
  if (h1 && !h2) {
    while (h1) {
      name = s1[0].frag; FragEntry fe;
      if (isMapped(s1)) w1.get(name, fe);
      printAll(s1);
      tmp = s1; h1 = r1.next(s1, &tmp);
    }
  }
  return out;
}

}  // namespace eqp
