// ORACLE (test infrastructure) — SAM side: decode, per-fragment statistics, best-set selection.
// Follows J/sam-source/{SamRecord,SamReader,FragmentEntry,SamProcessorFragmentEntry}.java.
#pragma once
#include "common.hpp"

namespace eqp {

// ---- SamRecord (J/sam-source/SamRecord.java:121-271)
struct SamRec {
  sv line;
  std::string qname, frag;      // originalQueryName / originalFragmentName
  int readIndex = -1, flag = 0, pos = -1, pnext = -1, seqLen = 0, nm = -1, hi = -1;
  sv rname, cigar, rnext;
  char strand = 0;              // '+', '-' or 0 (unmapped flag: strand stays "")

  int flagReadIndex() const {   // SamRecord.java:621-640
    int r = 1;
    if (flag % 2 == 1) { if (flag % 128 >= 64) r = 1; else if (flag % 256 >= 128) r = 2; }
    else if (ends_with(qname, "/1")) r = 1;
    else if (ends_with(qname, "/2")) r = 2;
    else if (readIndex != -1) r = readIndex;
    return r;
  }
  bool hasMate() const { return flag % 4 >= 2 && flag % 2 == 1; }           // :964-966
  bool mapped() const { return pos >= 0 && flag % 8 < 4; }                  // :980-982
  sv mateRef() const { return rnext == "=" ? rname : rnext; }              // :888-894
  bool isMate(const SamRec& s) const {                                      // :650-666
    if (hi >= 0) return s.hi == hi;
    return s.flagReadIndex() != flagReadIndex() && qname == s.qname && mateRef() == s.rname && pnext == s.pos &&
           rname == s.mateRef() && pos == s.pnext;
  }
};

inline SamRec parse_sam(sv line) {
  SamRec r; r.line = line;
  size_t p = 0; int i = 0; sv seq, qual; size_t tagStart = sv::npos;
  while (p < line.size()) {
    size_t e = line.find('\t', p);
    if (e == sv::npos) e = line.size();
    sv tok = line.substr(p, e - p);
    if (i == 0) {
      std::string q(tok);
      size_t sp = q.find(' ');
      if (sp != std::string::npos && sp > 0) {                              // :157-166
        q = q.substr(0, sp);
        if (ends_with(tok, "/1")) q += "/1";
        if (ends_with(tok, "/2")) q += "/2";
      }
      r.qname = q; r.frag = q;
      if (ends_with(r.frag, "/1")) { r.readIndex = 1; r.frag.resize(r.frag.size() - 2); }
      else if (ends_with(r.frag, "/2")) { r.readIndex = 2; r.frag.resize(r.frag.size() - 2); }
    } else if (i == 1) {
      r.flag = to_int(tok);
      int fri = r.flagReadIndex();
      if (r.readIndex == -1) r.readIndex = fri;
      else if (r.readIndex != fri) throw Error("Read index in flag: " + std::to_string(fri) + " differs from name read index:" + std::to_string(r.readIndex));
      if (r.flag % 8 < 4) r.strand = (r.flag % 32 >= 16) ? '-' : '+';        // :189-194
    } else if (i == 2) r.rname = tok;
    else if (i == 3) r.pos = tok == "*" ? -1 : to_int(tok);
    else if (i == 4) (void)to_int(tok);
    else if (i == 5) r.cigar = tok;
    else if (i == 6) r.rnext = tok;
    else if (i == 7) r.pnext = tok == "*" ? -1 : to_int(tok);
    else if (i == 9) seq = tok;
    else if (i == 10) qual = tok;
    else if (i >= 11) {
      if (tagStart == sv::npos) tagStart = p;
      if (starts_with(tok, "HI:i:")) r.hi = to_int(tok.substr(5));
    }
    if (i < 11 || tok.find(':') == sv::npos) i++;                           // :229-232 (tags do not advance i)
    p = e + 1;
  }
  if (i < 11) throw Error("Too few fields in SAM record (" + std::to_string(i) + ")");
  if (i > 12) throw Error("Too many fields in SAM record (" + std::to_string(i) + ")");
  if (qual != "*" && seq.size() != qual.size()) throw Error("sequence and qualities have different lengths");
  r.seqLen = (int)seq.size();
  // getEditDistance: first match of [nN]M:i:([0-9]+) in the joined tag string (:48,1052-1068)
  if (tagStart != sv::npos) {
    sv tags = line.substr(tagStart);
    for (size_t k = 0; k + 5 <= tags.size(); k++) {
      if ((tags[k] == 'N' || tags[k] == 'n') && tags.substr(k + 1, 4) == "M:i:") {
        size_t d = k + 5; int v = 0; bool any = false;
        while (d < tags.size() && tags[d] >= '0' && tags[d] <= '9') { v = v * 10 + (tags[d] - '0'); d++; any = true; }
        if (any) { r.nm = v; break; }
      }
    }
  }
  return r;
}

// ---- FragmentEntry (J/sam-source/FragmentEntry.java)
struct FragEntry {
  std::string name;
  int nPe = 0, nSr1 = 0, nSr2 = 0, edPe = 0, edSr1 = 0, edSr2 = 0;
  int readIndex = 1; bool hasMate = false;

  int nSr() const { return std::max(nSr1, nSr2); }
  int numAlignments() const { return nPe + nSr(); }                         // :595-609
  int sumEd() const { return edPe + edSr1 + edSr2; }
  double editDistance() const { return sumEd() * 1.0 / numAlignments(); }   // :591-593 (0/0 -> NaN)
  int multiplicity() const { return (hasMate || readIndex > 2) ? 2 : 1; }   // :677-679
  void setValues(int n, int ed, bool mate, int ri) {                        // :346-368
    if (mate) { nPe = n; edPe = ed; } else if (ri == 1) { nSr1 = n; edSr1 = ed; } else { nSr2 = n; edSr2 = ed; }
    hasMate = mate; readIndex = ri;
  }
  void addSingle(const SamRec& s) {                                         // :430-477 (readsMode false)
    name = s.frag; hasMate = false;
    if (s.readIndex != readIndex) readIndex = 3;
    if (s.readIndex == 1) { nSr1++; edSr1 += s.nm; } else { nSr2++; edSr2 += s.nm; }
  }
  void addPair(const SamRec& a, const SamRec& b) { name = a.frag; edPe += a.nm + b.nm; nPe++; hasMate = true; }   // :479-503
  std::string print() const {                                               // toPrintString :756-770
    std::string t = hasMate ? "paired-end" : "single-read";
    if (!hasMate && multiplicity() == 2) t += " (both)";
    return name + "\t" + std::to_string(numAlignments()) + "\t" + std::to_string(sumEd()) + "\t" + t;
  }
  int compareMultiplicity(const FragEntry& f) const {                       // :787-807
    if (hasMate) { if (!f.hasMate) return 1; } else if (f.hasMate) return -1;
    if (multiplicity() > 1) { if (f.multiplicity() <= 1) return 1; } else if (f.multiplicity() > 1) return -1;
    return 0;
  }
  int compareEditDistance(const FragEntry& f, double slack) const {         // :815-837
    int rv = compareMultiplicity(f);
    if (rv != 0) return rv;
    if (hasMate) slack = 2 * slack;
    if (f.editDistance() <= editDistance() && f.numAlignments() > 1.1 * numAlignments()) return -1;
    if (f.editDistance() < editDistance() - slack) return -1;
    return 1;
  }
};

inline FragEntry parse_frag_entry(sv line) {                                // parseFragmentEntryLine :176-262
  FragEntry f; auto t = split(line, '\t');
  std::string id(t[0]);
  if (ends_with(id, "/1/1") || ends_with(id, "/2/2")) id.resize(id.size() - 2);
  else if (ends_with(id, "/1/2") || ends_with(id, "/2/1")) throw Error("Wrong readId for read weight line: " + std::string(line));
  int ri = ends_with(id, "/2") ? 2 : 1;
  if (ends_with(id, "/1") || ends_with(id, "/2")) id.resize(id.size() - 2);
  f.name = id;
  if (t.size() < 2 || (t.size() == 2 && t[1].empty() && line.size() <= t[0].size() + 1)) throw Error("No weight for read " + id + " found.");
  int n = to_int(t[1]);
  int ed = 0; bool mate = false;
  if (t.size() >= 3 && !(t.size() == 3 && t[2].empty())) ed = to_int(t[2]);
  if (t.size() >= 4) { mate = t[3] == "paired-end"; if (t[3].find("both") != sv::npos) ri = 3; }
  f.setValues(n, ed, mate, ri);
  return f;
}

// ---- a1 + a3 + a4: SamReader.readSamFile driving SamProcessorFragmentEntry
// (J/sam-source/SamReader.java:139-301, SamProcessorFragmentEntry.java:100-217, ComputeReadWeightsSam.java:153-177)
template <class F>
inline void for_each_mapped_group(sv sam, F&& fn) {
  Lines L(sam); sv line; std::vector<SamRec> recs; std::string oldMapped;
  while (L.next(line)) {
    if (line.empty() || line[0] == '@') continue;
    SamRec r = parse_sam(line);
    if (!r.mapped()) continue;                                              // SamReader.java:197
    if (!oldMapped.empty() && r.frag != oldMapped) { fn(recs); recs.clear(); }
    oldMapped = r.frag;
    recs.push_back(std::move(r));
  }
  if (!recs.empty()) fn(recs);
}

// First-fit mate scan shared by a4 and a9 (SamRecord.findMate :582-608).
inline int find_mate(const std::vector<SamRec>& recs, size_t i, std::vector<char>& processed) {
  for (size_t j = i + 1; j < recs.size(); j++)
    if (!processed[j] && recs[i].isMate(recs[j])) { processed[j] = 1; return (int)j; }
  return -1;
}

inline std::string compute_read_weights_sam(sv sam, const std::unordered_set<std::string>* whitelist) {
  std::string out;
  for_each_mapped_group(sam, [&](std::vector<SamRec>& recs) {
    FragEntry fe; fe.name = "null";                      // FragmentEntry.fragmentName stays null until an add (:439-440)
    std::vector<char> processed(recs.size(), 0);
    for (size_t i = 0; i < recs.size(); i++) {
      if (processed[i]) continue;
      const SamRec& s = recs[i];
      bool ok = !whitelist || whitelist->count(std::string(s.rname));
      if (s.hasMate()) {
        int j = find_mate(recs, i, processed);
        if (j >= 0) { processed[i] = 1; if (ok) fe.addPair(s, recs[j]); }
        // else: silently dropped (SamProcessorFragmentEntry.java:192-197, warnings off)
      } else {
        processed[i] = 1;
        if (ok) fe.addSingle(s);
      }
    }
    out += fe.print(); out += '\n';
  });
  return out;
}

}  // namespace eqp
