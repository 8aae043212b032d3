// ORACLE (test infrastructure) — a12-a26: intersection-BED decode, transcript->genome mapping,
// alignment canonicalisation, multiplicity, conforming features, weighted accumulation.
// Follows J/bed-source/{BedRecord,Exon,GenomicInterval,GenomicAlignment,WeightObjectAlignment,
// WeightObject,ComputeReadWeights,ComputeCounts}.java and the mode flags of J/util-source/UtilLib.java:101-226.
#pragma once
#include <memory>
#include "common.hpp"

namespace eqp {

enum CountMode { MODE_NONE = 0, MODE_GENE = 1, MODE_EXON = 2, MODE_JUNCTION = 3 };

struct Params {                       // UtilLib statics + ComputeCounts locals, as one immutable struct
  CountMode mode = MODE_NONE;
  int overlapThreshold = 1;           // -O  (UtilLib.java:132)
  bool stranded = false;              // -s
  bool spliceConforming = true;       // -N clears (UtilLib.java:186)
  bool containment = false;           // -C
  int exonSlack = 0;                  // UtilLib.java:101-102
  bool useReadWeights = true;         // -u clears
  bool excludeAmbiguous = false;      // -U
  double readWeightThreshold = 0.01;  // -W
  bool outputZeroes = true;           // -n clears
  bool noAdjustShortcut = false;      // debug: run a17 even when it provably cannot change anything
};

inline bool is_tr(sv chr) { return starts_with(chr, "tr-"); }             // UtilLib.java:82-84
inline std::string strip_version(const std::string& s) {                  // Pattern "[.][0-9]+$"
  size_t dot = s.rfind('.');
  if (dot == std::string::npos || dot + 1 >= s.size()) return s;
  for (size_t i = dot + 1; i < s.size(); i++) if (s[i] < '0' || s[i] > '9') return s;
  return s.substr(0, dot);
}

// ---- Exon (J/bed-source/Exon.java:230-301, equality :366-376, toString :424-435)
struct Exon {
  std::string chr; int start = 0, end = 0; char strand = '+', ostrand = '+';
  bool isTrExon = false, isGenomic = false;
  std::string key() const { return chr + "/" + std::to_string(start) + "/" + std::to_string(end) + "/" + ostrand; }
  int length() const { return end - start + 1; }
};
inline Exon parse_exon(sv s, sv refId, bool haveRef) {
  auto t = tokenize(s, '/');
  Exon e;
  if (t.size() < 3) throw Error("Less than three fields are separated by / in exon: " + std::string(s));
  if (t.size() > 5) throw Error("Too many fields: " + std::string(s));
  size_t o = t.size() == 5 ? 1 : 0;
  e.chr = std::string(t[o]); e.start = to_int(t[o + 1]); e.end = to_int(t[o + 2]);
  e.strand = t.size() >= 4 ? (t[o + 3].empty() ? '+' : t[o + 3][0]) : '+';
  if (e.start == 0) e.start = 1;
  e.isTrExon = is_tr(e.chr);
  if (e.isTrExon) e.chr = strip_version(e.chr);
  e.ostrand = e.strand;
  if (haveRef) e.isGenomic = (e.chr == refId);
  return e;
}

// ---- Exon.loadCountObjectFile (:70-152): exon key -> feature ids; all feature ids sorted (TreeMap)
struct FeatureTable {
  std::unordered_map<std::string, std::vector<std::string>> byExon;
  std::set<std::string> ids;
  std::vector<std::pair<std::string, std::string>> rows3;                  // (printed id, internal id) in file order, 3-col maps
  int columns = 2;
  void load(sv text) {
    Lines L(text); sv line;
    while (L.next(line)) {
      auto t = tokenize(line, '\t');
      if (t.empty()) throw Error("No exon id found.");
      if (t[0] == "Exon Id") continue;
      if (t.size() < 2) throw Error("No count object for exon " + std::string(t[0]) + " found on loading.");
      std::string feat(t.size() >= 3 ? t[2] : t[1]);
      int fields = t.size() >= 3 ? 3 : 2;
      if (fields < columns) throw Error("Not all lines have " + std::to_string(columns) + " columns.");
      columns = std::max(columns, fields);
      if (fields == 3) rows3.push_back({std::string(t[1]), feat});
      ids.insert(feat);
      auto& v = byExon[parse_exon(t[0], "", false).key()];
      if (std::find(v.begin(), v.end(), feat) == v.end()) v.push_back(feat);
    }
  }
};

// ---- BedRecord (J/bed-source/BedRecord.java:118-343)
struct BedRec {
  std::string ref; int readStart = -1, readEnd = -1; std::string readIdFull; char alignStrand = 0;
  int eRefStart = -1, eRefEnd = -1; Exon exon; char eRefStrand = 0; int overlap = 0;
  int adaptedStart = -1, adaptedEnd = -1;
  std::string fragmentId, alignmentId, alignmentBaseId, readAlignmentId;
  int L = 0, I = 0, D = 0, fieldIndex = 0, maxFieldIndex = 0, readIndex = -1;
  bool paired = true, first = false, last = false, trExonAlignment = false;
  int genomeStart = 0, genomeEnd = 0, startDiff = -1, endDiff = -1; char intervalStrand = 0;

  bool isGenomicRecord() const { return ref == exon.chr; }                 // :932-934
  bool isJunction() const { return ref.find("junction") != std::string::npos; }
  void setFieldIndex(int v) { fieldIndex = v; if (v == 0) first = true; if (v == maxFieldIndex) last = true; }   // :894-909
  void setMaxFieldIndex(int v) { maxFieldIndex = v; if (fieldIndex == v) last = true; }                          // :911-922
};

inline BedRec parse_bed_record(sv line) {
  auto t = split(line, '\t');
  // Java's loop stops at end of string: a trailing empty field is not seen
  if (!t.empty() && t.back().empty() && !line.empty() && line.back() == '\t') t.pop_back();
  if (t.size() < 13) throw Error("ERROR: problem with bed record: " + std::string(line) + "  when reading field: " + std::to_string(t.size()) + " -> Too few BED entries: " + std::to_string(t.size()));
  BedRec b;
  b.ref = std::string(t[0]); b.readStart = to_int(t[1]) + 1; b.readEnd = to_int(t[2]); b.readIdFull = std::string(t[3]);
  b.alignStrand = t[5].empty() ? 0 : t[5][0];
  if (t[6] != t[0]) throw Error("Reference sequence id: " + b.ref + " does not equal: " + std::string(t[6]));
  b.eRefStart = to_int(t[7]) + 1; b.eRefEnd = to_int(t[8]);
  b.exon = parse_exon(t[9], b.ref, true); b.trExonAlignment = b.exon.isTrExon;
  b.eRefStrand = t[11].empty() ? 0 : t[11][0]; b.overlap = to_int(t[12]);
  int refDiff = b.eRefEnd - b.eRefStart + 1, genDiff = b.exon.end - b.exon.start + 1;
  if (b.eRefStart == 1 && refDiff < genDiff && !b.isGenomicRecord()) b.eRefStart -= genDiff - refDiff;   // :136-141
  b.adaptedStart = b.eRefStart; b.adaptedEnd = b.eRefEnd;
  if (ends_with(b.ref, ".inv-strand")) { b.ref.resize(b.ref.size() - 11); b.exon.strand = b.exon.strand == '+' ? '-' : '+'; }   // :154-161
  if (b.ref.find("junction") != std::string::npos) b.exon.strand = '+';                                  // :168-170
  if (!b.exon.isGenomic) b.ref = strip_version(b.ref);                                                     // :173-176
  const std::string& id = b.readIdFull;
  size_t a = id.rfind("-A");
  size_t l = a == std::string::npos ? id.find("-L") : id.find("-L", a);
  if (l != std::string::npos && l > 0 && id.find("-I", l) != std::string::npos && id.find("-D", l) != std::string::npos) {   // :188-217
    size_t i = id.find("-I", l), d = id.find("-D", i), f = id.find("-F", d), h = id.find("-", f + 1), s = id.find("/", h);
    if (a == std::string::npos || d == std::string::npos || f == std::string::npos || h == std::string::npos || s == std::string::npos)
      throw Error("ERROR: problem with bed record name: " + id);
    b.fragmentId = id.substr(0, a); b.alignmentId = id.substr(0, l);
    b.L = to_int(sv(id).substr(l + 2, i - l - 2)); b.I = to_int(sv(id).substr(i + 2, d - i - 2));
    b.D = to_int(sv(id).substr(d + 2, f - d - 2)); b.fieldIndex = to_int(sv(id).substr(f + 2, h - f - 2));
    b.maxFieldIndex = to_int(sv(id).substr(h + 1, s - h - 1));
    if (b.fieldIndex == 0) b.first = true;
    if (b.fieldIndex == b.maxFieldIndex) b.last = true;
  } else {                                                                                                 // :219-234
    std::string x = id;
    if (x.size() >= 3 && x[x.size() - 3] == '/' && (x[x.size() - 2] == 'S' || x[x.size() - 2] == 'P') && (x.back() == '1' || x.back() == '2')) {
      b.alignmentId = x.substr(0, x.size() - 3);
      size_t k = b.alignmentId.size();
      while (k > 0 && b.alignmentId[k - 1] >= '0' && b.alignmentId[k - 1] <= '9') k--;
      b.fragmentId = (k >= 2 && b.alignmentId.compare(k - 2, 2, "-A") == 0) ? b.alignmentId.substr(0, k - 2) : x;
    } else { b.alignmentId = x; b.fragmentId = x; }
    b.first = b.last = true;
  }
  if (id.size() < 2) throw Error("ERROR: problem with bed record name: " + id);
  if (id.back() == '1' || id.back() == '2') b.readIndex = id.back() - '0';
  bool sr = id[id.size() - 2] == 'S' && (id.back() == '1' || id.back() == '2');
  std::string tag = std::string(sr ? "/S" : "/P") + std::to_string(b.readIndex);
  b.readAlignmentId = b.alignmentId + tag;
  if (sr) { b.paired = false; b.alignmentBaseId = b.fragmentId + tag; } else b.alignmentBaseId = b.fragmentId;   // :246-258
  return b;
}

}  // namespace eqp
