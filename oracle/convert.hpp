// ORACLE (test infrastructure) — a8-a11: SAM -> BED6 blocks and the bedtools join.
#pragma once
#include "combine.hpp"

namespace eqp {

// ---- a8: ConvertSamBed.loadBedFile (J/sam-source/ConvertSamBed.java:47-98)
// base transcript id -> versioned ids, in first-seen order (the reference iterates a HashSet: SURVEY H10)
using VersionTable = std::unordered_map<std::string, std::vector<std::string>>;
inline VersionTable load_transcript_versions(sv transcriptBed) {
  VersionTable t; Lines L(transcriptBed); sv line;
  while (L.next(line)) {
    auto tok = tokenize(line, '\t');
    if (tok.empty()) continue;
    std::string ver(tok[0]), base = ver;
    if (ends_with(base, ".inv-strand")) base.resize(base.size() - 11);
    size_t dot = base.rfind('.');
    if (dot != std::string::npos && base.size() - dot <= 4) base.resize(dot);
    auto& v = t[base];
    if (std::find(v.begin(), v.end(), ver) == v.end()) v.push_back(ver);
  }
  return t;
}

// ---- a10: SamRecord.computeBedEntries (J/sam-source/SamRecord.java:316-475) + BedEntry (:76-88)
struct BedBlock { int start, end, score; std::string extra; char strand; };   // extra = "-L..-I..-D..-F..-.."
// strandDir: 0 = none, 1 = `-S forward` (getFragmentStrand, SamRecord.java:859-865), 2 = `-S backward` (getReverseFragmentStrand, :867-873)
inline std::vector<BedBlock> compute_bed_entries(const SamRec& r, int& bedEntryNum, const std::string& fragForErr, int strandDir = 0) {
  std::vector<BedBlock> out;
  char strand = r.strand;                                                   // :431-436
  if (strandDir && r.readIndex != (strandDir == 1 ? 1 : 2)) strand = strand == '+' ? '-' : '+';
  int bedStart = std::max(0, r.pos - 1);
  // split at [0-9]+N; then [0-9]+[A-Z=] tokens inside each segment
  struct Seg { int m = 0, ins = 0, del = 0, soft = 0, hard = 0; int intron = 0; };
  std::vector<Seg> segs(1);
  { long num = -1;
    for (char c : r.cigar) {
      if (c >= '0' && c <= '9') { num = (num < 0 ? 0 : num) * 10 + (c - '0'); continue; }
      if (num >= 0 && ((c >= 'A' && c <= 'Z') || c == '=')) {
        Seg& s = segs.back();
        if (c == 'N') { s.intron = (int)num; segs.emplace_back(); }
        else if (c == 'M' || c == 'X' || c == '=') s.m += (int)num;
        else if (c == 'I') s.ins += (int)num;
        else if (c == 'D') s.del += (int)num;
        else if (c == 'S') s.soft += (int)num;
        else if (c == 'H') s.hard += (int)num;
        else if (c != 'P') throw Error("ERROR: Unknown operation in CIAGR string for fragment " + fragForErr);
      }
      num = -1;
    }
  }
  int nSeg = (int)segs.size(), sumRef = 0, sumRead = 0, sumIntron = 0;
  for (int k = 0; k < nSeg; k++) {
    const Seg& s = segs[k];
    int start = bedStart + sumRef + sumIntron;
    int refLen = s.m + s.del, readLen = s.m + s.ins + s.soft;
    if (readLen > 0) {
      char buf[96];
      snprintf(buf, sizeof buf, "-L%d-I%d-D%d-F%d-%d", readLen, s.ins + s.soft, s.del, k, nSeg - 1);
      out.push_back({start, start + refLen, bedEntryNum, buf, strand});
      bedEntryNum++;
    }
    sumRef += refLen; sumRead += readLen; sumIntron += s.intron;
  }
  if (sumRead != r.seqLen)
    throw Error("Fragment " + fragForErr + " mapped to " + std::string(r.rname) + ": read length " + std::to_string(r.seqLen) +
                " differs from sum of matches, insertions, and soft clipped bases " + std::to_string(sumRead));
  return out;
}

// ---- a9: ConvertSamBed.main + SamProcessorBed.processSamRecords (:152-204) + SamRecord.printBedRecords (:484-573)
struct ConvertStats { long mappedFragments = 0; };
inline std::string convert_sam_bed(sv sam, const VersionTable* versions, ConvertStats* st = nullptr, int strandDir = 0) {
  std::string out; int bedEntryNum = 0;                                     // static SamRecord.bedEntryNum (:41)
  for_each_mapped_group(sam, [&](std::vector<SamRec>& recs) {
    if (st) st->mappedFragments++;
    const std::string fragId = recs[0].frag;
    int counter = 1;                                                        // new Counter(5) per fragment (SamProcessorBed.java:132-143)
    std::vector<char> processed(recs.size(), 0);
    for (size_t i = 0; i < recs.size(); i++) {
      if (processed[i]) continue;
      processed[i] = 1;
      const SamRec& s = recs[i];
      const SamRec* mate = nullptr;
      if (s.hasMate()) { int j = find_mate(recs, i, processed); if (j >= 0) mate = &recs[j]; }
      auto name = [&](const SamRec& x) { return fragId + (x.hasMate() ? "/P" : "/S") + std::to_string(x.flagReadIndex()); };
      auto be = compute_bed_entries(s, bedEntryNum, name(s), strandDir);
      std::vector<BedBlock> mbe;
      if (mate) mbe = compute_bed_entries(*mate, bedEntryNum, name(*mate), strandDir);
      std::vector<std::string> one{std::string(s.rname)};
      const std::vector<std::string>* ids = &one;
      if (versions) { auto it = versions->find(std::string(s.rname)); if (it != versions->end()) ids = &it->second; }
      if (mate && !(s.rname == mate->mateRef() && s.mateRef() == mate->rname))
        throw Error("Differing referenceName mate found for: " + std::string(s.line));
      for (const std::string& ref : *ids) {
        std::string aid = "-A" + zero_filled(counter, 5);
        auto put = [&](const BedBlock& b, const SamRec& x) {
          std::string nm = name(x);                                         // BedEntry.toString(alignmentId, ref) :198-212
          std::string id = nm.substr(0, nm.size() - 3) + aid + b.extra + nm.substr(nm.size() - 3);
          out += ref; out += '\t'; out += std::to_string(b.start); out += '\t'; out += std::to_string(b.end); out += '\t';
          out += id; out += '\t'; out += std::to_string(b.score); out += '\t'; if (b.strand) out += b.strand; out += '\n';
        };
        for (auto& b : be) put(b, s);
        if (mate) for (auto& b : mbe) put(b, *mate);
        counter++;
      }
      // counter.dec() after the id loop and counter.inc() per record cancel (SamRecord.java:565, SamProcessorBed.java:190)
    }
  });
  return out;
}

// ---- a11: bedtools intersect -wo [-s] -a stdin -b B | awk '$1==$7'  (CIB:938-940).
// bedtools (arq5x/bedtools2 >= 2.24.0, README.md:74) is not vendored; restated from its documented semantics:
// 0-based half-open, every A x B pair on the same chrom string with >= 1 bp overlap, A order kept,
// overlap = min(end) - max(start); -s = same strand.  B hits are emitted in B-file order.
struct BRow { int start, end; size_t fileIdx; sv line; char strand; };
struct BIndex {
  std::unordered_map<std::string, std::vector<BRow>> byChrom;                // sorted by start
  std::unordered_map<std::string, std::vector<int>> maxEnd;                  // running max of end
  explicit BIndex(sv bed) {
    Lines L(bed); sv line; size_t idx = 0;
    while (L.next(line)) {
      if (line.empty()) continue;
      auto t = split(line, '\t');
      if (t.size() < 6) throw Error("Too few BED entries in B file");
      byChrom[std::string(t[0])].push_back({to_int(t[1]), to_int(t[2]), idx++, line, t[5].empty() ? '.' : t[5][0]});
    }
    for (auto& kv : byChrom) {
      std::stable_sort(kv.second.begin(), kv.second.end(), [](const BRow& a, const BRow& b) { return a.start < b.start; });
      auto& me = maxEnd[kv.first]; int m = 0;
      for (auto& r : kv.second) { m = std::max(m, r.end); me.push_back(m); }
    }
  }
};
inline std::string intersect_wo(sv aBed, const BIndex& B, bool sameStrand) {
  std::string out; Lines L(aBed); sv line; std::vector<const BRow*> hits;
  while (L.next(line)) {
    auto t = split(line, '\t');
    auto it = B.byChrom.find(std::string(t[0]));
    if (it == B.byChrom.end()) continue;
    int as = to_int(t[1]), ae = to_int(t[2]); char astr = t.size() > 5 && !t[5].empty() ? t[5][0] : '.';
    const auto& rows = it->second; const auto& me = B.maxEnd.at(it->first);
    // rows with start < ae; walk back while the running max end can still exceed as
    size_t hi = std::lower_bound(rows.begin(), rows.end(), ae, [](const BRow& r, int v) { return r.start < v; }) - rows.begin();
    hits.clear();
    for (size_t k = hi; k-- > 0;) {
      if (me[k] <= as) break;
      if (rows[k].end > as && (!sameStrand || rows[k].strand == astr)) hits.push_back(&rows[k]);
    }
    std::sort(hits.begin(), hits.end(), [](const BRow* a, const BRow* b) { return a->fileIdx < b->fileIdx; });
    for (const BRow* h : hits) {
      int ov = std::min(ae, h->end) - std::max(as, h->start);
      if (ov < 1) continue;
      out.append(line); out += '\t'; out.append(h->line); out += '\t'; out += std::to_string(ov); out += '\n';
    }
  }
  return out;
}

}  // namespace eqp
