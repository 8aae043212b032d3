// Host-side annotation loader: parses the files compute-counts.sh reads and builds the device index
// (reference-name hash table, bucketed sorted interval index, exon -> feature CSR lists, a27 activation lists).
// Mirrors: ConvertSamBed.loadBedFile (J/sam-source/ConvertSamBed.java:47-98), Exon.loadCountObjectFile
// (J/bed-source/Exon.java:70-152), Exon(String,String) (:230-301), ComputeReadWeightsSam.readChromosomeNames (:44-58),
// filter-junction-map-files.sh:254-304 + createExtendedJunctionIds.py:93-114 + filterFile.py:131-194.
#pragma once
#include <zlib.h>
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <map>
#include <set>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>
#include "eqp_types.h"

namespace eqpb {

// Reference-name hash, defined over little-endian 4-byte groups of the name (the last group zero-padded) so that the device
// can feed it one aligned shared-memory word pair per step instead of one byte.  Two 32-bit lanes give the 64-bit key.
struct NameHash {
  uint32_t a = 0x811C9DC5u, b = 0x9E3779B9u;
  EQP_HD void add(uint32_t w) {
    uint32_t k = w * 0xCC9E2D51u; k = (k << 15) | (k >> 17); k *= 0x1B873593u;
    a ^= k; a = (a << 13) | (a >> 19); a = a * 5u + 0xE6546B64u;
    b = (b ^ w) * 0x85EBCA6Bu; b ^= b >> 15;
  }
  EQP_HD uint64_t finish(int n) {
    uint32_t x = a ^ (uint32_t)n; x ^= x >> 16; x *= 0x85EBCA6Bu; x ^= x >> 13; x *= 0xC2B2AE35u; x ^= x >> 16;
    uint32_t y = b + x; y ^= y >> 16; y *= 0x7FEB352Du; y ^= y >> 15; y *= 0x846CA68Bu; y ^= y >> 16;
    uint64_t h = ((uint64_t)x << 32) | y;
    return h ? h : 1;                                 // 0 marks an empty slot
  }
};
EQP_HD uint32_t name_group_bytes(const char* p, int nvalid) {     // bytes p[0..min(nvalid,4)) as a little-endian word, zero-padded
  uint32_t w = 0;
  for (int i = 0; i < 4; i++) if (i < nvalid) w |= (uint32_t)(uint8_t)p[i] << (8 * i);
  return w;
}
EQP_HD uint64_t name_hash(const char* p, int n) {
  NameHash H; for (int i = 0; i < n; i += 4) H.add(name_group_bytes(p + i, n - i));
  return H.finish(n);
}

EQP_HD uint64_t dict_key(uint64_t h, int len) { uint64_t k = (h & ~0xFFFFULL) | (uint64_t)(len & 0xFFFF); return k ? k : 0x10000ULL; }

struct AnnoError : std::runtime_error { int code; AnnoError(int c, const std::string& m) : std::runtime_error(m), code(c) {} };

struct AnnoHost {
  std::vector<DictEntry> dict; uint32_t dictMask = 0;
  std::string strPool; std::vector<uint32_t> refStrOff; std::vector<uint16_t> refStrLen; std::vector<uint8_t> refFlags;
  std::vector<uint32_t> refVerOff; std::vector<int32_t> refVerBucket;
  std::vector<uint32_t> refActOff; std::vector<int32_t> refAct;
  std::vector<uint32_t> bktRowOff; std::vector<uint8_t> bktFlags; std::vector<int32_t> bktRefId;
  std::vector<int32_t> rowStart, rowEnd, rowMaxEnd, rowExon; std::vector<uint8_t> rowFlags;
  std::vector<int32_t> exChr, exStart, exEnd; std::vector<uint8_t> exStrand;
  std::vector<RowA> rowA; std::vector<ExonA> exA; std::vector<ExonFeat> exF;                        // the device form of the row / exon tables
  std::vector<uint32_t> featOff[3], feat[3]; std::vector<std::string> featIds[3];
  std::vector<uint32_t> featAct; std::vector<std::string> actIds; int mapColumns[3] = {2, 2, 2};
  // rows of the count file in print order (ComputeCounts.java:875-950): 2-column maps print every id of the map in TreeSet order,
  // 3-column maps print one row per map line (file order; the junction map in the order filter-junction-map-files.sh:298-299 leaves it in)
  struct OutRow { std::string id; int32_t feature; int32_t act; uint8_t always; };
  std::vector<OutRow> outRows[3];
  std::vector<std::string> chrNames, trChrNames, bucketNames;

  static std::string slurp(const char* path) {
    if (!path) throw AnnoError(-1, "annotation path is null");
    gzFile f = gzopen(path, "rb");                    // plain or gzip (UtilLib.getBufferedReader: ".gz" => GZIPInputStream, UtilLib.java:322-383)
    if (!f) throw AnnoError(-3, std::string("Warning: file ") + path + " not found.");
    gzbuffer(f, 1 << 20);
    std::string s; std::vector<char> buf(1 << 22); int n;
    while ((n = gzread(f, buf.data(), (unsigned)buf.size())) > 0) s.append(buf.data(), (size_t)n);
    bool bad = n < 0; gzclose(f);
    if (bad) throw AnnoError(-3, std::string("Problem reading ") + path);
    return s;
  }
  template <class F> static void lines(const std::string& s, F&& fn) {
    size_t p = 0;
    while (p < s.size()) {
      size_t e = s.find('\n', p); if (e == std::string::npos) e = s.size();
      size_t l = e; if (l > p && s[l - 1] == '\r') l--;
      fn(s.substr(p, l - p)); p = e + 1;
    }
  }
  static std::vector<std::string> splitc(const std::string& s, char c) {
    std::vector<std::string> o; size_t p = 0;
    while (true) { size_t e = s.find(c, p); if (e == std::string::npos) { o.push_back(s.substr(p)); break; } o.push_back(s.substr(p, e - p)); p = e + 1; }
    return o;
  }
  static std::vector<std::string> tok(const std::string& s, char c) {   // StringTokenizer
    std::vector<std::string> o; for (auto& t : splitc(s, c)) if (!t.empty()) o.push_back(t); return o;
  }
  static std::string stripVersion(const std::string& s) {               // Pattern "[.][0-9]+$"
    size_t d = s.rfind('.'); if (d == std::string::npos || d + 1 >= s.size()) return s;
    for (size_t i = d + 1; i < s.size(); i++) if (s[i] < '0' || s[i] > '9') return s;
    return s.substr(0, d);
  }
  struct ExonKey { std::string chr; int s, e; char strand; bool tr; };
  static ExonKey parseExon(const std::string& id) {                     // Exon.java:230-301
    auto t = tok(id, '/');
    if (t.size() < 3 || t.size() > 5) throw AnnoError(-3, "Less than three fields are separated by / in exon: " + id);
    size_t o = t.size() == 5 ? 1 : 0; ExonKey k;
    k.chr = t[o]; k.s = atoi(t[o + 1].c_str()); k.e = atoi(t[o + 2].c_str()); k.strand = t.size() >= 4 ? t[o + 3][0] : '+';
    if (k.s == 0) k.s = 1;
    k.tr = k.chr.compare(0, 3, "tr-") == 0;
    if (k.tr) k.chr = stripVersion(k.chr);
    return k;
  }
  static std::string keyStr(const ExonKey& k) { return k.chr + "/" + std::to_string(k.s) + "/" + std::to_string(k.e) + "/" + k.strand; }

  int32_t addRef(std::unordered_map<std::string, int32_t>& idx, const std::string& name) {
    auto it = idx.find(name); if (it != idx.end()) return it->second;
    int32_t r = (int32_t)refStrOff.size(); idx[name] = r;
    refStrOff.push_back((uint32_t)strPool.size()); refStrLen.push_back((uint16_t)name.size()); strPool += name; while (strPool.size() & 3) strPool.push_back('\0'); refFlags.push_back(0);   // 4-byte aligned, zero-padded: lookup_ref compares words
    return r;
  }

  void load(const char* combinedBed, const char* transcriptBed, const char* geneMap, const char* exonMap, const char* junctionMap,
            const char* junctionTranscriptMap, const char* jjCoordinateMap, const char* equivalentsMap, const char* fai) {
    // ---- B file: buckets, rows, distinct exons.  Rows equal in (start,end,strand) keep file order, so the device-side
    // first-wins dedupe (TreeSet<BedRecord>, BedRecord.java:351-387) can pick the lowest row index.
    std::unordered_map<std::string, int32_t> bucketIdx, exonIdx, chrIdxTmp;
    struct Row { int s, e, exon; char strand; size_t file; };
    std::vector<std::vector<Row>> rows; std::vector<ExonKey> exons;
    { std::string b = slurp(combinedBed); size_t fileIdx = 0;
      lines(b, [&](const std::string& l) {
        if (l.empty()) return;
        auto t = splitc(l, '\t'); if (t.size() < 6) throw AnnoError(-3, "Too few BED entries in " + std::string(combinedBed));
        auto bi = bucketIdx.find(t[0]);
        if (bi == bucketIdx.end()) { bi = bucketIdx.emplace(t[0], (int32_t)rows.size()).first; rows.emplace_back(); bucketNames.push_back(t[0]); }
        ExonKey k = parseExon(t[3]);
        std::string ks = keyStr(k);
        auto ei = exonIdx.find(ks);
        if (ei == exonIdx.end()) { ei = exonIdx.emplace(ks, (int32_t)exons.size()).first; exons.push_back(k); }
        rows[bi->second].push_back({atoi(t[1].c_str()), atoi(t[2].c_str()), ei->second, t[5].empty() ? '.' : t[5][0], fileIdx++});
      }); }
    { std::set<std::string> chrs, trs; for (auto& k : exons) (k.tr ? trs : chrs).insert(k.chr);
      for (auto& c : chrs) { chrIdxTmp[c] = (int32_t)chrNames.size(); chrNames.push_back(c); }
      for (auto& c : trs) { chrIdxTmp[c] = -2 - (int32_t)trChrNames.size(); trChrNames.push_back(c); } }
    for (auto& k : exons) { exChr.push_back(chrIdxTmp[k.chr]); exStart.push_back(k.s); exEnd.push_back(k.e); exStrand.push_back((uint8_t)k.strand); }
    std::unordered_map<std::string, int32_t> refIdTmp;
    bktRowOff.push_back(0);
    for (size_t b = 0; b < rows.size(); b++) {
      auto& v = rows[b];
      std::stable_sort(v.begin(), v.end(), [](const Row& x, const Row& y) { return x.s != y.s ? x.s < y.s : x.e != y.e ? x.e < y.e : x.strand != y.strand ? x.strand < y.strand : x.file < y.file; });
      const std::string& name = bucketNames[b];
      bool inv = name.size() > 11 && name.compare(name.size() - 11, 11, ".inv-strand") == 0;
      bktFlags.push_back((name.find("junction") != std::string::npos ? 1 : 0) | (inv ? 2 : 0));
      { bool genomicBucket = false; for (auto& x : v) genomicBucket |= !exons[x.exon].tr && exons[x.exon].chr == name;   // BedRecord.java:154-176
        std::string rid = inv ? name.substr(0, name.size() - 11) : name; if (!genomicBucket) rid = stripVersion(rid);
        bktRefId.push_back(refIdTmp.emplace(rid, (int32_t)refIdTmp.size()).first->second); }
      int maxEnd = 0;
      for (size_t i = 0; i < v.size(); i++) {
        maxEnd = std::max(maxEnd, v[i].e);
        rowStart.push_back(v[i].s); rowEnd.push_back(v[i].e); rowMaxEnd.push_back(maxEnd); rowExon.push_back(v[i].exon);
        rowFlags.push_back((!exons[v[i].exon].tr && exons[v[i].exon].chr == name ? 1 : 0) | (v[i].strand == '+' ? 0 : v[i].strand == '-' ? 2 : 4));
      }
      bktRowOff.push_back((uint32_t)rowStart.size());
    }
    rowA.resize(rowStart.size()); exA.resize(exChr.size());
    for (size_t i = 0; i < rowA.size(); i++) rowA[i] = RowA{rowStart[i], rowEnd[i], rowMaxEnd[i], ((uint32_t)rowExon[i] << 3) | (uint32_t)(rowFlags[i] & 7)};
    for (size_t i = 0; i < exA.size(); i++) exA[i] = ExonA{exChr[i], exStart[i], exEnd[i], (uint32_t)exStrand[i]};
    if (exA.size() >= (1u << 29)) throw AnnoError(-3, "too many distinct exons");
    // ---- reference-name dictionary: bucket names, base ids of transcript_exons.bed, genome.fa.fai names
    std::unordered_map<std::string, int32_t> refIdx;
    std::map<std::string, std::vector<std::string>> versions;           // base id -> versioned ids, first-seen order
    { std::string t = slurp(transcriptBed);
      lines(t, [&](const std::string& l) {
        auto f = tok(l, '\t'); if (f.empty()) return;
        std::string ver = f[0], base = ver;
        if (base.size() > 11 && base.compare(base.size() - 11, 11, ".inv-strand") == 0) base.resize(base.size() - 11);
        size_t dot = base.rfind('.'); if (dot != std::string::npos && base.size() - dot <= 4) base.resize(dot);
        auto& v = versions[base]; if (std::find(v.begin(), v.end(), ver) == v.end()) v.push_back(ver);
      }); }
    for (auto& n : bucketNames) addRef(refIdx, n);
    for (auto& kv : versions) addRef(refIdx, kv.first);
    { std::string f = slurp(fai); lines(f, [&](const std::string& l) { auto t = tok(l, '\t'); if (!t.empty()) refFlags[addRef(refIdx, t[0])] |= 1; }); }
    int32_t nRef = (int32_t)refStrOff.size();
    std::vector<std::string> refNames(nRef); for (auto& kv : refIdx) refNames[kv.second] = kv.first;
    refVerOff.push_back(0);
    for (int32_t r = 0; r < nRef; r++) {
      auto vi = versions.find(refNames[r]);
      if (vi != versions.end()) { for (auto& v : vi->second) { auto b = bucketIdx.find(v); if (b != bucketIdx.end()) refVerBucket.push_back(b->second); } }
      else { auto b = bucketIdx.find(refNames[r]); if (b != bucketIdx.end()) refVerBucket.push_back(b->second); }
      refVerOff.push_back((uint32_t)refVerBucket.size());
    }
    uint32_t cap = 16; while (cap < (uint32_t)nRef * 2u) cap <<= 1;
    dict.assign(cap, DictEntry{0, -1, 0}); dictMask = cap - 1;
    for (int32_t r = 0; r < nRef; r++) {
      uint64_t h = name_hash(refNames[r].data(), (int)refNames[r].size()); uint32_t s = (uint32_t)h & dictMask;
      while (dict[s].key) s = (s + 1) & dictMask;
      dict[s] = DictEntry{dict_key(h, (int)refNames[r].size()), r, refStrOff[r]};
    }
    // ---- exon -> feature lists.  2 columns: exon, feature id.  3 columns: exon, printed id, internal id (Exon.java:107-125): counted under
    // the internal id, printed per map line under the printed id; the junction map's rows are switched on per chunk by their column-2 id
    // (filter-junction-map-files.sh:286-296 filters `-I 1`), so the activation index of a junction row is the index of that id.
    const char* maps[3] = {geneMap, exonMap, junctionMap};
    std::unordered_map<std::string, uint32_t> jActIdx;
    for (int m = 0; m < 3; m++) {
      std::string text = slurp(maps[m]); std::set<std::string> ids, acts;
      struct MapRow { std::string exon, printed, feat, line; bool always; };
      std::vector<MapRow> rowsM; int columns = 2;
      auto readMap = [&](const std::string& txt, bool isAlways) {
        lines(txt, [&](const std::string& l) {
          auto t = tok(l, '\t'); if (t.empty()) throw AnnoError(-3, "No exon id found.");
          if (t[0] == "Exon Id") return;
          if (t.size() < 2) throw AnnoError(-3, "No count object for exon " + t[0] + " found on loading.");
          int fields = t.size() >= 3 ? 3 : 2;
          // Exon.java:126-131 rejects a line with fewer columns than an earlier one; the opposite order passes the load and then fails in
          // the 3-column writer ("No third column found", ComputeCounts.java:921-923): either way the run aborts
          if (fields != columns && !rowsM.empty()) throw AnnoError(-3, fields < columns ? "Not all lines have " + std::to_string(columns) + " columns." : "No third column found for exon " + rowsM[0].exon + ".");
          columns = fields;
          rowsM.push_back({keyStr(parseExon(t[0])), t[1], fields == 3 ? t[2] : t[1], l, isAlways});
          ids.insert(rowsM.back().feat); acts.insert(t[1]);
        });
      };
      if (m == 2) readMap(slurp(junctionTranscriptMap), true);           // the filter puts these rows first (`cat` order is irrelevant: sort -u follows)
      readMap(text, false);
      mapColumns[m] = columns;
      featIds[m].assign(ids.begin(), ids.end());
      std::unordered_map<std::string, uint32_t> fidx; for (size_t i = 0; i < featIds[m].size(); i++) fidx[featIds[m][i]] = (uint32_t)i;
      if (m == 2) { actIds.assign(acts.begin(), acts.end()); for (size_t i = 0; i < actIds.size(); i++) jActIdx[actIds[i]] = (uint32_t)i; }
      struct Ent { uint32_t f, act; };
      std::vector<std::vector<Ent>> per(exons.size());
      std::set<std::pair<std::string, std::string>> alwaysRows;
      if (m == 2) for (auto& r : rowsM) if (r.always) alwaysRows.insert({r.exon, r.printed + "\t" + r.feat});
      for (auto& r : rowsM) {
        auto ei = exonIdx.find(r.exon); if (ei == exonIdx.end()) continue;   // exon never reachable through the B file
        bool alw = m == 2 && alwaysRows.count({r.exon, r.printed + "\t" + r.feat});
        uint32_t f = fidx[r.feat] | (alw ? 0x80000000u : 0u), act = m == 2 ? jActIdx[r.printed] : 0;
        auto& v = per[ei->second]; bool have = false;
        for (auto& x : v) if ((x.f & 0x7fffffffu) == (f & 0x7fffffffu) && x.act == act) { x.f |= f; have = true; }
        if (!have) v.push_back({f, act});
      }
      featOff[m].push_back(0);
      for (auto& v : per) {                                             // by feature id: the junction run intersects two lists by merging (junction_pair)
        std::stable_sort(v.begin(), v.end(), [](const Ent& a, const Ent& b) { return (a.f & 0x7fffffffu) < (b.f & 0x7fffffffu); });
        for (auto& x : v) { feat[m].push_back(x.f); if (m == 2 && columns == 3) featAct.push_back(x.act); }
        featOff[m].push_back((uint32_t)feat[m].size()); }
      // print order
      if (columns == 2) {
        std::unordered_set<std::string> alwaysIds; for (auto& r : rowsM) if (r.always) alwaysIds.insert(r.feat);   // ids of exon_junction_transcript.map are in every filtered map
        for (size_t i = 0; i < featIds[m].size(); i++) outRows[m].push_back({featIds[m][i], (int32_t)i, m == 2 ? (int32_t)jActIdx[featIds[m][i]] : 0, (uint8_t)(m != 2 || alwaysIds.count(featIds[m][i]) ? 1 : 0)});
      }
      else if (m != 2) { for (auto& r : rowsM) outRows[m].push_back({r.printed, (int32_t)fidx[r.feat], 0, 1}); }
      else {                                                              // sort -u of whole lines, then a stable sort on the text from column 2 on
        std::map<std::string, MapRow*> uniq; for (auto& r : rowsM) { auto it = uniq.find(r.line); if (it == uniq.end()) uniq[r.line] = &r; else if (r.always) it->second->always = true; }
        std::vector<MapRow*> v; for (auto& kv : uniq) v.push_back(kv.second);
        std::stable_sort(v.begin(), v.end(), [](const MapRow* a, const MapRow* b) { return a->line.substr(a->line.find('\t') + 1) < b->line.substr(b->line.find('\t') + 1); });
        for (auto* r : v) outRows[m].push_back({r->printed, (int32_t)fidx[r->feat], (int32_t)jActIdx[r->printed], (uint8_t)(r->always ? 1 : 0)});
      }
    }
    exF.resize(exons.size());
    for (size_t x = 0; x < exons.size(); x++) {
      uint32_t n[3]; for (int m = 0; m < 3; m++) n[m] = featOff[m][x + 1] - featOff[m][x];
      if (n[0] > 255 || n[1] > 255 || n[2] > 65535) throw AnnoError(-3, "an exon maps to more than 255 genes / exons or 65535 junctions");
      exF[x].g = n[0] == 1 ? (feat[0][featOff[0][x]] & 0x7fffffffu) : featOff[0][x];
      exF[x].e = n[1] == 1 ? (feat[1][featOff[1][x]] & 0x7fffffffu) : featOff[1][x];
      exF[x].jOff = featOff[2][x]; exF[x].n = n[0] | (n[1] << 8) | (n[2] << 16);
    }
    // ---- a27: which junction features does a reference id seen in the chunk keep alive?
    std::unordered_map<std::string, std::vector<std::string>> jjExt;    // internal junction id -> external ids
    { std::string t = slurp(jjCoordinateMap);
      lines(t, [&](const std::string& l) { auto c = splitc(l, ':'); auto f = splitc(l, '\t'); if (f.size() >= 2) jjExt[c[0]].push_back(f[1]); }); }
    std::unordered_map<std::string, std::vector<std::string>> equiv;
    { std::string t = slurp(equivalentsMap);
      lines(t, [&](const std::string& l) { auto f = splitc(l, '\t'); if (f.size() >= 2) equiv[f[0]].push_back(f[1]); }); }
    refActOff.push_back(0);
    for (int32_t r = 0; r < nRef; r++) {
      const std::string& name = refNames[r];
      if (name.find('-') != std::string::npos) {
        auto c = splitc(name, ':'); std::vector<std::string> j; if (c.size() == 3) j = splitc(c[0], '-');
        if (c.size() != 3 || j.size() != 4) refFlags[r] |= 2;          // createExtendedJunctionIds.py would raise
        else {
          std::vector<std::string> idx;
          if (!c[1].empty()) for (auto& x : splitc(c[1], '-')) idx.push_back(x);
          idx.push_back(j[2]); idx.push_back(j[3]);
          if (!c[2].empty()) for (auto& x : splitc(c[2], '-')) idx.push_back(x);
          std::set<std::string> ext;
          for (size_t i = 0; i + 1 < idx.size(); i++) {
            auto e = jjExt.find(j[0] + "-" + j[1] + "-" + idx[i] + "-" + idx[i + 1]);
            if (e != jjExt.end()) ext.insert(e->second.begin(), e->second.end());
          }
          std::set<std::string> all = ext;
          for (auto& x : ext) { auto q = equiv.find(x); if (q != equiv.end()) all.insert(q->second.begin(), q->second.end()); }
          for (auto& x : all) { auto f = jActIdx.find(x); if (f != jActIdx.end()) refAct.push_back((int32_t)f->second); }
        }
      }
      refActOff.push_back((uint32_t)refAct.size());
    }
  }

  AnnoView view() const {
    AnnoView v{};
    v.dict = dict.data(); v.dictMask = dictMask; v.strPool = strPool.data();
    v.refStrOff = refStrOff.data(); v.refStrLen = refStrLen.data(); v.refFlags = refFlags.data();
    v.refVerOff = refVerOff.data(); v.refVerBucket = refVerBucket.data(); v.refActOff = refActOff.data(); v.refAct = refAct.data();
    v.nRef = (int32_t)refStrOff.size();
    v.bktRowOff = bktRowOff.data(); v.bktFlags = bktFlags.data(); v.bktRefId = bktRefId.data(); v.nBucket = (int32_t)bktFlags.size();
    v.row = rowA.data(); v.ex = exA.data(); v.exf = exF.data(); v.nExon = (int32_t)exChr.size(); v.nTrChr = (int32_t)trChrNames.size();
    v.featAct = featAct.empty() ? nullptr : featAct.data(); v.nAct = (int32_t)actIds.size();
    for (int m = 0; m < 3; m++) { v.feat[m] = feat[m].data(); v.nFeat[m] = (int32_t)featIds[m].size(); }
    return v;
  }
};

}  // namespace eqpb
