// worldgen — seeded synthetic EQP world: annotation files + read-ordered SAM streams.
//
// Nothing like this exists in the reference (its annotation/test blobs are missing,
// /root/reference/.MISSING_LARGE_BLOBS).  File formats follow SURVEY.md App. A.1, which
// infers them from the reference code that parses them:
//   <base>_<R>_combined_exons.bed     BedRecord.java:288-331, Exon.java:230-301
//   <base>_transcript_exons.bed       ConvertSamBed.java:58-77 (only col 1 is read)
//   <base>_exon_{gene,exon}.map       Exon.java:70-152 (header "Exon Id" skipped)
//   <base>_exon_junction.map, _exon_junction_transcript.map,
//   <base>_<R>_junction_junction_coordinate.map, _equivalent_junctions.map
//                                     filter-junction-map-files.sh:254-304
//   genome.fa.fai                     ComputeReadWeightsSam.java:44-58 (col 1 = whitelist)
//   SAM streams shaped like `bowtie2 --reorder -k N` output (bowtie2-align.sh:344-357):
//   read-ordered, unaligned reads present (flag 4), NM:i/AS:i tags, no NH.
//
// Usage:
//   worldgen anno  <outdir> <seed> <n_genes>
//   worldgen reads <annodir> <outdir> <seed> <first_frag> <n_frags> <se|pe> [spliced_frac]
// `reads` re-derives the world from <annodir>/world.seed so annotation and reads agree.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

static const int RADIUS = 100;       // bin/setup.sh:5
static const int READ_LEN = 101;     // BASELINE.json configs: 101 bp

struct Rng {                         // splitmix64: reproducible across platforms
  uint64_t s;
  explicit Rng(uint64_t seed) : s(seed * 0x9E3779B97F4A7C15ULL + 0x1234567ULL) {}
  uint64_t next() {
    uint64_t z = (s += 0x9E3779B97F4A7C15ULL);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
  }
  double uni() { return (next() >> 11) * (1.0 / 9007199254740992.0); }
  int below(int n) { return (int)(next() % (uint64_t)n); }
  double normal() {
    double u1 = uni() + 1e-300, u2 = uni();
    return std::sqrt(-2.0 * std::log(u1)) * std::cos(6.283185307179586 * u2);
  }
  int geom(double p, int cap) { int k = 0; while (k < cap && uni() > p) k++; return k; }
};

struct GExon { int s, e; };                           // 1-based closed genome coords
struct Transcript { int gene; std::vector<int> ex; int len; std::string id; };
struct JRef {                                         // one junction reference
  int gene; int a, b;                                 // exon indices (genomic order), a<b
  std::vector<int> pex, ps, pe;                       // pieces: exon idx, genome start/end
  int len; std::string id; std::string ext;           // ext = external junction id
  bool annotated;
};
struct Gene {
  int chr; char strand; std::string id;
  std::vector<GExon> ex;                              // distinct exons, genomic order
  std::vector<int> tr;                                // transcript indices
  std::vector<int> jr;                                // junction-reference indices
  double expr;
};
struct World {
  std::vector<std::string> chrs; std::vector<int> chrLen; int nWhite;
  std::vector<Gene> genes; std::vector<Transcript> trs; std::vector<JRef> jrs;
  std::vector<double> cumExpr;
};

static std::string fmt(const char* f, ...) __attribute__((format(printf, 1, 2)));
#include <cstdarg>
static std::string fmt(const char* f, ...) {
  char buf[512]; va_list ap; va_start(ap, f); vsnprintf(buf, sizeof buf, f, ap); va_end(ap);
  return buf;
}

static void build_world(World& w, uint64_t seed, int nGenes) {
  Rng r(seed);
  int nChr = std::max(2, std::min(25, nGenes / 40 + 2));
  for (int c = 0; c < nChr; c++) {
    w.chrs.push_back(c < 22 ? fmt("chr%d", c + 1) : c == 22 ? "chrX" : c == 23 ? "chrY" : "chrM");
    w.chrLen.push_back(0);
  }
  w.nWhite = nChr;                                    // scaffolds below are NOT in genome.fa.fai
  w.chrs.push_back("GL000191.1"); w.chrLen.push_back(0);
  std::vector<int> cursor(w.chrs.size(), 10000), lastOnChr(w.chrs.size(), -1);
  for (int g = 0; g < nGenes; g++) {
    Gene G; G.id = fmt("ENSG%011d", g + 1);
    G.chr = (r.uni() < 0.01) ? nChr : r.below(nChr);
    G.strand = (r.uni() < 0.5) ? '+' : '-';
    int nEx = 1 + std::min(29, (int)(-8.0 * std::log(r.uni() + 1e-12)));
    // 8% of genes overlap the previous gene on the chromosome (shared region, maybe shared exon)
    int pos = cursor[G.chr];
    int prev = lastOnChr[G.chr];
    bool overlapPrev = r.uni() < 0.08 && prev >= 0;
    bool shareExon = overlapPrev && r.uni() < 0.4;      // identical exon in two genes
    if (overlapPrev) pos = w.genes[prev].ex.back().s + (shareExon ? 0 : 5 + r.below(10));
    for (int k = 0; k < nEx; k++) {
      int len = (int)std::exp(std::log(130.0) + 0.8 * r.normal());
      len = std::max(20, std::min(6000, len));
      if (k == 0 && shareExon) len = w.genes[prev].ex.back().e - pos + 1;
      GExon e{pos, pos + len - 1};
      G.ex.push_back(e);
      int intron = (int)std::exp(std::log(1500.0) + 0.9 * r.normal());
      pos = e.e + 1 + std::max(70, std::min(200000, intron));
    }
    cursor[G.chr] = std::max(cursor[G.chr], pos + 2000);
    lastOnChr[G.chr] = g;
    G.expr = std::exp(2.0 * r.normal());
    // transcripts: 1 + Geom(0.3) capped 30, each a subset of exons (first always full set)
    int nTr = 1 + r.geom(0.3, 29);
    for (int t = 0; t < nTr; t++) {
      Transcript T; T.gene = g; T.len = 0;
      for (int k = 0; k < nEx; k++)
        if (t == 0 || nEx == 1 || r.uni() < 0.8) T.ex.push_back(k);
      if (T.ex.empty()) T.ex.push_back(r.below(nEx));
      for (int k : T.ex) T.len += G.ex[k].e - G.ex[k].s + 1;
      bool dup = false;
      for (int ti : G.tr) if (w.trs[ti].ex == T.ex) dup = true;
      if (dup) continue;
      T.id = fmt("ENST%011d", (int)w.trs.size() + 1);
      G.tr.push_back((int)w.trs.size());
      w.trs.push_back(T);
    }
    w.genes.push_back(G);
  }
  for (size_t c = 0; c < w.chrs.size(); c++) w.chrLen[c] = cursor[c] + 10000;
  // junction references: annotated (consecutive in a transcript) + novel pairs with b-a<=3
  for (int g = 0; g < nGenes; g++) {
    Gene& G = w.genes[g]; int nEx = (int)G.ex.size();
    std::vector<std::vector<char>> ann(nEx, std::vector<char>(nEx, 0));
    for (int ti : G.tr) { auto& ex = w.trs[ti].ex; for (size_t k = 0; k + 1 < ex.size(); k++) ann[ex[k]][ex[k + 1]] = 1; }
    for (int a = 0; a < nEx; a++) for (int b = a + 1; b < nEx; b++) {
      if (!ann[a][b] && b - a > 3) continue;
      if (G.ex[b].s <= G.ex[a].e + 1) continue;
      JRef J; J.gene = g; J.a = a; J.b = b; J.annotated = ann[a][b];
      // left side: last RADIUS bases of a, extended by whole exons a-1, a-2.. while short
      std::vector<int> lext, rext; int have = G.ex[a].e - G.ex[a].s + 1; int k = a - 1;
      while (have < RADIUS && k >= 0) { lext.insert(lext.begin(), k); have += G.ex[k].e - G.ex[k].s + 1; k--; }
      for (int x : lext) { J.pex.push_back(x); J.ps.push_back(G.ex[x].s); J.pe.push_back(G.ex[x].e); }
      J.pex.push_back(a); J.ps.push_back(lext.empty() ? std::max(G.ex[a].s, G.ex[a].e - RADIUS + 1) : G.ex[a].s); J.pe.push_back(G.ex[a].e);
      have = G.ex[b].e - G.ex[b].s + 1; k = b + 1;
      while (have < RADIUS && k < nEx) { rext.push_back(k); have += G.ex[k].e - G.ex[k].s + 1; k++; }
      J.pex.push_back(b); J.ps.push_back(G.ex[b].s); J.pe.push_back(rext.empty() ? std::min(G.ex[b].e, G.ex[b].s + RADIUS - 1) : G.ex[b].e);
      for (int x : rext) { J.pex.push_back(x); J.ps.push_back(G.ex[x].s); J.pe.push_back(G.ex[x].e); }
      J.len = 0; for (size_t p = 0; p < J.pex.size(); p++) J.len += J.pe[p] - J.ps[p] + 1;
      std::string le, re;
      for (int x : lext) le += (le.empty() ? "" : "-") + fmt("%02d", x + 1);
      for (int x : rext) re += (re.empty() ? "" : "-") + fmt("%02d", x + 1);
      J.id = G.id + fmt("-junction-%02d-%02d:", a + 1, b + 1) + le + ":" + re;
      J.ext = fmt("JCT_%s_%d_%d_%c", w.chrs[G.chr].c_str(), G.ex[a].e, G.ex[b].s, G.strand);
      G.jr.push_back((int)w.jrs.size());
      w.jrs.push_back(J);
    }
  }
  double acc = 0;
  for (auto& G : w.genes) { acc += G.expr; w.cumExpr.push_back(acc); }
}
