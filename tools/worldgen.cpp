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
//   worldgen anno  <outdir> <seed> <n_genes> [key=value ...]
//   worldgen reads <annodir> <outdir> <seed> <first_frag> <n_frags> <se|pe> [spliced_frac]
//   worldgen reads <annodir> <outdir> <seed> <first_frag> <n_frags> <se|pe> genomic      one genomic-<sr|pe>.sam of a spliced genome aligner
// `reads` re-derives the world from <annodir>/world.seed so annotation and reads agree.
// Annotation shape knobs (SURVEY.md 8d "Ensembl76-shaped"; defaults in Knobs):
//   ld=F     fraction of transcripts with one exon whose transcript length differs from its genome length by +-1..3 bp
//            (BedRecord.java:417-452; a positive difference makes GenomicInterval.adjustChrStartEnd effective, :284-421)
//   tr=F     fraction of transcripts carrying a `tr-` filler exon between two exons (Exon.java:230-301, GenomicAlignment.java:233-234)
//   trlong=F fraction of those fillers that are longer than a read (a read half can then consist of the filler alone)
//   inv=F    fraction of transcripts listed as <id>.inv-strand (BedRecord.java:154-161)
//   ver=F    fraction of transcripts with a versioned id <id>.<n>; ver2=F fraction with TWO versioned ids (ConvertSamBed.java:47-98)
//   scaf=N   number of scaffolds (in genome.fa.fai) next to the chromosomes
//   map3=0|1 write the junction maps with 3 columns: exon, printed id, internal id (Exon.java:107-125)
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
struct Transcript {
  int gene; std::vector<int> ex; int len; std::string id;
  std::vector<int> tl;                                // transcript-space length of each exon (tl[i] - genome length = lengthDifference)
  int fillAfter = -1, fillLen = 0; bool fillInMap = false;   // tr- filler exon after ex[fillAfter] (genomic order), its length, listed in exon_gene.map
  bool inv = false;                                   // listed as <id>.inv-strand: the transcript runs against its exons' strand
  std::vector<std::string> names;                     // column-1 names in the BED files (versioned ids); SAM RNAME is `id`
};
struct Knobs { double ld = 0.005, tr = 0.01, trlong = 0.0, inv = 0.002, ver = 0.7, ver2 = 0.02; int scaf = 150, map3 = 0; };
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
  Knobs k;
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
  int nMain = std::max(2, std::min(25, nGenes / 40 + 2));
  for (int c = 0; c < nMain; c++) {
    w.chrs.push_back(c < 22 ? fmt("chr%d", c + 1) : c == 22 ? "chrX" : c == 23 ? "chrY" : "chrM");
    w.chrLen.push_back(0);
  }
  int nScaf = std::max(0, std::min(w.k.scaf, nGenes / 20));
  for (int c = 0; c < nScaf; c++) { w.chrs.push_back(fmt("KI27%04d.1", c + 1)); w.chrLen.push_back(0); }
  int nChr = nMain + nScaf;
  w.chrs.push_back("GL000191.1"); w.chrLen.push_back(0);
  w.nWhite = nChr + 1;                                // the decoy used for repeat copies is NOT in genome.fa.fai
  std::vector<int> cursor(w.chrs.size(), 10000), lastOnChr(w.chrs.size(), -1);
  for (int g = 0; g < nGenes; g++) {
    Gene G; G.id = fmt("ENSG%011d", g + 1);
    { double u = r.uni(); G.chr = u < 0.01 ? nChr : (u < 0.03 && nScaf > 0) ? nMain + r.below(nScaf) : r.below(nMain); }
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
      bool dup = false;
      for (int ti : G.tr) if (w.trs[ti].ex == T.ex) dup = true;
      if (dup) continue;
      for (int k : T.ex) T.tl.push_back(G.ex[k].e - G.ex[k].s + 1);
      Rng q(seed * 7919ULL + (uint64_t)w.trs.size() * 104729ULL + 17);   // own stream: the knobs do not disturb the gene layout
      if (q.uni() < w.k.tr && T.ex.size() >= 2) {      // tr- filler exon between two exons of the transcript
        T.fillAfter = q.below((int)T.ex.size() - 1);
        T.fillLen = q.uni() < w.k.trlong ? 110 + q.below(60) : 5 + q.below(26);
        T.fillInMap = q.uni() < 0.5;
      }
      // one exon whose transcript length differs by +-1..3 bp; not together with a filler: an alignment with both a tr- record and a
      // length difference gets different coordinates in the junction run (the device reports DERR_TR_LENDIFF)
      if (q.uni() < w.k.ld && T.fillAfter < 0) {
        int i = q.below((int)T.ex.size()); int d = (1 + q.below(3)) * (q.uni() < 0.5 ? 1 : -1);
        if (T.tl[i] + d >= 15) T.tl[i] += d;
      }
      T.inv = q.uni() < w.k.inv;
      for (int l : T.tl) T.len += l;
      T.len += T.fillLen;
      T.id = fmt("ENST%011d", (int)w.trs.size() + 1);
      { double u = q.uni(); int v = 1 + q.below(9);
        if (u < w.k.ver2) { T.names.push_back(T.id + fmt(".%d", v)); T.names.push_back(T.id + fmt(".%d", v + 1)); }
        else if (u < w.k.ver2 + w.k.ver) T.names.push_back(T.id + fmt(".%d", v));
        else T.names.push_back(T.id);
        if (T.inv) for (auto& n : T.names) n += ".inv-strand"; }
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

// ---------------------------------------------------------------- annotation files
static std::string exon_key(const World& w, const Gene& G, int k) {     // map key (no gene prefix)
  return fmt("%s/%d/%d/%c", w.chrs[G.chr].c_str(), G.ex[k].s, G.ex[k].e, G.strand);
}
static std::string tr_name(const Transcript& T) {                      // chromosome part of the transcript's tr- exon ids
  std::string n = T.names[0]; size_t p = n.find(".inv-strand"); if (p != std::string::npos) n.resize(p); return n;
}
static FILE* xopen(const std::string& p, const char* m) {
  FILE* f = fopen(p.c_str(), m);
  if (!f) { fprintf(stderr, "worldgen: cannot open %s\n", p.c_str()); exit(1); }
  return f;
}

static void write_annotation(const World& w, const std::string& dir, uint64_t seed, int nGenes) {
  const std::string base = dir + "/ensembl_rna_hs";
  FILE* fs = xopen(dir + "/world.seed", "w");
  fprintf(fs, "%llu %d %.17g %.17g %.17g %.17g %.17g %.17g %d %d\n", (unsigned long long)seed, nGenes, w.k.ld, w.k.tr, w.k.trlong, w.k.inv, w.k.ver, w.k.ver2, w.k.scaf, w.k.map3);
  fclose(fs);
  FILE* fai = xopen(dir + "/genome.fa.fai", "w");
  for (int c = 0; c < w.nWhite; c++) fprintf(fai, "%s\t%d\t0\t60\t61\n", w.chrs[c].c_str(), w.chrLen[c]);
  fclose(fai);
  FILE* bed = xopen(base + fmt("_%d_combined_exons.bed", RADIUS), "w");
  FILE* tbed = xopen(base + "_transcript_exons.bed", "w");
  // transcript rows: exons in transcript (5'->3') order, transcript coordinates, 0-based half-open.  An .inv-strand transcript
  // runs against its exons' strand (BedRecord.java:154-161 inverts the exon strand for the mapping, the exon id keeps it).
  for (auto& T : w.trs) {
    const Gene& G = w.genes[T.gene]; int n = (int)T.ex.size();
    bool flip = (G.strand == '-') != T.inv;
    for (const std::string& name : T.names) {
      int off = 0;
      for (int i = 0; i < n; i++) {
        int gi = flip ? n - 1 - i : i; int k = T.ex[gi]; int len = T.tl[gi];
        if (flip && T.fillAfter == gi) {                 // filler sits between ex[fillAfter] and ex[fillAfter + 1] (genomic order)
          fprintf(bed, "%s\t%d\t%d\t%s/tr-%s/%d/%d/%c\t0\t%c\n", name.c_str(), off, off + T.fillLen, G.id.c_str(), tr_name(T).c_str(), off + 1, off + T.fillLen, G.strand, G.strand);
          off += T.fillLen;
        }
        fprintf(bed, "%s\t%d\t%d\t%s/%s\t0\t%c\n", name.c_str(), off, off + len, G.id.c_str(), exon_key(w, G, k).c_str(), G.strand);
        fprintf(tbed, "%s\t%d\t%d\t%s/%s\t0\t%c\n", name.c_str(), off, off + len, G.id.c_str(), exon_key(w, G, k).c_str(), G.strand);
        off += len;
        if (!flip && T.fillAfter == gi) {
          fprintf(bed, "%s\t%d\t%d\t%s/tr-%s/%d/%d/%c\t0\t%c\n", name.c_str(), off, off + T.fillLen, G.id.c_str(), tr_name(T).c_str(), off + 1, off + T.fillLen, G.strand, G.strand);
          off += T.fillLen;
        }
      }
    }
  }
  // junction rows: pieces in genome order, junction coordinates
  for (auto& J : w.jrs) {
    const Gene& G = w.genes[J.gene]; int off = 0;
    for (size_t p = 0; p < J.pex.size(); p++) {
      int len = J.pe[p] - J.ps[p] + 1;
      fprintf(bed, "%s\t%d\t%d\t%s/%s\t0\t+\n", J.id.c_str(), off, off + len, G.id.c_str(), exon_key(w, G, J.pex[p]).c_str());
      off += len;
    }
  }
  // genome rows: one per distinct (chr,start,end,strand); first gene wins (avoids SURVEY H5 corner)
  {
    std::vector<std::string> seen;
    std::vector<std::pair<std::string, std::string>> rows;
    for (auto& G : w.genes) for (size_t k = 0; k < G.ex.size(); k++) {
      std::string key = exon_key(w, G, (int)k);
      rows.push_back({key, fmt("%s\t%d\t%d\t%s/%s\t0\t%c\n", w.chrs[G.chr].c_str(), G.ex[k].s - 1, G.ex[k].e, G.id.c_str(), key.c_str(), G.strand)});
    }
    std::stable_sort(rows.begin(), rows.end(), [](auto& a, auto& b) { return a.first < b.first; });
    for (size_t i = 0; i < rows.size(); i++)
      if (i == 0 || rows[i].first != rows[i - 1].first) fputs(rows[i].second.c_str(), bed);
  }
  fclose(bed); fclose(tbed);
  // <base>_genome_exons.bed: the B file of compute-genomic-counts.sh (:775): the genome rows only
  { FILE* gb = xopen(base + "_genome_exons.bed", "w");
    std::vector<std::pair<std::string, std::string>> rows;
    for (auto& G : w.genes) for (size_t k = 0; k < G.ex.size(); k++) {
      std::string key = exon_key(w, G, (int)k);
      rows.push_back({key, fmt("%s\t%d\t%d\t%s/%s\t0\t%c\n", w.chrs[G.chr].c_str(), G.ex[k].s - 1, G.ex[k].e, G.id.c_str(), key.c_str(), G.strand)});
    }
    std::stable_sort(rows.begin(), rows.end(), [](auto& a, auto& b) { return a.first < b.first; });
    for (size_t i = 0; i < rows.size(); i++) if (i == 0 || rows[i].first != rows[i - 1].first) fputs(rows[i].second.c_str(), gb);
    fclose(gb); }
  FILE* eg = xopen(base + "_exon_gene.map", "w"); FILE* ee = xopen(base + "_exon_exon.map", "w");
  fprintf(eg, "Exon Id\tGene Id\n"); fprintf(ee, "Exon Id\tExon Id\n");
  int exonNo = 0;
  for (auto& G : w.genes) for (size_t k = 0; k < G.ex.size(); k++) {
    std::string key = exon_key(w, G, (int)k);
    fprintf(eg, "%s\t%s\n", key.c_str(), G.id.c_str());
    fprintf(ee, "%s\tENSE%011d\n", key.c_str(), ++exonNo);
  }
  for (auto& T : w.trs) if (T.fillAfter >= 0 && T.fillInMap) {       // tr- exon ids of half the fillers are known to the gene map
    const Gene& G = w.genes[T.gene]; int n = (int)T.ex.size(); bool flip = (G.strand == '-') != T.inv; int off = 0;
    for (int i = 0; i < n; i++) {
      int gi = flip ? n - 1 - i : i;
      if (flip && T.fillAfter == gi) break;
      off += T.tl[gi];
      if (!flip && T.fillAfter == gi) break;
    }
    fprintf(eg, "tr-%s/%d/%d/%c\t%s\n", tr_name(T).c_str(), off + 1, off + T.fillLen, G.strand, G.id.c_str());
  }
  fclose(eg); fclose(ee);
  FILE* ej = xopen(base + "_exon_junction.map", "w"); FILE* ejt = xopen(base + "_exon_junction_transcript.map", "w");
  FILE* jj = xopen(base + fmt("_%d_junction_junction_coordinate.map", RADIUS), "w");
  FILE* eq = xopen(base + "_equivalent_junctions.map", "w");
  std::vector<std::pair<std::string, std::string>> extByCoord;         // (coord w/o strand, ext id)
  int jNo = 0;
  for (auto& J : w.jrs) {
    const Gene& G = w.genes[J.gene];
    std::string tail = J.ext;                                          // 3-column variant: exon, printed id, internal id (Exon.java:107-125)
    if (w.k.map3) tail += fmt("\tJI%08d", ++jNo);
    for (int k : {J.a, J.b}) {
      fprintf(ej, "%s\t%s\n", exon_key(w, G, k).c_str(), tail.c_str());
      if (J.annotated) fprintf(ejt, "%s\t%s\n", exon_key(w, G, k).c_str(), tail.c_str());
    }
    fprintf(jj, "%s-junction-%02d-%02d:%s:%d:%d\t%s\n", G.id.c_str(), J.a + 1, J.b + 1, w.chrs[G.chr].c_str(), G.ex[J.a].e, G.ex[J.b].s, J.ext.c_str());
    extByCoord.push_back({fmt("%s_%d_%d", w.chrs[G.chr].c_str(), G.ex[J.a].e, G.ex[J.b].s), J.ext});
  }
  std::sort(extByCoord.begin(), extByCoord.end());
  for (size_t i = 0; i + 1 < extByCoord.size(); i++)
    if (extByCoord[i].first == extByCoord[i + 1].first && extByCoord[i].second != extByCoord[i + 1].second) {
      fprintf(eq, "%s\t%s\n", extByCoord[i].second.c_str(), extByCoord[i + 1].second.c_str());
      fprintf(eq, "%s\t%s\n", extByCoord[i + 1].second.c_str(), extByCoord[i].second.c_str());
    }
  fclose(ej); fclose(ejt); fclose(jj); fclose(eq);
}

// ---------------------------------------------------------------- read placement
struct Blk { int ex, gs, ge; };                       // exon idx, genome start/end (1-based closed)
struct Place { int gene, tr, go; std::vector<Blk> b; bool clean; };   // go: offset in genomic-order transcript coords; clean: no filler, no length difference touched

// Footprint of `span` transcript bases at genomic-order offset `go` of transcript T.  An exon whose transcript length differs from
// its genome length by d carries its extra (d > 0) / missing (d < 0) bases at its genomic start.
static bool place(const World& w, int ti, int go, int span, Place& P) {
  const Transcript& T = w.trs[ti]; const Gene& G = w.genes[T.gene];
  if (go < 0 || go + span > T.len) return false;
  P.gene = T.gene; P.tr = ti; P.go = go; P.b.clear(); P.clean = true;
  int off = 0, need = span, cur = go;
  for (size_t i = 0; i < T.ex.size(); i++) {
    int k = T.ex[i]; int glen = G.ex[k].e - G.ex[k].s + 1, tl = T.tl[i], d = tl - glen;
    if (cur < off + tl && need > 0) {
      int x0 = cur - off, take = std::min(need, tl - x0);
      int a = std::max(0, std::min(glen - 1, x0 - d)), z = std::max(0, std::min(glen - 1, x0 + take - 1 - d));
      P.b.push_back({k, G.ex[k].s + a, G.ex[k].s + z});
      if (d != 0) P.clean = false;
      need -= take; cur += take;
    }
    off += tl;
    if (T.fillAfter == (int)i) {
      if (cur < off + T.fillLen && need > 0) { int take = std::min(need, off + T.fillLen - cur); need -= take; cur += take; P.clean = false; }
      off += T.fillLen;
    }
  }
  return need == 0;
}
// Offset of the same genomic footprint in another transcript of the gene, or -1.
static int iso_offset(const World& w, int ti, const Place& P, int span) {
  const Transcript& T = w.trs[ti]; const Gene& G = w.genes[T.gene];
  if (ti == P.tr) return P.go;
  if (!P.clean) return -1;
  int off = 0; size_t j = 0; int res = -1;
  for (size_t i = 0; i < T.ex.size(); i++) {
    int k = T.ex[i]; int glen = G.ex[k].e - G.ex[k].s + 1;
    if (j < P.b.size() && k == P.b[j].ex) {
      if (j == 0) { res = off + (P.b[0].gs - G.ex[k].s) + (T.tl[i] - glen); if (res < off) return -1; }
      j++;
      if (j == P.b.size()) return res + span <= T.len ? res : -1;
      if (T.fillAfter == (int)i) return -1;              // a filler separates the two exons in this transcript
    } else if (j > 0) return -1;                       // blocks must sit in consecutive exons
    off += T.tl[i];
    if (T.fillAfter == (int)i) off += T.fillLen;
  }
  return -1;
}
// Offset of the footprint inside junction reference J, or -1.
static int jref_offset(const JRef& J, const Place& P) {
  int off = 0;
  for (size_t p = 0; p < J.pex.size(); p++) {
    if (J.pex[p] == P.b[0].ex && P.b[0].gs >= J.ps[p] && P.b[0].ge <= J.pe[p]) {
      int res = off + (P.b[0].gs - J.ps[p]);
      for (size_t j = 1; j < P.b.size(); j++) {
        size_t q = p + j;
        if (q >= J.pex.size() || J.pex[q] != P.b[j].ex || P.b[j].gs != J.ps[q] || P.b[j].ge > J.pe[q]) return -1;
        if (P.b[j - 1].ge != J.pe[q - 1]) return -1;
      }
      return res;
    }
    off += J.pe[p] - J.ps[p] + 1;
  }
  return -1;
}

// ---------------------------------------------------------------- SAM emission
struct Out {
  std::string buf; FILE* f;
  void flush() { if (!buf.empty()) { fwrite(buf.data(), 1, buf.size(), f); buf.clear(); } }
};
static char g_seq[1 << 16], g_qual[1 << 16];
struct ReadShape { const char* cigar; int span; int nmExtra; };
static const ReadShape SHAPES[4] = {{"101M", 101, 0}, {"50M1I50M", 100, 1}, {"50M1D51M", 102, 1}, {"98M3S", 98, 0}};

static void emit(Out& o, int frag, int flag, const char* rname, int pos, const char* cigar,
                 const char* rnext, int pnext, int tlen, int nm, int seqOff) {
  char line[700];
  int n;
  if (flag & 4)
    n = snprintf(line, sizeof line, "F%09d\t%d\t*\t0\t0\t*\t*\t0\t0\t%.*s\t%.*s\tYT:Z:U%c\n", frag, flag,
                 READ_LEN, g_seq + seqOff, READ_LEN, g_qual + seqOff, (flag & 1) ? 'P' : 'U');
  else
    n = snprintf(line, sizeof line, "F%09d\t%d\t%s\t%d\t%d\t%s\t%s\t%d\t%d\t%.*s\t%.*s\tAS:i:%d\tXN:i:0\tXM:i:%d\tXO:i:0\tXG:i:0\tNM:i:%d\tMD:Z:%d\tYT:Z:U%c\n",
                 frag, flag, rname, pos, nm == 0 ? 42 : 1, cigar, rnext, pnext, tlen,
                 READ_LEN, g_seq + seqOff, READ_LEN, g_qual + seqOff, -6 * nm, nm, nm, READ_LEN, (flag & 1) ? 'P' : 'U');
  o.buf.append(line, n);
  if (o.buf.size() > (1u << 22)) o.flush();
}

static int pick_gene(const World& w, Rng& r) {
  double x = r.uni() * w.cumExpr.back();
  return (int)(std::lower_bound(w.cumExpr.begin(), w.cumExpr.end(), x) - w.cumExpr.begin());
}
static int draw_nm(Rng& r) { double u = r.uni(); return u < 0.67 ? 0 : u < 0.94 ? 1 : u < 0.99 ? 2 : 3; }

struct Aln { std::string ref; int pos, rev, nm; };
struct MateAlns { std::vector<Aln> t, j, g; };

// All alignments a bowtie2 -k run would report for footprint P of one locus.
static void locus_alns(const World& w, const Place& P, int span, int rv, bool boundaryJ, int nm, MateAlns& M) {
  const Gene& G = w.genes[P.gene]; int gflip = (G.strand == '-');
  for (int ti : G.tr) {
    int off = iso_offset(w, ti, P, span);
    if (off < 0) continue;
    int tflip = gflip ^ (w.trs[ti].inv ? 1 : 0);
    int o = tflip ? w.trs[ti].len - off - span : off;
    M.t.push_back({w.trs[ti].id, o + 1, rv ^ (w.trs[ti].inv ? 1 : 0), nm});
  }
  if (!P.clean) return;                                // the genome and the junction references do not contain this sequence
  if (P.b.size() == 1) M.g.push_back({w.chrs[G.chr], P.b[0].gs, rv ^ gflip, nm});
  int nb = 0;
  for (int ji : G.jr) {
    const JRef& J = w.jrs[ji];
    if (P.b.size() == 1 && (!boundaryJ || nb >= 4)) continue;
    int off = jref_offset(J, P);
    if (off < 0 || off + span > J.len) continue;
    M.j.push_back({J.id, off + 1, rv ^ gflip, nm});
    if (P.b.size() == 1) nb++;
  }
}

static bool random_place(const World& w, Rng& r, int span, Place& P, int gene = -1) {
  for (int attempt = 0; attempt < 50; attempt++) {
    int g = gene >= 0 ? gene : pick_gene(w, r);
    const Gene& G = w.genes[g];
    int ti = G.tr[r.below((int)G.tr.size())];
    if (w.trs[ti].len < span) { if (gene >= 0 && attempt > 10) return false; continue; }
    if (place(w, ti, r.below(w.trs[ti].len - span + 1), span, P)) return true;
  }
  return false;
}

static void emit_file(Out& o, int frag, bool pe, std::vector<Aln>* a, size_t cap, const char* cigar[2], int seqOff[2]) {
  if (!pe) {
    if (a[0].empty()) { emit(o, frag, 4, "*", 0, "*", "*", 0, 0, 0, seqOff[0]); return; }
    for (size_t i = 0; i < a[0].size() && i < cap; i++)
      emit(o, frag, a[0][i].rev ? 16 : 0, a[0][i].ref.c_str(), a[0][i].pos, cigar[0], "*", 0, 0, a[0][i].nm, seqOff[0]);
    return;
  }
  std::vector<char> used0(a[0].size(), 0), used1(a[1].size(), 0);
  size_t n = 0;
  for (size_t i = 0; i < a[0].size() && n < cap; i++)
    for (size_t k = 0; k < a[1].size(); k++) {
      if (used1[k] || a[1][k].ref != a[0][i].ref) continue;
      int d = a[1][k].pos - a[0][i].pos;
      if (d < -500 || d > 500) continue;
      const Aln &x = a[0][i], &y = a[1][k];
      int tl = d >= 0 ? d + READ_LEN : d - READ_LEN;
      emit(o, frag, 1 + 2 + 64 + (x.rev ? 16 : 0) + (y.rev ? 32 : 0), x.ref.c_str(), x.pos, cigar[0], "=", y.pos, tl, x.nm, seqOff[0]);
      emit(o, frag, 1 + 2 + 128 + (y.rev ? 16 : 0) + (x.rev ? 32 : 0), y.ref.c_str(), y.pos, cigar[1], "=", x.pos, -tl, y.nm, seqOff[1]);
      used0[i] = used1[k] = 1; n++;
      break;
    }
  bool any[2] = {n > 0, n > 0};
  for (int m = 0; m < 2; m++) {
    std::vector<char>& used = m ? used1 : used0; size_t c = 0;
    for (size_t i = 0; i < a[m].size() && c < cap && n == 0; i++) {     // bowtie2 reports unpaired hits only without a concordant pair
      if (used[i]) continue;
      emit(o, frag, 1 + 8 + (m ? 128 : 64) + (a[m][i].rev ? 16 : 0), a[m][i].ref.c_str(), a[m][i].pos, cigar[m], "*", 0, 0, a[m][i].nm, seqOff[m]);
      any[m] = true; c++;
    }
  }
  for (int m = 0; m < 2; m++)
    if (!any[m]) emit(o, frag, 1 + 4 + (m ? 128 : 64) + (any[1 - m] ? 0 : 8), "*", 0, "*", "*", 0, 0, 0, seqOff[m]);
}

static void gen_reads(const World& w, const std::string& dir, uint64_t seed, int first, int n, bool pe, double splicedBoost) {
  Rng init(seed ^ 0xABCDEFULL);
  for (size_t i = 0; i < sizeof g_seq; i++) { g_seq[i] = "ACGT"[init.below(4)]; g_qual[i] = (char)(35 + init.below(38)); }
  const char* sfx = pe ? "pe" : "sr";
  Out T{"", xopen(dir + "/transcript-" + sfx + ".sam", "w")};
  Out J{"", xopen(dir + "/junction-" + sfx + ".sam", "w")};
  Out Gn{"", xopen(dir + "/genome-" + sfx + ".sam", "w")};
  if (first == 1) {                                    // header only in chunk C001 (bowtie2-align.sh:462-467)
    T.buf += "@HD\tVN:1.0\tSO:unsorted\n@PG\tID:bowtie2\tPN:bowtie2\tVN:2.0.5\n";
    J.buf += "@HD\tVN:1.0\tSO:unsorted\n@PG\tID:bowtie2\tPN:bowtie2\tVN:2.0.5\n";
  }
  std::vector<int> novel;                              // junction refs that no transcript contains
  for (size_t i = 0; i < w.jrs.size(); i++) if (!w.jrs[i].annotated) novel.push_back((int)i);
  for (int f = 0; f < n; f++) {
    int frag = first + f;
    Rng r(seed * 1000003ULL + (uint64_t)frag);          // per-fragment stream: chunks are independent
    MateAlns M[2];
    int shape[2] = {0, 0};
    for (int m = 0; m < 2; m++) { double u = r.uni(); shape[m] = u < 0.96 ? 0 : u < 0.97 ? 1 : u < 0.98 ? 2 : 3; }
    const char* cigar[2] = {SHAPES[shape[0]].cigar, SHAPES[shape[1]].cigar};
    int span[2] = {SHAPES[shape[0]].span, SHAPES[shape[1]].span};
    int seqOff[2] = {r.below(60000), r.below(60000)};
    int nMates = pe ? 2 : 1;
    double cat = r.uni();
    if (cat < 0.03) {
      // unaligned everywhere
    } else if (cat < 0.08) {                           // intronic / intergenic: genome only
      const Gene& G = w.genes[pick_gene(w, r)];
      int lo = std::max(1, G.ex.front().s - 300), hi = G.ex.back().e + 300;
      int p = lo + r.below(hi - lo + 1), rv = r.below(2);
      M[0].g.push_back({w.chrs[G.chr], p, rv, draw_nm(r)});
      if (pe && r.uni() < 0.8) M[1].g.push_back({w.chrs[G.chr], p + 100 + r.below(250), 1 - rv, draw_nm(r)});
    } else if (cat < 0.12 && !novel.empty()) {         // novel junction: junction reference only
      const JRef& Jr = w.jrs[novel[r.below((int)novel.size())]];
      int joint = 0; for (size_t p = 0; p < Jr.pex.size() && Jr.pex[p] != Jr.b; p++) joint += Jr.pe[p] - Jr.ps[p] + 1;
      int lo = std::max(0, joint - span[0] + 8), hi = std::min(joint - 8, Jr.len - span[0]);
      if (hi >= lo) M[0].j.push_back({Jr.id, lo + r.below(hi - lo + 1) + 1, r.below(2), draw_nm(r)});
      if (pe && r.uni() < 0.5) {                       // mate 2 from a transcript of the same gene
        Place P; if (random_place(w, r, span[1], P, Jr.gene)) locus_alns(w, P, span[1], r.below(2), false, draw_nm(r), M[1]);
      }
    } else {                                           // transcript-derived, maybe multi-mapped
      int nLoci = 1; double u = r.uni();
      if (u < 0.10) nLoci = 2 + r.below(4); else if (u < 0.115) nLoci = 6 + r.below(15);
      bool wantSpliced = r.uni() < splicedBoost;
      for (int l = 0; l < nLoci; l++) {
        Place P[2]; int rv = r.below(2); int nm0 = draw_nm(r) + (l > 0 ? r.below(2) : 0);
        bool ok = false;
        for (int attempt = 0; attempt < 8 && !ok; attempt++) {
          if (!random_place(w, r, span[0], P[0])) continue;
          ok = !wantSpliced || P[0].b.size() > 1 || attempt == 7;
        }
        if (!ok) continue;
        bool boundaryJ = r.uni() < 0.5;
        locus_alns(w, P[0], span[0], rv, boundaryJ, nm0 + SHAPES[shape[0]].nmExtra, M[0]);
        if (pe && r.uni() > 0.06) {                    // 6%: only one mate aligns
          const Transcript& Tr = w.trs[P[0].tr];
          int ins = std::max(READ_LEN, std::min(500, (int)(300 + 50 * r.normal())));
          int go2 = rv ^ (w.genes[P[0].gene].strand == '-') ? P[0].go - (ins - span[0]) : P[0].go + ins - span[1];
          go2 = std::max(0, std::min(Tr.len - span[1], go2));
          if (place(w, P[0].tr, go2, span[1], P[1]))
            locus_alns(w, P[1], span[1], 1 - rv, boundaryJ, draw_nm(r) + SHAPES[shape[1]].nmExtra, M[1]);
        }
      }
      if (r.uni() < 0.01) {                            // highly repetitive: many genome-only copies
        int extra = 21 + r.below(100);
        for (int e = 0; e < extra; e++) {
          Place P; if (!random_place(w, r, span[0], P) || P.b.size() != 1 || !P.clean) continue;
          bool decoy = r.uni() < 0.1;                    // exercises the -c whitelist (ComputeReadWeightsSam.java:168-172)
          M[0].g.push_back({decoy ? std::string("GL000999.1_decoy") : w.chrs[w.genes[P.gene].chr], P.b[0].gs, r.below(2), 1 + r.below(2)});
        }
      }
      if (pe && r.uni() < 0.03) {                      // mates best in different files (CombineSamFiles.java:312-349)
        M[1] = MateAlns(); M[0].g.clear(); M[0].j.clear();
        const Gene& G = w.genes[pick_gene(w, r)];
        M[1].g.push_back({w.chrs[G.chr], std::max(1, G.ex.front().s - 200 + r.below(400)), r.below(2), 0});
      }
    }
    std::vector<Aln> a[2];
    a[0] = M[0].t; a[1] = M[1].t; emit_file(T, frag, pe, a, 2000, cigar, seqOff);
    a[0] = M[0].j; a[1] = M[1].j; emit_file(J, frag, pe, a, 400, cigar, seqOff);
    a[0] = M[0].g; a[1] = M[1].g; emit_file(Gn, frag, pe, a, 100, cigar, seqOff);
    (void)nMates;
  }
  T.flush(); J.flush(); Gn.flush(); fclose(T.f); fclose(J.f); fclose(Gn.f);
}

// ---------------------------------------------------------------- genomic-aligner output (compute-genomic-counts.sh)
// One name-sorted SAM as a spliced genome aligner (tophat2 / STAR / hisat) writes it: reads aligned to the chromosomes, introns as
// N operations, every alignment of a fragment numbered by an HI:i tag that both mates of a pair share (SamRecord.java:233-240,658-660),
// NH:i present (and ignored by the reference).  ids stay Fnnnnnnnnn (the reference renumbers foreign ids itself, SamReader.java:157-186).
static std::string spliced_cigar(const Place& P, int span, int shape) {
  std::string c; int used = 0;
  for (size_t i = 0; i < P.b.size(); i++) {
    int len = P.b[i].ge - P.b[i].gs + 1; used += len;
    if (i) c += fmt("%dN", P.b[i].gs - P.b[i - 1].ge - 1);
    if (i == 0 && shape == 1 && len > 12) c += fmt("%dM1I%dM", 6, len - 7);          // an insertion in the first block (read length = span)
    else if (i + 1 == P.b.size() && shape == 2 && len > 12) c += fmt("%dM1D%dM", len - 6, 5);   // a deletion in the last block: 1 reference base is skipped
    else c += fmt("%dM", len);
  }
  (void)span; (void)used;
  return c;
}
static void gen_genomic(const World& w, const std::string& dir, uint64_t seed, int first, int n, bool pe) {
  Rng init(seed ^ 0xABCDEFULL);
  for (size_t i = 0; i < sizeof g_seq; i++) { g_seq[i] = "ACGT"[init.below(4)]; g_qual[i] = (char)(35 + init.below(38)); }
  Out S{"", xopen(dir + "/genomic-" + (pe ? "pe" : "sr") + ".sam", "w")};
  if (first == 1) { S.buf += "@HD\tVN:1.0\tSO:queryname\n"; for (size_t c = 0; c < w.chrs.size(); c++) S.buf += fmt("@SQ\tSN:%s\tLN:%d\n", w.chrs[c].c_str(), w.chrLen[c]); }
  auto put = [&](int frag, int flag, const char* rname, int pos, const std::string& cigar, int seqLen, const char* rnext, int pnext, int tlen, int nm, int nh, int hi, int seqOff) {
    char line[900]; int k;
    if (flag & 4) k = snprintf(line, sizeof line, "F%09d\t%d\t*\t0\t0\t*\t*\t0\t0\t%.*s\t%.*s\tYT:Z:UU\n", frag, flag, seqLen, g_seq + seqOff, seqLen, g_qual + seqOff);
    else k = snprintf(line, sizeof line, "F%09d\t%d\t%s\t%d\t%d\t%s\t%s\t%d\t%d\t%.*s\t%.*s\tAS:i:%d\tNM:i:%d\tNH:i:%d\tHI:i:%d\n", frag, flag, rname, pos, nh == 1 ? 50 : 3,
                      cigar.c_str(), rnext, pnext, tlen, seqLen, g_seq + seqOff, seqLen, g_qual + seqOff, -6 * nm, nm, nh, hi);
    S.buf.append(line, k); if (S.buf.size() > (1u << 22)) S.flush();
  };
  for (int f = 0; f < n; f++) {
    int frag = first + f; Rng r(seed * 1000003ULL + (uint64_t)frag);
    int seqOff[2] = {r.below(60000), r.below(60000)};
    double cat = r.uni();
    struct GA { int chr; int pos[2]; std::string cig[2]; int len[2]; int rev; int nm[2]; bool two; };
    std::vector<GA> alns;
    if (cat >= 0.03) {
      int nLoci = 1; double u = r.uni(); if (u < 0.10) nLoci = 2 + r.below(4); else if (u < 0.115) nLoci = 6 + r.below(10);
      for (int l = 0; l < nLoci; l++) {
        GA a; a.two = false; a.rev = r.below(2);
        if (cat < 0.10) {                                                  // intronic / intergenic, unspliced
          const Gene& G = w.genes[pick_gene(w, r)]; a.chr = G.chr;
          int lo = std::max(1, G.ex.front().s - 300), hi = G.ex.back().e + 300;
          a.pos[0] = lo + r.below(hi - lo + 1); a.cig[0] = "101M"; a.len[0] = READ_LEN; a.nm[0] = draw_nm(r);
          if (pe && r.uni() < 0.8) { a.two = true; a.pos[1] = a.pos[0] + 100 + r.below(250); a.cig[1] = "101M"; a.len[1] = READ_LEN; a.nm[1] = draw_nm(r); }
        } else {                                                           // from a transcript: the footprint's blocks become M runs, the introns N runs
          Place P[2]; int sh[2] = {0, 0}; for (int m = 0; m < 2; m++) { double v = r.uni(); sh[m] = v < 0.95 ? 0 : v < 0.975 ? 1 : 2; }
          bool ok = false;
          for (int attempt = 0; attempt < 8 && !ok; attempt++) ok = random_place(w, r, READ_LEN, P[0]) && P[0].clean;
          if (!ok) continue;
          a.chr = w.genes[P[0].gene].chr; a.pos[0] = P[0].b[0].gs; a.cig[0] = spliced_cigar(P[0], 0, sh[0]); a.nm[0] = draw_nm(r) + (sh[0] ? 1 : 0);
          a.len[0] = a.cig[0].find('D') != std::string::npos ? READ_LEN - 1 : READ_LEN;   // "xM1D5M" reads one base less than the footprint's reference span
          if (pe && r.uni() > 0.06) {
            const Transcript& Tr = w.trs[P[0].tr]; int ins = std::max(READ_LEN, std::min(500, (int)(300 + 50 * r.normal())));
            int go2 = std::max(0, std::min(Tr.len - READ_LEN, P[0].go + ins - READ_LEN));
            if (place(w, P[0].tr, go2, READ_LEN, P[1]) && P[1].clean) { a.two = true; a.pos[1] = P[1].b[0].gs; a.cig[1] = spliced_cigar(P[1], 0, 0); a.len[1] = READ_LEN; a.nm[1] = draw_nm(r); }
          }
        }
        alns.push_back(a);
      }
    }
    if (alns.empty()) {
      if (!pe) put(frag, 4, "*", 0, "*", READ_LEN, "*", 0, 0, 0, 0, 0, seqOff[0]);
      else { put(frag, 77, "*", 0, "*", READ_LEN, "*", 0, 0, 0, 0, 0, seqOff[0]); put(frag, 141, "*", 0, "*", READ_LEN, "*", 0, 0, 0, 0, 0, seqOff[1]); }
      continue;
    }
    int nh = (int)alns.size();
    for (int k = 0; k < nh; k++) {
      const GA& a = alns[k]; const char* chr = w.chrs[a.chr].c_str(); int sec = k ? 256 : 0;
      if (!pe) put(frag, (a.rev ? 16 : 0) | sec, chr, a.pos[0], a.cig[0], a.len[0], "*", 0, 0, a.nm[0], nh, k + 1, seqOff[0]);
      else if (a.two) {
        int tl = a.pos[1] - a.pos[0] + READ_LEN;
        put(frag, 1 + 2 + 64 + (a.rev ? 16 : 32) + sec, chr, a.pos[0], a.cig[0], a.len[0], "=", a.pos[1], tl, a.nm[0], nh, k + 1, seqOff[0]);
        put(frag, 1 + 2 + 128 + (a.rev ? 32 : 16) + sec, chr, a.pos[1], a.cig[1], a.len[1], "=", a.pos[0], -tl, a.nm[1], nh, k + 1, seqOff[1]);
      } else put(frag, 1 + 8 + 64 + (a.rev ? 16 : 0) + sec, chr, a.pos[0], a.cig[0], a.len[0], "*", 0, 0, a.nm[0], nh, k + 1, seqOff[0]);
    }
  }
  S.flush(); fclose(S.f);
}

int main(int argc, char** argv) {
  if (argc >= 5 && !strcmp(argv[1], "anno")) {
    World w; uint64_t seed = strtoull(argv[3], 0, 10); int ng = atoi(argv[4]);
    for (int i = 5; i < argc; i++) {
      const char* eq = strchr(argv[i], '='); std::string key(argv[i], eq ? eq - argv[i] : strlen(argv[i])); double v = eq ? atof(eq + 1) : 0;
      if (key == "ld") w.k.ld = v; else if (key == "tr") w.k.tr = v; else if (key == "trlong") w.k.trlong = v; else if (key == "inv") w.k.inv = v;
      else if (key == "ver") w.k.ver = v; else if (key == "ver2") w.k.ver2 = v; else if (key == "scaf") w.k.scaf = (int)v; else if (key == "map3") w.k.map3 = (int)v;
      else { fprintf(stderr, "worldgen: unknown knob %s\n", argv[i]); return 2; }
    }
    build_world(w, seed, ng); write_annotation(w, argv[2], seed, ng);
    size_t nLd = 0, nTr = 0, nInv = 0, nV2 = 0;
    for (auto& T : w.trs) { const Gene& G = w.genes[T.gene]; for (size_t i = 0; i < T.ex.size(); i++) if (T.tl[i] != G.ex[T.ex[i]].e - G.ex[T.ex[i]].s + 1) { nLd++; break; } nTr += T.fillAfter >= 0; nInv += T.inv; nV2 += T.names.size() > 1; }
    fprintf(stderr, "worldgen: %zu genes, %zu transcripts (%zu with a length difference, %zu with a tr- exon, %zu inv-strand, %zu with two versions), %zu junction references, %zu reference sequences\n",
            w.genes.size(), w.trs.size(), nLd, nTr, nInv, nV2, w.jrs.size(), w.chrs.size());
    return 0;
  }
  if (argc >= 8 && !strcmp(argv[1], "reads")) {
    unsigned long long seed; int ng; World w;
    FILE* fs = xopen(std::string(argv[2]) + "/world.seed", "r");
    if (fscanf(fs, "%llu %d %lf %lf %lf %lf %lf %lf %d %d", &seed, &ng, &w.k.ld, &w.k.tr, &w.k.trlong, &w.k.inv, &w.k.ver, &w.k.ver2, &w.k.scaf, &w.k.map3) != 10) return 1;
    fclose(fs);
    build_world(w, seed, ng);
    if (argc > 8 && !strcmp(argv[8], "genomic")) gen_genomic(w, argv[3], strtoull(argv[4], 0, 10), atoi(argv[5]), atoi(argv[6]), !strcmp(argv[7], "pe"));
    else gen_reads(w, argv[3], strtoull(argv[4], 0, 10), atoi(argv[5]), atoi(argv[6]), !strcmp(argv[7], "pe"), argc > 8 ? atof(argv[8]) : 0.0);
    return 0;
  }
  fprintf(stderr, "usage: worldgen anno <outdir> <seed> <n_genes> [ld=F tr=F trlong=F inv=F ver=F ver2=F scaf=N map3=0|1] | reads <annodir> <outdir> <seed> <first_frag> <n_frags> <se|pe> [spliced_frac | genomic]\n");
  return 2;
}
