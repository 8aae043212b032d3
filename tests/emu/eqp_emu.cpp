// TEST-ONLY host emulation of the device pipeline: runs the very same per-item device functions
// (eqp-cluster_b200/csrc/*.h) in plain loops on the CPU so their logic can be diffed against the oracle
// in a container without a GPU.  Never linked into libeqpcount.so; the product path has no CPU fallback.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "../../eqp-cluster_b200/csrc/eqp_pipeline.h"
using namespace eqpb;

struct HostBackend {
  std::vector<void*> allocs;
  template <class T> T* alloc(uint64_t n) { void* p = calloc(n ? n : 1, sizeof(T)); allocs.push_back(p); return (T*)p; }
  void memset(void* p, int v, uint64_t bytes) { ::memset(p, v, bytes); }
  template <class Tag = st::misc, class F> void launch(uint64_t n, F f) { for (uint64_t i = 0; i < n; i++) f(i); }
  template <class Tag, int MINB, class F> void launch_occ(uint64_t n, F f) { for (uint64_t i = 0; i < n; i++) f(i); }
  template <class Tag = st::misc, class F> void launch_counted(const uint32_t* count, uint64_t, F f) { for (uint64_t i = 0; i < *count; i++) f(i); }
  void scan(uint32_t* p, uint64_t n) { uint32_t acc = 0; for (uint64_t i = 0; i < n; i++) { uint32_t v = p[i]; p[i] = acc; acc += v; } p[n] = acc; }
  void to_host(void* dst, const void* src, uint64_t bytes) { memcpy(dst, src, bytes); }
  void to_device(void* dst, const void* src, uint64_t bytes) { memcpy(dst, src, bytes); }
  void count_newlines(const char* text, uint64_t len, uint64_t nTiles, uint32_t* cnt) {
    for (uint64_t t = 0; t < nTiles; t++) { uint64_t b = t * LINE_TILE, e = b + LINE_TILE < len ? b + LINE_TILE : len; uint32_t c = 0; for (uint64_t i = b; i < e; i++) c += text[i] == '\n'; cnt[t] = c; }
  }
  void fill_line_offsets(const char* text, uint64_t len, uint64_t nTiles, const uint32_t* cnt, uint64_t* lo, uint32_t nLines, char last) {
    for (uint64_t t = 0; t < nTiles; t++) {
      uint64_t b = t * LINE_TILE, e = b + LINE_TILE < len ? b + LINE_TILE : len; uint32_t k = cnt[t];
      if (t == 0) lo[0] = 0;
      for (uint64_t i = b; i < e; i++) if (text[i] == '\n') lo[++k] = i + 1;
      if (e == len && last != '\n') lo[nLines] = len + 1;
    }
  }
  void stage(const char*) {}
  static bool env_flag(const char* name) { const char* e = getenv(name); return e && e[0] == '1'; }
  static const bool kFusedDecode = false;
  void decode_fused(const AnnoView&, SamView&, uint8_t*&, const char*, uint64_t, uint32_t*, uint32_t*) {}
  void fork() {}
  void branch(int) {}
  void join() {}
  uint32_t order_by_work(const uint32_t* key, uint64_t n, uint32_t* perm) {
    uint32_t k = 0;
    for (int b = 255; b >= 1; b--) for (uint64_t i = 0; i < n; i++) { uint32_t v = key[i] > 255 ? 255 : key[i]; if ((int)v == b) perm[k++] = (uint32_t)i; }
    return k;
  }
  void decode(const AnnoView& A, const SamView& S, uint32_t* err) { for (uint32_t i = 0; i < S.nLines; i++) decode_line(A, S, i, err); }
  ~HostBackend() { for (void* p : allocs) free(p); }
};

static std::string slurp(const std::string& p) { return AnnoHost::slurp(p.c_str()); }

// `eqp_emu decode <annodir> <sam>`: decode every line on its own and print the SoA fields (or the error code) — decoder unit tests
static int decode_only(const char* annoDir, const char* samPath) {
  std::string anno = annoDir, base = anno + "/ensembl_rna_hs";
  AnnoHost H;
  H.load((base + "_100_combined_exons.bed").c_str(), (base + "_transcript_exons.bed").c_str(), (base + "_exon_gene.map").c_str(),
         (base + "_exon_exon.map").c_str(), (base + "_exon_junction.map").c_str(), (base + "_exon_junction_transcript.map").c_str(),
         (base + "_100_junction_junction_coordinate.map").c_str(), (base + "_equivalent_junctions.map").c_str(), (anno + "/genome.fa.fai").c_str());
  std::string T = slurp(samPath);
  HostBackend be; RunParams P{}; Pipeline<HostBackend> pl(be, H.view(), P);
  pl.index_lines(0, T.data(), T.size());
  const SamView& S = pl.C.sam[0]; AnnoView A = H.view();
  for (uint32_t i = 0; i < S.nLines; i++) {
    uint32_t err[2] = {0, 0};
    decode_line(A, S, i, err);
    if (err[0]) { printf("%u\tERR\t%u\n", i, err[0]); continue; }
    if (!(S.rf[i] & RF_REC)) { printf("%u\tSKIP\n", i); continue; }
    std::string ref = S.ref[i] >= 0 ? std::string(A.strPool + A.refStrOff[S.ref[i]], A.refStrLen[S.ref[i]]) : (S.ref[i] == -1 ? "*" : "?");
    printf("%u\tREC\t%u\t%u\t%s\t%d\t%d\t%d\t%.*s\t%u\t%u\n", i, S.frag[i], S.flag[i], ref.c_str(), S.pos[i], S.pnext[i], S.nm[i],
           (int)S.cigLen[i], S.text + S.lineOff[i] + S.cigOff[i], S.seqLen[i], S.rf[i]);
  }
  return 0;
}

int main(int argc, char** argv) {
  if (argc == 4 && !strcmp(argv[1], "decode")) { try { return decode_only(argv[2], argv[3]); } catch (const std::exception& e) { fprintf(stderr, "ERROR: %s\n", e.what()); return 1; } }
  if (argc < 5) { fprintf(stderr, "usage: eqp_emu <annodir> <chunkdir> <se|pe> <outdir>\n"); return 2; }
  std::string anno = argv[1], chunk = argv[2], sfx = !strcmp(argv[3], "pe") ? "pe" : "sr", outd = argv[4], base = anno + "/ensembl_rna_hs";
  try {
    AnnoHost H;
    H.load((base + "_100_combined_exons.bed").c_str(), (base + "_transcript_exons.bed").c_str(), (base + "_exon_gene.map").c_str(),
           (base + "_exon_exon.map").c_str(), (base + "_exon_junction.map").c_str(), (base + "_exon_junction_transcript.map").c_str(),
           (base + "_100_junction_junction_coordinate.map").c_str(), (base + "_equivalent_junctions.map").c_str(), (anno + "/genome.fa.fai").c_str());
    bool genomicIn = false; for (int i = 5; i < argc; i++) genomicIn |= !strcmp(argv[i], "--genomic");
    std::string T = slurp(chunk + (genomicIn ? "/genomic-" : "/transcript-") + sfx + ".sam"), J = genomicIn ? std::string() : slurp(chunk + "/junction-" + sfx + ".sam"), G = genomicIn ? std::string() : slurp(chunk + "/genome-" + sfx + ".sam");
    RunParams P{}; P.overlap[0] = 1; P.overlap[1] = 5; P.overlap[2] = 8;
    for (int m = 0; m < 3; m++) { P.spliceConforming[m] = 1; P.containment[m] = 0; P.unambiguous[m] = 0; }
    P.unweighted = 0;
    for (int i = 5; i < argc; i++) {                                     // same switches as the oracle driver
      if (!strcmp(argv[i], "--unweighted")) P.unweighted = 1;
      else if (!strcmp(argv[i], "--unambig")) P.unambiguous[0] = P.unambiguous[1] = 1;
      else if (!strcmp(argv[i], "--containment")) P.containment[0] = 1;
      else if (!strcmp(argv[i], "--nonsplice")) P.spliceConforming[0] = P.spliceConforming[1] = 0;
      else if (!strcmp(argv[i], "--genomic")) P.genomic = 1;
      else if (!strcmp(argv[i], "--stranded=forward")) P.stranded = 1;
      else if (!strcmp(argv[i], "--stranded=backward")) P.stranded = 2;
    }
    P.modes = P.genomic ? 3 : 7; P.weightThreshold = 0.01; P.slack = 0.25;
    HostBackend be; Pipeline<HostBackend> pl(be, H.view(), P);
    const char* text[3] = {T.data(), J.data(), G.data()}; uint64_t len[3] = {T.size(), J.size(), G.size()};
    ChunkResult R{};
    int rc = pl.run(text, len, R);
    if (rc) { fprintf(stderr, "ERROR: device error %u detail %u\n", R.err[0], R.err[1]); return 1; }
    const char* names[3] = {"gene", "exon", "junction"};
    for (int m = 0; m < 3; m++) {
      FILE* f = fopen((outd + "/" + names[m] + ".tsv").c_str(), "w");
      for (size_t i = 0; i < H.featIds[m].size(); i++) {
        if (m == 2 && R.fixed[m][i] == 0) continue;
        fprintf(f, "%s\t%llu\t%.17g\n", H.featIds[m][i].c_str(), R.fixed[m][i], R.counts[m][i]);
      }
      fclose(f);
    }
    // the count files' rows in print order (ComputeCounts.java:875-950): gene and exon print every row, the junction run (-n) non-zero rows of the filtered map
    for (int m = 0; m < 3; m++) {
      FILE* f = fopen((outd + "/" + names[m] + ".rows").c_str(), "w");
      for (auto& r : H.outRows[m]) {
        if (m == 2 && !(r.always || R.junctionActive[r.act])) continue;
        if (m == 2 && R.fixed[m][r.feature] == 0) continue;
        fprintf(f, "%s\t%.17g\n", r.id.c_str(), R.counts[m][r.feature]);
      }
      fclose(f);
    }
    { FILE* f = fopen((outd + "/weights.tsv").c_str(), "w");
      for (uint32_t s = 0; s < pl.C.nSlots; s++) if (pl.V.pack[s].a20) fprintf(f, "F%09u\t%u\t%u\n", pl.C.base + s, pl.V.pack[s].a20, pl.V.pack[s].nFinal);
      fclose(f); }
    { FILE* f = fopen((outd + "/samweights.tsv").c_str(), "w");
      const char* fn[4] = {"T", "J", "G", "TJ"};
      for (int k = 0; k < 4; k++) for (uint32_t s = 0; s < pl.C.nSlots; s++) {
        const FragStat& st = k < 3 ? pl.C.st[k][s] : pl.C.stTJ[s];
        if (!(st.flags & FS_EXISTS)) continue;
        const char* cls = fe_mate(st) ? "paired-end" : (fe_mult(st) == 2 ? "single-read (both)" : "single-read");
        fprintf(f, "%s\tF%09u\t%d\t%d\t%s\n", fn[k], pl.C.base + s, fe_n(st), fe_sum(st), cls);
      }
      fclose(f); }
    { FILE* f = fopen((outd + "/stats.txt").c_str(), "w");
      fprintf(f, "NUMBER_MAPPED_READS=%llu\nNUMBER_EXPRESSED_READS=%llu\nNUM_WEIGHT_OBJECTS=%llu\nBLOCKS=%u\nHITS=%u\nALIGNMENTS=%u\nSUM_WEIGHTS_GENE=%.4f\nSUM_WEIGHTS_EXON=%.4f\nSUM_WEIGHTS_JUNCTION=%.4f\n", R.stats[9], R.stats[1], R.stats[0], pl.nBlocks, pl.nHits, pl.nAln,
              (double)R.stats[10] / 4294967296.0, (double)R.stats[11] / 4294967296.0, (double)R.stats[12] / 4294967296.0);
      fclose(f); }
    return 0;
  } catch (const std::exception& e) { fprintf(stderr, "ERROR: %s\n", e.what()); return 1; }
}
