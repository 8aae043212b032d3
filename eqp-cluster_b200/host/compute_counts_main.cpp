// eqp_compute_counts — host-side drop-in for the body of compute-counts.sh (exon-pipeline-scripts/bin/process-scripts/
// compute-counts.sh:133-1108): same options, same positional arguments, same count-file names and the same log lines, with
// the ~25 java/bedtools/python processes replaced by ONE call into libeqpcount.so (include/eqp_count.h).
// The reference's host language for this path is Java; no JDK exists in the build image, so this mirror is C++ (INTEGRATION.md
// shows the equivalent Java main + JNI/Panama binding).  No counting happens here: without the library or a GPU it exits 1.
//
// usage: eqp_compute_counts [-a|-g|-e|-j] [-E n] [-J n] [-W w] [-sr] [--unweighted] [-s dir] [-o prefix] [-nj] [--unambig]
//                           [-r] [-C] [-N] [-M mem] [--device n] [--lib path] <project dir> <transcript.sam[.gz]> <junction.sam[.gz]> <genome.sam[.gz]>
#include <dlfcn.h>
#include <sys/stat.h>
#include <cstdint>
#include <mutex>
#include <thread>
#include <vector>
#include "eqp_count.h"
#include "host_util.hpp"

using namespace eqphost;

#define FN(name) decltype(&::name) p_##name = (decltype(&::name))dlsym(lib, #name); if (!p_##name) { fprintf(stderr, "ERROR: %s missing from library\n", #name); return 1; }

static std::string dirname_of(const std::string& p) { size_t s = p.rfind('/'); return s == std::string::npos ? "." : p.substr(0, s); }
static std::string basename_of(const std::string& p) { size_t s = p.rfind('/'); return s == std::string::npos ? p : p.substr(s + 1); }
static bool ends_with(const std::string& s, const std::string& t) { return s.size() >= t.size() && s.compare(s.size() - t.size(), t.size(), t) == 0; }
static std::string strip_sam_ext(std::string b) {                         // sed -e 's/.gz$//' | sed -e 's/.[bs]am$//'  (CC:424)
  if (ends_with(b, ".gz")) b.resize(b.size() - 3);
  if (ends_with(b, ".sam") || ends_with(b, ".bam")) b.resize(b.size() - 4);
  return b;
}
static std::string replace_first(std::string s, const std::string& a, const std::string& b) { size_t k = s.find(a); if (k != std::string::npos) s.replace(k, a.size(), b); return s; }
static std::string pick(const std::string& plain) {                       // map files may be stored gzipped (CC:600, filter-junction-map-files.sh:207-236)
  struct stat st; if (stat(plain.c_str(), &st) == 0) return plain;
  std::string gz = plain + ".gz"; if (stat(gz.c_str(), &st) == 0) return gz;
  return plain;
}

int main(int argc, char** argv) {
  bool gene = false, exon = false, junction = false, singleRead = false, nonZeroJunctions = false; int device = 0;
  std::string outPrefix, libPath, strandDir; const char* env;
  std::string fileBase = (env = getenv("FILE_BASE")) ? env : "ensembl_rna_hs";          // bin/setup.sh:3-12
  int readLength = (env = getenv("READ_LENGTH")) ? atoi(env) : 100, radius = (env = getenv("RADIUS")) ? atoi(env) : 100;
  bool pairedEnd = !((env = getenv("PAIRED_END")) && !strcmp(env, "FALSE"));
  eqp_params P; memset(&P, 0, sizeof P); P.gene_overlap = 1; P.exon_overlap = 5; P.junction_overlap = -1; P.read_weight_threshold = 0.01; P.distance_slack = 0.25;
  int i = 1;
  for (; i < argc && argv[i][0] == '-' && argv[i][1]; i++) {
    std::string a = argv[i];
    auto val = [&]() -> const char* { if (i + 1 >= argc) { fprintf(stderr, "ERROR: option %s needs a value\n", a.c_str()); exit(1); } return argv[++i]; };
    if (a == "-a") gene = exon = junction = true; else if (a == "-g") gene = true; else if (a == "-e") exon = true; else if (a == "-j") junction = true;
    else if (a == "-E") P.exon_overlap = atoi(val()); else if (a == "-J") P.junction_overlap = atoi(val()); else if (a == "-W") P.read_weight_threshold = atof(val());
    else if (a == "-sr") singleRead = true; else if (a == "--unweighted" || a == "-u") P.unweighted = 1; else if (a == "-s") strandDir = val();
    else if (a == "-o") outPrefix = val(); else if (a == "-nj") nonZeroJunctions = true; else if (a == "--unambig" || a == "-unambig") P.unambiguous = 1;
    else if (a == "-r") {} else if (a == "-C") P.containment = 1; else if (a == "-N") P.nonsplice_conforming = 1; else if (a == "-O") {} else if (a == "-M") val();
    else if (a == "--device") device = atoi(val()); else if (a == "--lib") libPath = val();
    else if (a == "-js") { fprintf(stderr, "ERROR: -js (junction SAM counts) is not part of the compute-counts default flow and is not supported\n"); return 1; }
    else { fprintf(stderr, "ERROR: unknown option %s\n", a.c_str()); return 1; }
  }
  if (argc - i < 4) { fprintf(stderr, "Usage: %s <options> <project dir> <transcript SAM file> <junction SAM file> <genome SAM-file>\n", argv[0]); return 1; }
  std::string project = argv[i], tPath = argv[i + 1], jPath = argv[i + 2], gPath = argv[i + 3];
  if (singleRead) pairedEnd = false;
  if (P.junction_overlap < 0) P.junction_overlap = readLength - radius > 8 ? readLength - radius : 8;   // CC:391-401
  P.paired_end = pairedEnd; P.stranded = strandDir.empty() ? 0 : (strandDir == "backward" ? 2 : 1);
  P.modes = (gene ? 1 : 0) | (exon ? 2 : 0) | (junction ? 4 : 0);
  if (!P.modes) { fprintf(stderr, "ERROR: none of -a, -g, -e, -j given\n"); return 1; }
  // log header scraped by getJobStatus.py / extractPipelineMetaData.py (CC:641-646)
  printf("READ_WEIGHT_THRESHOLD=%s\nEXON_OVERLAP=%d\nJUNCTION_OVERLAP=%d\nSTRAND_SPECIFIC=%s\nSTRAND_SPECIFIC_DIRECTION=%s\nPAIRED_END=%s\n",
         java_double(P.read_weight_threshold).c_str(), P.exon_overlap, P.junction_overlap, strandDir.empty() ? "FALSE" : "TRUE", strandDir.c_str(), pairedEnd ? "TRUE" : "FALSE");
  // naming convention checks and output names (CC:424-487, 655-696, 915-1026)
  std::string samDir = dirname_of(tPath), tBase = strip_sam_ext(basename_of(tPath));
  if (tBase.find(pairedEnd ? "-pe" : "-sr") == std::string::npos && !(pairedEnd && tBase.find("-mixed") != std::string::npos)) {
    printf("%s does not conform to naming convention (%s) ... exiting.\n", tBase.c_str(), pairedEnd ? "-pe|-mixed" : "-sr"); return 1; }
  std::string combinedBase = replace_first(tBase, "-transcript", "-combined"); if (combinedBase == tBase) combinedBase = tBase + "-combined";
  std::string countDir = replace_first(samDir, "/sam-files", "/count-files");
  if (outPrefix.empty()) mkdir(countDir.c_str(), 0775);
  std::string outBase = outPrefix.empty() ? countDir + "/" + combinedBase : outPrefix;
  try {
    std::string here = dirname_of(argv[0]);
    if (libPath.empty()) libPath = here + "/libeqpcount.so";
    void* lib = dlopen(libPath.c_str(), RTLD_NOW | RTLD_LOCAL);
    if (!lib) { fprintf(stderr, "ERROR: cannot load %s: %s (there is no CPU fallback)\n", libPath.c_str(), dlerror()); return 1; }
    FN(eqp_create) FN(eqp_load_annotation) FN(eqp_chunk_begin) FN(eqp_chunk_feed) FN(eqp_chunk_end) FN(eqp_map_columns) FN(eqp_output_row_count) FN(eqp_get_output_row)
    FN(eqp_last_error) FN(eqp_destroy)
    eqp_ctx* ctx = nullptr;
    if (p_eqp_create(device, &ctx)) { fprintf(stderr, "ERROR: %s\n", p_eqp_last_error(nullptr)); return 1; }
    std::string ex = project + "/exon-pipeline-files", bed = ex + "/bed-files/" + fileBase, map = ex + "/map-files/" + fileBase, R = std::to_string(radius);
    std::string f0 = bed + "_" + R + "_combined_exons.bed", f1 = bed + "_transcript_exons.bed", f2 = map + "_exon_gene.map", f3 = map + "_exon_exon.map",
                f4 = pick(map + "_exon_junction.map"), f5 = pick(map + "_exon_junction_transcript.map"), f6 = pick(map + "_" + R + "_junction_junction_coordinate.map"),
                f7 = pick(map + "_equivalent_junctions.map"), f8 = ex + "/genome-files/genome.fa.fai";
    eqp_annotation_files af{f0.c_str(), f1.c_str(), f2.c_str(), f3.c_str(), f4.c_str(), f5.c_str(), f6.c_str(), f7.c_str(), f8.c_str()};
    if (p_eqp_load_annotation(ctx, &af)) { fprintf(stderr, "ERROR: %s\n", p_eqp_last_error(ctx)); return 1; }
    // `zcat <file> | java ...` (CIB:644-662) becomes: one inflate thread per input, each feeding the device through eqp_chunk_feed as the
    // text comes out of zlib, so inflating, the host-to-device copies and (for the later inputs) each other overlap.  The context is not
    // thread-safe: feed calls are serialised by a mutex (they only copy into a page-locked staging buffer and queue an async copy).
    eqp_counts C; bool counted = false;
    // Three `.gz` inputs: read them whole and let eqp_count_chunk_gz inflate them side by side on the device.
    { const std::string* paths[3] = {&tPath, &jPath, &gPath}; auto p_count_gz = (decltype(&::eqp_count_chunk_gz))dlsym(lib, "eqp_count_chunk_gz");
      bool allGz = p_count_gz && !getenv("EQP_HOST_INFLATE");
      for (int f = 0; f < 3; f++) allGz = allGz && paths[f]->size() > 3 && !paths[f]->compare(paths[f]->size() - 3, 3, ".gz");
      if (allGz) {
        std::vector<uint8_t> z[3]; bool readOk[3] = {true, true, true}; std::vector<std::thread> rd;
        for (int f = 0; f < 3; f++) rd.emplace_back([&, f] {
          FILE* fp = fopen(paths[f]->c_str(), "rb"); if (!fp) { readOk[f] = false; return; }
          uint8_t buf[1 << 16]; size_t n; while ((n = fread(buf, 1, sizeof buf, fp)) > 0) z[f].insert(z[f].end(), buf, buf + n); fclose(fp); });
        for (auto& t : rd) t.join();
        for (int f = 0; f < 3; f++) if (!readOk[f]) { fprintf(stderr, "ERROR: cannot open %s\n", paths[f]->c_str()); return 1; }
        const int rc = p_count_gz(ctx, &P, z[0].data(), z[0].size(), z[1].data(), z[1].size(), z[2].data(), z[2].size(), &C);
        if (rc == EQP_OK) counted = true;
        else if (rc != EQP_ERR_UNSUPPORTED) { fprintf(stderr, "ERROR: %s\n", p_eqp_last_error(ctx)); return 1; }
      } }
    if (!counted) {
    if (p_eqp_chunk_begin(ctx, 0, 0, 0)) { fprintf(stderr, "ERROR: %s\n", p_eqp_last_error(ctx)); return 1; }
    // A `.gz` input goes to the device COMPRESSED and is inflated there (eqp_chunk_feed_gz); a file that decoder does not take (several gzip
    // members, BGZF) or EQP_HOST_INFLATE=1 leaves it to the zlib thread below.
    bool onDevice[3] = {false, false, false};
    { const std::string* paths[3] = {&tPath, &jPath, &gPath}; auto p_feed_gz = (decltype(&::eqp_chunk_feed_gz))dlsym(lib, "eqp_chunk_feed_gz");
      for (int f = 0; f < 3 && p_feed_gz && !getenv("EQP_HOST_INFLATE"); f++) {
        const std::string& path = *paths[f];
        if (path.size() < 3 || path.compare(path.size() - 3, 3, ".gz")) continue;
        FILE* fp = fopen(path.c_str(), "rb"); if (!fp) { fprintf(stderr, "ERROR: cannot open %s\n", path.c_str()); return 1; }
        std::vector<uint8_t> z; { uint8_t buf[1 << 16]; size_t n; while ((n = fread(buf, 1, sizeof buf, fp)) > 0) z.insert(z.end(), buf, buf + n); } fclose(fp);
        const int rc = z.size() >= 18 ? p_feed_gz(ctx, f, z.data(), z.size()) : EQP_ERR_UNSUPPORTED;
        if (rc == EQP_OK) onDevice[f] = true;
        else if (rc != EQP_ERR_UNSUPPORTED) { fprintf(stderr, "ERROR: %s\n", p_eqp_last_error(ctx)); return 1; }
      } }
    { std::mutex mu; std::string failed; const std::string* paths[3] = {&tPath, &jPath, &gPath}; std::vector<std::thread> th;
      for (int f = 0; f < 3; f++) if (!onDevice[f]) th.emplace_back([&, f] {
        try {
          stream_maybe_gz(*paths[f], [&](const char* buf, size_t n) {
            std::lock_guard<std::mutex> lk(mu);
            if (failed.empty() && p_eqp_chunk_feed(ctx, f, (const uint8_t*)buf, n)) failed = p_eqp_last_error(ctx);
            return failed.empty();
          });
        } catch (const std::exception& e) { std::lock_guard<std::mutex> lk(mu); if (failed.empty()) failed = e.what(); }
      });
      for (auto& t : th) t.join();
      if (!failed.empty()) { fprintf(stderr, "ERROR: %s\n", failed.c_str()); return 1; } }
    if (p_eqp_chunk_end(ctx, &P, &C)) { fprintf(stderr, "ERROR: %s\n", p_eqp_last_error(ctx)); return 1; }
    }
    printf("NUMBER_MAPPED_READS=%lld\n", (long long)C.number_mapped_reads);                      // ConvertSamBed.java:266, CIB:975-976
    const char* names[3] = {"gene", "exon", "junction"}; const double* vec[3] = {C.gene, C.exon, C.junction};
    const uint64_t* fx[3] = {C.gene_fixed, C.exon_fixed, C.junction_fixed};
    for (int m = 0; m < 3; m++) {
      if (!(P.modes & (1 << m))) continue;
      std::string path = outBase + "-" + names[m] + ".cnt";
      FILE* f = fopen(path.c_str(), "w"); if (!f) { fprintf(stderr, "ERROR: cannot write %s\n", path.c_str()); return 1; }
      // rows in the reference's print order (ComputeCounts.java:875-950): 2-column map = ids in TreeSet order, "0" when absent; 3-column map =
      // one row per map line under its column-2 id, "0.0" when absent; -n (the junction run under -nj, CC:1034-1037) drops the zero rows
      int64_t n = p_eqp_output_row_count(ctx, m); bool nonZeroOnly = m == 2 && nonZeroJunctions; const char* zero = p_eqp_map_columns(ctx, m) == 3 ? "0.0" : "0";
      for (int64_t k = 0; k < n; k++) {
        eqp_output_row r; if (p_eqp_get_output_row(ctx, m, k, &r)) { fprintf(stderr, "ERROR: %s\n", p_eqp_last_error(ctx)); return 1; }
        if (m == 2 && !(r.always || C.junction_active[r.act])) continue;                                 // not in this chunk's filtered junction map (a27)
        if (fx[m][r.feature] == 0) { if (!nonZeroOnly) fprintf(f, "%s\t%s\n", r.id, zero); continue; }
        fprintf(f, "%s\t%s\n", r.id, java_double(vec[m][r.feature]).c_str());
      }
      fclose(f);
      double nw = (double)C.num_weight_objects[m];                                                       // ComputeCounts.java:986-989
      fprintf(stderr, "Num weight objects in BED file: %lld, num weight objects included: %lld (%.2f%%), num weight objects excluded: %lld (%.2f%%).\n", (long long)C.num_weight_objects[m],
              (long long)C.num_weight_objects_included[m], C.num_weight_objects_included[m] * 100.0 / nw, (long long)C.num_weight_objects_excluded[m], C.num_weight_objects_excluded[m] * 100.0 / nw);
      fprintf(stderr, "Sum of weights contributing to count objects: %.2f\n", C.sum_weights[m]);
    }
    if (P.modes & 1) printf("NUMBER_EXPRESSED_READS=%lld\n", (long long)C.number_expressed_reads);        // ComputeCounts.java:991-993
    p_eqp_destroy(ctx);
    printf("Compute counts successfully completed.\n");                                                  // CC:1108
    return 0;
  } catch (const std::exception& e) { fprintf(stderr, "ERROR: %s\n", e.what()); return 1; }
}
