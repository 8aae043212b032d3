// ORACLE (test infrastructure) — driver restating compute-counts.sh (CC:733-1037),
// compute-intersection-bed-file.sh (CIB:639-940) and filter-junction-map-files.sh:254-304.
// Writes every reference intermediate so tests can compare stage by stage.
#include <chrono>
#include "convert.hpp"
#include "count4.hpp"
using namespace eqp;

static std::string cut_f3_sort_u(std::initializer_list<sv> sams) {        // `cut -f 3 | sort -u` (CIB:830,888-898)
  std::set<std::string> ids;
  for (sv sam : sams) { Lines L(sam); sv line; while (L.next(line)) { auto t = split(line, '\t'); ids.insert(std::string(t.size() >= 3 ? t[2] : sv())); } }
  std::string out; for (auto& s : ids) out += s + "\n"; return out;
}

// filter-junction-map-files.sh:254-304 (+ createExtendedJunctionIds.py:93-114, filterFile.py:131-194)
static std::string filter_junction_map(sv refIds, sv exonJunction, sv exonJunctionTranscript, sv jjCoordinate, sv equivalents) {
  std::unordered_set<std::string> extended;
  { Lines L(refIds); sv line;
    while (L.next(line)) {
      if (line.find('-') == sv::npos) continue;
      auto c = split(line, ':');
      if (c.size() != 3) throw Error("createExtendedJunctionIds: need junctionId:leftExt:rightExt in " + std::string(line));
      auto j = split(c[0], '-');
      if (j.size() != 4) throw Error("createExtendedJunctionIds: need gene-fill-i-j in " + std::string(line));
      std::vector<sv> idx;
      if (!c[1].empty()) for (sv x : split(c[1], '-')) idx.push_back(x);
      idx.push_back(j[2]); idx.push_back(j[3]);
      if (!c[2].empty()) for (sv x : split(c[2], '-')) idx.push_back(x);
      for (size_t i = 0; i + 1 < idx.size(); i++)
        extended.insert(std::string(j[0]) + "-" + std::string(j[1]) + "-" + std::string(idx[i]) + "-" + std::string(idx[i + 1]));
    } }
  std::set<std::string> external;
  { Lines L(jjCoordinate); sv line;
    while (L.next(line)) { auto c = split(line, ':'); if (extended.count(std::string(c[0]))) { auto t = split(line, '\t'); if (t.size() >= 2) external.insert(std::string(t[1])); } } }
  { std::set<std::string> extra; Lines L(equivalents); sv line;
    while (L.next(line)) { auto t = split(line, '\t'); if (external.count(std::string(t[0])) && t.size() >= 2) extra.insert(std::string(t[1])); }
    external.insert(extra.begin(), extra.end()); }
  std::set<std::string> rows;
  { Lines L(exonJunctionTranscript); sv line; while (L.next(line)) rows.insert(std::string(line)); }
  { Lines L(exonJunction); sv line; while (L.next(line)) { auto t = split(line, '\t'); if (t.size() >= 2 && external.count(std::string(t[1]))) rows.insert(std::string(line)); } }
  std::vector<std::string> v(rows.begin(), rows.end());
  std::stable_sort(v.begin(), v.end(), [](const std::string& a, const std::string& b) { return a.substr(a.find('\t') + 1) < b.substr(b.find('\t') + 1); });
  std::string out; for (auto& s : v) out += s + "\n"; return out;
}

// compute-genomic-counts.sh:1040-1058,1126-1152 — one SAM of a spliced genome aligner: ComputeReadWeightsSam (no -c) for the weights,
// ConvertSamBed (no -T) | bedtools intersect -b <base>_genome_exons.bed, then ComputeCounts -g / -e with that weight file.  The junction
// run of that script filters its map through ExtractSplicedExonExonIds and a GTF (:1210-1290) and is not part of this restatement.
static int run_genomic(int argc, char** argv) {
  std::string anno = argv[2], samPath = argv[3], outd = argv[4];
  int strandDir = 0; bool stranded = false, unweighted = false, unambig = false, containment = false, nonsplice = false, quiet = false;
  for (int i = 5; i < argc; i++) {
    std::string a = argv[i];
    if (a == "--stranded=forward") { stranded = true; strandDir = 1; } else if (a == "--stranded=backward") { stranded = true; strandDir = 2; }
    else if (a == "--unweighted") unweighted = true; else if (a == "--unambig") unambig = true; else if (a == "--containment") containment = true;
    else if (a == "--nonsplice") nonsplice = true; else if (a == "--quiet") quiet = true;
  }
  try {
    const std::string base = anno + "/ensembl_rna_hs";
    std::string G = read_file(samPath), bFile = read_file(base + "_genome_exons.bed");
    FeatureTable ftGene, ftExon; ftGene.load(read_file(base + "_exon_gene.map")); ftExon.load(read_file(base + "_exon_exon.map"));
    BIndex B(bFile);
    auto t0 = std::chrono::steady_clock::now();
    std::string wG = compute_read_weights_sam(G, nullptr);
    ConvertStats cs;
    std::string bed = convert_sam_bed(G, nullptr, &cs, strandDir);
    std::string inter = intersect_wo(bed, B, stranded);
    Params pg; pg.mode = MODE_GENE; pg.stranded = stranded; pg.useReadWeights = !unweighted; pg.excludeAmbiguous = unambig; pg.containment = containment; pg.spliceConforming = !nonsplice;
    Params pe = pg; pe.mode = MODE_EXON; pe.overlapThreshold = 5; pe.containment = false;
    CountResult rg = compute_counts(inter, wG, ftGene, pg), re = compute_counts(inter, wG, ftExon, pe);
    double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    write_file(outd + "/genomic-sam.wgt", wG); write_file(outd + "/combined.bed", bed); write_file(outd + "/intersection.bed", inter);
    write_file(outd + "/gene.cnt", format_counts(rg, ftGene, pg)); write_file(outd + "/exon.cnt", format_counts(re, ftExon, pe)); write_file(outd + "/junction.cnt", "");
    char buf[512];
    snprintf(buf, sizeof buf, "NUMBER_MAPPED_READS=%ld\nNUMBER_EXPRESSED_READS=%ld\nNUM_WEIGHT_OBJECTS=%ld\nORACLE_SECONDS=%.6f\nSUM_WEIGHTS_GENE=%.4f\nSUM_WEIGHTS_EXON=%.4f\nSUM_WEIGHTS_JUNCTION=0.0000\n",
             cs.mappedFragments, rg.stats.included, rg.stats.numWeightObjects, secs, rg.stats.sumObjectWeight, re.stats.sumObjectWeight);
    write_file(outd + "/stats.txt", buf);
    if (!quiet) fputs(buf, stdout);
    return 0;
  } catch (const std::exception& e) { fprintf(stderr, "ERROR: %s\n", e.what()); return 1; }
}

int main(int argc, char** argv) {
  if (argc >= 5 && !strcmp(argv[1], "genomic")) return run_genomic(argc, argv);
  if (argc < 6 || strcmp(argv[1], "run")) {
    fprintf(stderr, "usage: eqp_oracle genomic <annodir> <sam> <outdir> [--stranded=forward|backward] [--unweighted] [--unambig] [--containment] [--nonsplice] [--quiet]\n");
    fprintf(stderr, "usage: eqp_oracle run <annodir> <chunkdir> <se|pe> <outdir> [--stranded=forward|backward] [--unweighted] [--unambig] [--containment] [--nonsplice] [--no-adjust-shortcut] [--quiet]\n");
    return 2;
  }
  std::string anno = argv[2], chunk = argv[3], lay = argv[4], outd = argv[5];
  int strandDir = 0; bool stranded = false, unweighted = false, unambig = false, containment = false, nonsplice = false, noShortcut = false, quiet = false;
  for (int i = 6; i < argc; i++) {
    std::string a = argv[i];
    if (a == "--stranded=forward") { stranded = true; strandDir = 1; } else if (a == "--stranded=backward") { stranded = true; strandDir = 2; } else if (a == "--unweighted") unweighted = true; else if (a == "--unambig") unambig = true;
    else if (a == "--containment") containment = true; else if (a == "--nonsplice") nonsplice = true;
    else if (a == "--no-adjust-shortcut") noShortcut = true; else if (a == "--quiet") quiet = true;
  }
  try {
    const std::string sfx = lay == "pe" ? "pe" : "sr", base = anno + "/ensembl_rna_hs";
    std::string T = read_file(chunk + "/transcript-" + sfx + ".sam"), J = read_file(chunk + "/junction-" + sfx + ".sam"), G = read_file(chunk + "/genome-" + sfx + ".sam");
    std::string bFile = read_file(base + "_100_combined_exons.bed"), tBed = read_file(base + "_transcript_exons.bed");
    std::unordered_set<std::string> white;
    { std::string fai = read_file(anno + "/genome.fa.fai"); Lines L(fai); sv line; while (L.next(line)) { auto t = tokenize(line, '\t'); if (!t.empty()) white.insert(std::string(t[0])); } }
    FeatureTable ftGene, ftExon, ftJunction;
    ftGene.load(read_file(base + "_exon_gene.map")); ftExon.load(read_file(base + "_exon_exon.map"));
    VersionTable versions = load_transcript_versions(tBed);
    BIndex B(bFile);
    auto t0 = std::chrono::steady_clock::now();
    std::string wT = compute_read_weights_sam(T, nullptr), wJ = compute_read_weights_sam(J, nullptr), wG = compute_read_weights_sam(G, &white);
    std::string wTJ = combine_read_weight_files(wT, wJ, nullptr, true);
    std::string samTJ = combine_sam_files(T, J, wT, wJ);
    std::string samC = combine_sam_files(samTJ, G, wTJ, wG);
    ConvertStats cs;
    std::string bed = convert_sam_bed(samC, &versions, &cs, strandDir);       // ConvertSamBed -S <direction> (CIB:366,912)
    std::string inter = intersect_wo(bed, B, stranded);
    std::string wBed = compute_read_weights(inter);
    std::string wFinal = combine_read_weight_files(wBed, wG, &wTJ, false);
    std::string refIds = cut_f3_sort_u({samC, J});
    std::string fmap = filter_junction_map(refIds, read_file(base + "_exon_junction.map"), read_file(base + "_exon_junction_transcript.map"),
                                           read_file(base + "_100_junction_junction_coordinate.map"), read_file(base + "_equivalent_junctions.map"));
    ftJunction.load(fmap);
    Params pg; pg.mode = MODE_GENE; pg.stranded = stranded; pg.useReadWeights = !unweighted; pg.excludeAmbiguous = unambig; pg.containment = containment;
    pg.spliceConforming = !nonsplice; pg.noAdjustShortcut = noShortcut;
    Params pe = pg; pe.mode = MODE_EXON; pe.overlapThreshold = 5; pe.containment = false;                 // CC:959-960
    Params pj = pg; pj.mode = MODE_JUNCTION; pj.overlapThreshold = 8; pj.outputZeroes = false; pj.excludeAmbiguous = false;
    pj.containment = false; pj.spliceConforming = true;                                                  // CC:1034-1035
    CountResult rg = compute_counts(inter, wFinal, ftGene, pg), re = compute_counts(inter, wFinal, ftExon, pe), rj = compute_counts(inter, wFinal, ftJunction, pj);
    std::string cg = format_counts(rg, ftGene, pg), ce = format_counts(re, ftExon, pe), cj = format_counts(rj, ftJunction, pj);
    double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    write_file(outd + "/transcript-sam.wgt", wT); write_file(outd + "/junction-sam.wgt", wJ); write_file(outd + "/genome-sam.wgt", wG);
    write_file(outd + "/trans-junction-sam.wgt", wTJ); write_file(outd + "/combined.sam", samC); write_file(outd + "/combined.bed", bed);
    write_file(outd + "/intersection.bed", inter); write_file(outd + "/bed.wgt", wBed); write_file(outd + "/final.wgt", wFinal);
    write_file(outd + "/reference.ids", refIds); write_file(outd + "/exon_junction_filtered.map", fmap);
    write_file(outd + "/gene.cnt", cg); write_file(outd + "/exon.cnt", ce); write_file(outd + "/junction.cnt", cj);
    char buf[512];
    snprintf(buf, sizeof buf, "NUMBER_MAPPED_READS=%ld\nNUMBER_EXPRESSED_READS=%ld\nNUM_WEIGHT_OBJECTS=%ld\nORACLE_SECONDS=%.6f\nSHAPE_SNAPS=%ld\nSUM_WEIGHTS_GENE=%.4f\nSUM_WEIGHTS_EXON=%.4f\nSUM_WEIGHTS_JUNCTION=%.4f\n",
             cs.mappedFragments, rg.stats.included, rg.stats.numWeightObjects, secs, shape_counters().snaps, rg.stats.sumObjectWeight, re.stats.sumObjectWeight, rj.stats.sumObjectWeight);
    write_file(outd + "/stats.txt", buf);
    if (!quiet) fputs(buf, stdout);
    return 0;
  } catch (const std::exception& e) {
    fprintf(stderr, "ERROR: %s\n", e.what());
    return 1;
  }
}
