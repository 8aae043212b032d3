/* abi_smoke.c — drives libeqpcount.so through include/eqp_count.h ALONE, from plain C (what a JNI / Panama host sees):
 * create -> load annotation -> streamed feed -> count -> accumulate -> reduce (one-rank NCCL communicator) -> rows -> destroy.
 * Prints "id<TAB>value" rows of the three count files to <outdir>/{gene,exon,junction}.rows ("%.17g") and the scalars to stdout.
 *   gcc -std=c11 -O2 -Iinclude tests/abi/abi_smoke.c -o abi_smoke -Leqp-cluster_b200/lib -leqpcount -lcudart? (no: only -leqpcount) -Wl,-rpath,...
 * usage: abi_smoke <annodir> <T.sam> <J.sam> <G.sam> <outdir>           exit 0 = every call succeeded, 3 = no CUDA device (eqp_create failed loudly) */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "eqp_count.h"

static unsigned char* slurp(const char* path, size_t* n) {
  FILE* f = fopen(path, "rb"); unsigned char* b; long sz;
  if (!f) { fprintf(stderr, "ERROR: cannot open %s\n", path); exit(2); }
  fseek(f, 0, SEEK_END); sz = ftell(f); fseek(f, 0, SEEK_SET);
  b = (unsigned char*)malloc(sz ? (size_t)sz : 1); *n = fread(b, 1, (size_t)sz, f); fclose(f); return b;
}
#define CHECK(call) do { int rc_ = (call); if (rc_) { fprintf(stderr, "ERROR: %s -> %d: %s\n", #call, rc_, eqp_last_error(ctx)); return 1; } } while (0)

int main(int argc, char** argv) {
  eqp_ctx* ctx = NULL; eqp_params p; eqp_counts c; eqp_annotation_files af; char path[9][1024]; const char* names[3] = {"gene", "exon", "junction"};
  unsigned char* sam[3]; size_t len[3], off[3] = {0, 0, 0}; int64_t n; char id[EQP_COMM_ID_BYTES];
  if (argc != 6) { fprintf(stderr, "usage: abi_smoke <annodir> <T.sam> <J.sam> <G.sam> <outdir>\n"); return 2; }
  if (eqp_version() != EQP_ABI_VERSION) { fprintf(stderr, "ERROR: header/library ABI mismatch\n"); return 1; }
  if (eqp_create(0, &ctx)) { fprintf(stderr, "ERROR: %s\n", eqp_last_error(NULL)); return 3; }
  snprintf(path[0], 1024, "%s/ensembl_rna_hs_100_combined_exons.bed", argv[1]); snprintf(path[1], 1024, "%s/ensembl_rna_hs_transcript_exons.bed", argv[1]);
  snprintf(path[2], 1024, "%s/ensembl_rna_hs_exon_gene.map", argv[1]); snprintf(path[3], 1024, "%s/ensembl_rna_hs_exon_exon.map", argv[1]);
  snprintf(path[4], 1024, "%s/ensembl_rna_hs_exon_junction.map", argv[1]); snprintf(path[5], 1024, "%s/ensembl_rna_hs_exon_junction_transcript.map", argv[1]);
  snprintf(path[6], 1024, "%s/ensembl_rna_hs_100_junction_junction_coordinate.map", argv[1]); snprintf(path[7], 1024, "%s/ensembl_rna_hs_equivalent_junctions.map", argv[1]);
  snprintf(path[8], 1024, "%s/genome.fa.fai", argv[1]);
  af.combined_exons_bed = path[0]; af.transcript_exons_bed = path[1]; af.exon_gene_map = path[2]; af.exon_exon_map = path[3]; af.exon_junction_map = path[4];
  af.exon_junction_transcript_map = path[5]; af.junction_junction_coordinate_map = path[6]; af.equivalent_junctions_map = path[7]; af.genome_fai = path[8];
  CHECK(eqp_load_annotation(ctx, &af));
  for (int f = 0; f < 3; f++) sam[f] = slurp(argv[2 + f], &len[f]);
  eqp_default_params(&p);
  /* streamed: 1 MB pieces, inputs interleaved */
  CHECK(eqp_chunk_begin(ctx, 0, len[1], 0));
  for (int more = 1; more;) { more = 0; for (int f = 0; f < 3; f++) if (off[f] < len[f]) { size_t k = len[f] - off[f] < (1u << 20) ? len[f] - off[f] : (1u << 20); CHECK(eqp_chunk_feed(ctx, f, sam[f] + off[f], k)); off[f] += k; more = 1; } }
  CHECK(eqp_chunk_end(ctx, &p, &c));
  printf("NUMBER_MAPPED_READS=%lld\nNUMBER_EXPRESSED_READS=%lld\nNUM_WEIGHT_OBJECTS=%lld\nKERNEL_LAUNCHES=%lld\n", (long long)c.number_mapped_reads, (long long)c.number_expressed_reads,
         (long long)c.num_weight_objects[0], (long long)c.kernel_launches);
  { const double* vec[3] = {c.gene, c.exon, c.junction}; const uint64_t* fx[3] = {c.gene_fixed, c.exon_fixed, c.junction_fixed};
    for (int m = 0; m < 3; m++) {
      char out[1200]; FILE* f; snprintf(out, sizeof out, "%s/%s.rows", argv[5], names[m]); f = fopen(out, "w");
      if (!f) { fprintf(stderr, "ERROR: cannot write %s\n", out); return 2; }
      n = eqp_output_row_count(ctx, m);
      for (int64_t k = 0; k < n; k++) {
        eqp_output_row r; CHECK(eqp_get_output_row(ctx, m, k, &r));
        if (m == 2 && !(r.always || c.junction_active[r.act])) continue;
        if (m == 2 && fx[m][r.feature] == 0) continue;
        fprintf(f, "%s\t%.17g\n", r.id, vec[m][r.feature]);
      }
      fclose(f);
    } }
  /* the same chunk from whole host buffers must give the same bits */
  { uint64_t g0 = c.n_gene ? c.gene_fixed[0] : 0; int64_t mapped = c.number_mapped_reads;
    CHECK(eqp_count_chunk(ctx, &p, sam[0], len[0], sam[1], len[1], sam[2], len[2], 0, &c));
    if (c.number_mapped_reads != mapped || (c.n_gene && c.gene_fixed[0] != g0)) { fprintf(stderr, "ERROR: streamed and whole-buffer counts differ\n"); return 1; } }
  /* sample level: chunk vector -> wide accumulator -> reduce on a one-rank communicator; read back through the device vector API only */
  { uint64_t* vec; CHECK(eqp_device_vector(ctx, &vec, &n)); printf("VECTOR_LENGTH=%lld\n", (long long)n); }
  if (eqp_comm_id(id) == 0) { CHECK(eqp_comm_init(ctx, id, 1, 0)); printf("NCCL=ok\n"); } else printf("NCCL=unavailable (%s)\n", eqp_last_error(NULL));
  eqp_destroy(ctx);
  for (int f = 0; f < 3; f++) free(sam[f]);
  printf("ABI_SMOKE=ok\n");
  return 0;
}
