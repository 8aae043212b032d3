// Development tool: times the line index + SAM decode kernels alone on one SAM file (not part of the product or the tests).
// usage: decode_bench <annodir> <sam> [repeats]
#include <cstdio>
#include "../eqp-cluster_b200/csrc/eqp_cuda_backend.cuh"
using namespace eqpb;
int main(int argc, char** argv) {
  if (argc < 3) return 2;
  std::string anno = argv[1], base = anno + "/ensembl_rna_hs"; int rep = argc > 3 ? atoi(argv[3]) : 5;
  AnnoHost H;
  H.load((base + "_100_combined_exons.bed").c_str(), (base + "_transcript_exons.bed").c_str(), (base + "_exon_gene.map").c_str(), (base + "_exon_exon.map").c_str(),
         (base + "_exon_junction.map").c_str(), (base + "_exon_junction_transcript.map").c_str(), (base + "_100_junction_junction_coordinate.map").c_str(),
         (base + "_equivalent_junctions.map").c_str(), (anno + "/genome.fa.fai").c_str());
  std::string T = AnnoHost::slurp(argv[2]);
  cudaStream_t st; cudaStreamCreate(&st);
  AnnoView A = H.view();
  DictEntry* dDict; char* dPool; cudaMalloc(&dDict, H.dict.size() * sizeof(DictEntry)); cudaMemcpy(dDict, H.dict.data(), H.dict.size() * sizeof(DictEntry), cudaMemcpyHostToDevice);
  cudaMalloc(&dPool, H.strPool.size() + 16); cudaMemcpy(dPool, H.strPool.data(), H.strPool.size(), cudaMemcpyHostToDevice);
  A.dict = dDict; A.strPool = dPool;
  char* dText; cudaMalloc(&dText, T.size() + 16); cudaMemcpy(dText, T.data(), T.size(), cudaMemcpyHostToDevice);
  CudaBackend be(st, 148); RunParams P{}; Pipeline<CudaBackend> pl(be, A, P);
  uint32_t* err = be.alloc<uint32_t>(32); be.memset(err, 0, 128);
  pl.index_lines(0, dText, T.size());
  for (int r = 0; r < rep; r++) {
    be.decodeMs = 0; be.decode(A, pl.C.sam[0], err);
    printf("decode %d: %.3f ms  %.1f GB/s  (%u lines, %zu bytes)\n", r, be.decodeMs, T.size() / be.decodeMs / 1e6, pl.C.sam[0].nLines, T.size());
  }
  uint32_t h[2]; be.to_host(h, err, 8); printf("err %u %u\n", h[0], h[1]);
  return 0;
}
