// Chunk pipeline: the order of stages that replaces compute-counts.sh:733-1037 for one chunk.
// Templated on a backend that provides alloc/memset/launch/scan, so that the identical per-item device functions
// run (a) as CUDA kernels in libeqpcount.so and (b) in the host emulation harness under tests/emu (test-only).
#pragma once
#include <string>
#include <vector>
#include "eqp_accumulate.h"

#if defined(__CUDACC__)
#define EQP_LAMBDA __host__ __device__
#else
#define EQP_LAMBDA
#endif

namespace eqpb {

// Single-address reductions: the active lanes that pass the same p (grouped with match.any: lanes of different loop iterations may
// sit on the same instruction) combine in the warp and one lane issues the atomic.  Same-address atomics serialise in L2 (~0.85 clk per lane): 10^8 of them cost as much as a whole stage.
EQP_HD void atomic_min_u32(uint32_t* p, uint32_t v) {
#if defined(__CUDA_ARCH__)
  uint32_t m = __match_any_sync(__activemask(), (unsigned long long)p), r = __reduce_min_sync(m, v);
  if ((threadIdx.x & 31) == (uint32_t)(__ffs((int)m) - 1)) atomicMin(p, r);
#else
  if (v < *p) *p = v;
#endif
}
EQP_HD void atomic_max_u32(uint32_t* p, uint32_t v) {
#if defined(__CUDA_ARCH__)
  uint32_t m = __match_any_sync(__activemask(), (unsigned long long)p), r = __reduce_max_sync(m, v);
  if ((threadIdx.x & 31) == (uint32_t)(__ffs((int)m) - 1)) atomicMax(p, r);
#else
  if (v > *p) *p = v;
#endif
}
EQP_HD void counter_add(unsigned long long* p, uint32_t v) {
#if defined(__CUDA_ARCH__)
  uint32_t m = __match_any_sync(__activemask(), (unsigned long long)p), r = __reduce_add_sync(m, v);
  if (r && (threadIdx.x & 31) == (uint32_t)(__ffs((int)m) - 1)) atomicAdd(p, (unsigned long long)r);
#else
  *p += v;
#endif
}

EQP_HD uint32_t atomic_inc_u32(uint32_t* p) {
#if defined(__CUDA_ARCH__)
  return atomicAdd(p, 1u);
#else
  return (*p)++;
#endif
}
// append to a work list (order irrelevant); rare events only
EQP_HD void list_push(uint32_t* count, uint32_t* list, uint32_t v) {
#if defined(__CUDA_ARCH__)
  list[atomicAdd(count, 1u)] = v;
#else
  list[(*count)++] = v;
#endif
}

EQP_HD void counter_add64(unsigned long long* p, uint64_t v) {          // v < 2^42: two pieces that cannot wrap in a 32-lane sum
#if defined(__CUDA_ARCH__)
  uint32_t m = __match_any_sync(__activemask(), (unsigned long long)p);
  uint32_t r0 = __reduce_add_sync(m, (uint32_t)(v & 0x3FFFFFu)), r1 = __reduce_add_sync(m, (uint32_t)(v >> 22));
  unsigned long long r = (unsigned long long)r0 + ((unsigned long long)r1 << 22);
  if (r && (threadIdx.x & 31) == (uint32_t)(__ffs((int)m) - 1)) atomicAdd(p, r);
#else
  *p += v;
#endif
}

// stage tags: they only name the kernels (profiler listings)
namespace st { struct misc; struct frag_table; struct frag_stats; struct select_tj_weights; struct select_stage1; struct select_stage2; struct seen_refs; struct junction_active;
               struct blocks_order; struct blocks_count; struct blocks_fill; struct join_count; struct join_fill; struct work_class; struct alignment_build; struct work_lists; struct fragment_adjust;
               struct fragment_weight; struct fragment_weight_gene; struct fragment_count; struct fragment_count_large; }

struct ChunkResult {
  std::vector<double> counts[3]; std::vector<unsigned long long> fixed[3]; std::vector<uint8_t> junctionActive;
  unsigned long long stats[16]; uint32_t nRecords[3]; uint32_t err[2];
};

static const uint32_t LINE_TILE = 1024;               // bytes per newline-count item

template <class BE>
struct Pipeline {
  BE& be; AnnoView A; RunParams P;
  ChunkView C{}; CountView V{};
  uint32_t nBlocks = 0, nAln = 0, nHits = 0;
  uint32_t* scal = nullptr;                           // device scalars: [0..1] err, [2..7] min/max frag per file, [8..] presence bounds
  uint32_t* tileCnt[3] = {nullptr, nullptr, nullptr};
  unsigned long long* accOut[3] = {nullptr, nullptr, nullptr};   // caller-owned accumulators (nFeat[m] + 1 entries); allocated here when null

  Pipeline(BE& b, const AnnoView& a, const RunParams& p) : be(b), A(a), P(p) {}

  // line index of the host emulation (tests/emu); the CUDA backend builds it inside the tokenizer kernel
  void index_lines(int f, const char* text, uint64_t len) {
    SamView& S = C.sam[f]; S.text = text; S.len = len;
    uint64_t nTiles = (len + LINE_TILE - 1) / LINE_TILE;
    uint32_t* cnt = be.template alloc<uint32_t>(nTiles + 1); tileCnt[f] = cnt;
    be.count_newlines(text, len, nTiles, cnt);
    be.scan(cnt, nTiles);
    uint32_t nl = 0; be.to_host(&nl, cnt + nTiles, sizeof nl);
    char last = 0; if (len) be.to_host(&last, text + len - 1, 1);
    S.nLines = len == 0 ? 0 : (last == '\n' ? nl : nl + 1);
    S.lineOff = be.template alloc<uint64_t>((uint64_t)S.nLines + 1);
    uint64_t* lo = S.lineOff; uint32_t nLines = S.nLines;
    be.fill_line_offsets(text, len, nTiles, cnt, lo, nLines, last);
    uint64_t n = S.nLines;
    S.frag = be.template alloc<uint32_t>(n); S.ref = be.template alloc<int32_t>(n); S.pos = be.template alloc<int32_t>(n);
    S.rnext = be.template alloc<int32_t>(n); S.pnext = be.template alloc<int32_t>(n); S.flag = be.template alloc<uint16_t>(n);
    S.nm = be.template alloc<int16_t>(n); S.cigOff = be.template alloc<uint16_t>(n); S.cigLen = be.template alloc<uint16_t>(n);
    S.seqLen = be.template alloc<uint16_t>(n); S.cigM = be.template alloc<uint16_t>(n); S.rf = be.template alloc<uint8_t>(n); C.proc[f] = be.template alloc<uint8_t>(n);
    S.hi = P.genomic ? be.template alloc<int32_t>(n) : nullptr;
  }

  int run(const char* const text[3], const uint64_t len[3], ChunkResult& R) {
    be.stage(nullptr);
    scal = be.template alloc<uint32_t>(128); be.memset(scal, 0, 128 * sizeof(uint32_t));
    C.err = scal;
    const AnnoView A = this->A; const RunParams P = this->P;
    // ---- a1/a2: line index + decode (+ the fragment id range of each input)
    { uint32_t init[6] = {0xFFFFFFFFu, 0, 0xFFFFFFFFu, 0, 0xFFFFFFFFu, 0}; be.to_device(scal + 2, init, sizeof init); }
    for (int f = 0; f < 3; f++) {
      uint32_t* mm = scal + 2 + 2 * f;
      if constexpr (BE::kFusedDecode) be.decode_fused(A, C.sam[f], C.proc[f], text[f], len[f], scal, mm);   // CUDA: one pass, line index + records + id range
      else {                                                             // host emulation (tests/emu): line index, then the per-line decoder
        index_lines(f, text[f], len[f]); be.stage("line-index"); be.decode(A, C.sam[f], scal);
        const SamView S = C.sam[f];
        be.launch(S.nLines, [=] EQP_LAMBDA(uint64_t i) { if (S.rf[i] & RF_REC) { atomic_min_u32(mm, S.frag[i]); atomic_max_u32(mm + 1, S.frag[i]); } });
      }
      be.stage("decode");
      R.nRecords[f] = C.sam[f].nLines;
    }
    uint32_t h[8]; be.to_host(h, scal, sizeof h);
    if (h[0]) { R.err[0] = h[0]; R.err[1] = h[1]; return -1; }
    uint32_t lo = 0xFFFFFFFFu, hi = 0; bool anyRec = false;
    for (int f = 0; f < 3; f++) if (h[2 + 2 * f] != 0xFFFFFFFFu) { anyRec = true; lo = std::min(lo, h[2 + 2 * f]); hi = std::max(hi, h[3 + 2 * f]); }
    C.base = anyRec ? lo : 0; C.nSlots = anyRec ? hi - lo + 1 : 0;
    uint64_t totalLines = (uint64_t)C.sam[0].nLines + C.sam[1].nLines + C.sam[2].nLines;
    if ((uint64_t)C.nSlots > 8 * totalLines + 1024) { R.err[0] = DERR_CAPACITY; R.err[1] = 0; return -1; }
    const uint32_t N = C.nSlots;
    for (int f = 0; f < 3; f++) { C.ff[f] = be.template alloc<FragFile>(N); be.memset(C.ff[f], 0, (uint64_t)N * sizeof(FragFile)); C.st[f] = be.template alloc<FragStat>(N); }
    C.stTJ = be.template alloc<FragStat>(N); C.setTJ = be.template alloc<SetDesc>(N); C.setC = be.template alloc<SetDesc>(N);
    C.auxSlot = be.template alloc<uint32_t>((uint64_t)N + 1); C.blkOff = be.template alloc<uint32_t>((uint64_t)N + 1); C.alnOff = be.template alloc<uint32_t>((uint64_t)N + 1);
    C.seenRef = be.template alloc<uint8_t>((uint64_t)A.nRef + 1); be.memset(C.seenRef, 0, (uint64_t)A.nRef + 1);
    C.slotTr = be.template alloc<uint8_t>((uint64_t)N + 1); C.slotLd = be.template alloc<uint8_t>((uint64_t)N + 1);
    be.memset(C.slotTr, 0, (uint64_t)N + 1); be.memset(C.slotLd, 0, (uint64_t)N + 1);
    // one fragment adds at most 1.0 = 2^38 units to a feature, so a u64 accumulator holds 2^26 fragments' worth (include/eqp_count.h)
    if ((uint64_t)N >= (1ull << 26)) { R.err[0] = DERR_ACC_OVERFLOW; R.err[1] = N; return -1; }
    be.stage("frag-range");
    // ---- fragment table: [first,end) line range per (file, slot); order check (CombineSamFiles.java:191-193)
    for (int f = 0; f < 3; f++) { const SamView S = C.sam[f]; FragFile* ff = C.ff[f]; uint32_t base = C.base; uint32_t* err = scal;
      be.template launch<st::frag_table>(S.nLines, [=] EQP_LAMBDA(uint64_t i) {
        if (!(S.rf[i] & RF_REC)) return;
        uint32_t fr = S.frag[i];
        uint64_t pi = i, ni = i + 1;                                     // nearest record lines around i (header / empty lines skipped)
        while (pi > 0 && !(S.rf[pi - 1] & RF_REC)) pi--;
        while (ni < S.nLines && !(S.rf[ni] & RF_REC)) ni++;
        bool prevRec = pi > 0, nextRec = ni < S.nLines;
        if (prevRec && S.frag[pi - 1] > fr) raise(err, DERR_ORDER, fr);
        if (!prevRec || S.frag[pi - 1] != fr) ff[fr - base].first = (uint32_t)i;
        if (!nextRec || S.frag[ni] != fr) ff[fr - base].end = (uint32_t)i + 1;
      }); }
    be.stage("frag-table");
    // ---- a3/a4: per-fragment statistics in each input
    for (int f = 0; f < 3; f++) be.memset(C.proc[f], 0, C.sam[f].nLines);
    { const ChunkView C = this->C;
      be.template launch<st::frag_stats>((uint64_t)N * 3, [=] EQP_LAMBDA(uint64_t k) {
        int f = (int)(k / N); uint32_t s = (uint32_t)(k % N); FragFile ff = C.ff[f][s]; FragStat st{};
        fragment_stats(A, C.sam[f], ff.first, ff.end, f == FILE_G, C.proc[f], st, ff, C.err, C.base + s);
        C.st[f][s] = st; C.ff[f][s].bits = ff.bits;
      }); }
    be.stage("frag-stats");
    // ---- a5-a7: T+J weights, stage-1 selection, presence bounds, stage-2 selection, aux lookup table
    // presence bounds of the inputs (first / last slot with a record, any record): known from the fragment id range of each file;
    // those of the stage-1 output are reduced on the device, in a cache line of their own (nothing else reads or writes it meanwhile)
    const bool anyT = h[2] != 0xFFFFFFFFu, anyJ = h[4] != 0xFFFFFFFFu, anyG = h[6] != 0xFFFFFFFFu;
    const uint32_t firstT = anyT ? h[2] - C.base : 0xFFFFFFFFu, lastT = anyT ? h[3] - C.base : 0, firstJ = anyJ ? h[4] - C.base : 0xFFFFFFFFu, firstG = anyG ? h[6] - C.base : 0xFFFFFFFFu;
    uint32_t* bnd = scal + 64;                         // [0] firstTJ [1] lastTJ [2] anyTJ
    { uint32_t init[3] = {0xFFFFFFFFu, 0, 0}; be.to_device(bnd, init, sizeof init); }
    { const ChunkView C = this->C;
      be.template launch<st::select_tj_weights>(N, [=] EQP_LAMBDA(uint64_t s) {
        C.stTJ[s] = combine_weights_ed(C.st[FILE_T][s], C.st[FILE_J][s], P.slack);
      });
      be.template launch<st::select_stage1>(N, [=] EQP_LAMBDA(uint64_t s) {
        Side a = side_of_file(FILE_T, C.ff[FILE_T][s], C.st[FILE_T][s]), b = side_of_file(FILE_J, C.ff[FILE_J][s], C.st[FILE_J][s]);
        SetDesc r = combine_stage(a, b, (uint32_t)s, firstT, firstJ, lastT, anyT, anyJ, P.slack, C.err, C.base + (uint32_t)s);
        C.setTJ[s] = r;
        if (sd_n(r)) { atomic_min_u32(bnd + 0, (uint32_t)s); atomic_max_u32(bnd + 1, (uint32_t)s); bnd[2] = 1; }
      });
      be.template launch<st::select_stage2>(N, [=] EQP_LAMBDA(uint64_t s) {
        FragFile ffs[3] = {C.ff[0][s], C.ff[1][s], C.ff[2][s]};
        Side a = side_of_set(C.setTJ[s], ffs, C.stTJ[s]), b = side_of_file(FILE_G, C.ff[FILE_G][s], C.st[FILE_G][s]);
        C.setC[s] = combine_stage(a, b, (uint32_t)s, bnd[0], firstG, bnd[1], bnd[2] != 0, anyG, P.slack, C.err, C.base + (uint32_t)s);
      });
      // aux: first slot >= s with a T+J weight entry (FragmentEntryFactory.getNextFragmentEntry(query), :75-106)
      const uint32_t CH = 512; uint64_t nCh = ((uint64_t)N + CH - 1) / CH;
      uint32_t* chNext = be.template alloc<uint32_t>(nCh + 1);
      be.launch(nCh, [=] EQP_LAMBDA(uint64_t c) {
        uint64_t b = c * CH, e = b + CH < N ? b + CH : N; uint32_t nx = 0xFFFFFFFFu;
        for (uint64_t s = e; s-- > b;) { if (C.stTJ[s].flags & FS_EXISTS) nx = (uint32_t)s; C.auxSlot[s] = nx; }
        chNext[c] = nx;
      });
      be.launch(1, [=] EQP_LAMBDA(uint64_t) { uint32_t nx = 0xFFFFFFFFu; for (uint64_t c = nCh; c-- > 0;) { uint32_t own = chNext[c]; chNext[c] = nx; if (own != 0xFFFFFFFFu) nx = own; } });
      be.launch(N, [=] EQP_LAMBDA(uint64_t s) { if (C.auxSlot[s] == 0xFFFFFFFFu) C.auxSlot[s] = chNext[s / CH]; });
    }
    be.stage("select");
    // ---- a27: reference ids seen (combined SAM + junction SAM) -> active junction features
    V.junctionActive = be.template alloc<uint8_t>((uint64_t)A.nAct + 1); be.memset((void*)V.junctionActive, 0, (uint64_t)A.nAct + 1);
    { const ChunkView C = this->C; uint8_t* act = (uint8_t*)V.junctionActive; const SamView J = C.sam[FILE_J];
      be.launch(J.nLines, [=] EQP_LAMBDA(uint64_t i) { if ((J.rf[i] & RF_REC) && J.ref[i] >= 0) C.seenRef[J.ref[i]] = 1; });
      be.template launch<st::seen_refs>(N, [=] EQP_LAMBDA(uint64_t s) {
        SetDesc sd = C.setC[s];
        for (int k = 0; k < sd_n(sd); k++) { int f = sd_file(sd, k), flt = sd_filter(sd, k); FragFile ff = C.ff[f][s];
          for (uint32_t i = ff.first; i < ff.end; i++) { uint8_t rf = C.sam[f].rf[i]; if ((rf & RF_REC) && (!flt || rf_read_index(rf) == flt) && C.sam[f].ref[i] >= 0) C.seenRef[C.sam[f].ref[i]] = 1; } }
      });
      be.template launch<st::junction_active>((uint64_t)A.nRef, [=] EQP_LAMBDA(uint64_t r) {
        if (!C.seenRef[r]) return;
        if (A.refFlags[r] & 2) raise(C.err, DERR_BAD_JUNCTION_ID, (uint32_t)r);
        for (uint32_t k = A.refActOff[r]; k < A.refActOff[r + 1]; k++) act[A.refAct[k]] = 1;
      }); }
    be.stage("junction-active");
    // ---- a8-a10: blocks (count, scan, fill)
    V.stats = be.template alloc<unsigned long long>(16); be.memset(V.stats, 0, 16 * sizeof(unsigned long long));
    for (int f = 0; f < 3; f++) be.memset(C.proc[f], 0, C.sam[f].nLines);
    // The two passes walk a fragment's record set in one thread; record counts have a long tail (multi-mappers), and in slot order a warp runs at
    // 4.7 active lanes of 32 (ncu, profiles/r03d).  Visiting the fragments in descending order of their record count (EQP_VAR_BLOCKS_SORTED) fixes the
    // lanes but scatters the reads of the record arrays: measured 25.5 ms against 21.0 ms in slot order (profiles/r03h), so slot order stays; the
    // list below only drops the fragments without records.
    uint32_t* bkey = be.template alloc<uint32_t>((uint64_t)N + 1); uint32_t* bperm = be.template alloc<uint32_t>((uint64_t)N + 1);
    be.memset(C.blkOff, 0, ((uint64_t)N + 1) * 4); be.memset(C.alnOff, 0, ((uint64_t)N + 1) * 4);
    { const ChunkView C = this->C;
      be.template launch<st::blocks_order>(N, [=] EQP_LAMBDA(uint64_t s) {
        SetDesc sd = C.setC[s]; uint32_t n = 0;
        for (int k = 0; k < sd_n(sd); k++) { const FragFile ff = C.ff[sd_file(sd, k)][s]; n += ff.end - ff.first; }
#ifdef EQP_VAR_BLOCKS_SORTED
        bkey[s] = n < 255u ? n : 255u;
#else
        bkey[s] = n ? 1u : 0u;
#endif
      }); }
    const uint32_t nB = be.order_by_work(bkey, N, bperm);
    { const ChunkView C = this->C; unsigned long long* stats = V.stats;
      be.template launch<st::blocks_count>(nB, [=] EQP_LAMBDA(uint64_t i) { const uint32_t s = bperm[i]; uint32_t nb, na; fragment_blocks(A, C, s, false, P.stranded, nb, na); C.blkOff[s] = nb; C.alnOff[s] = na;
        SetCursor c{0, 0xFFFFFFFFu}; int f; counter_add(&stats[9], set_next(C, C.setC[s], s, c, f) ? 1u : 0u); }); }
    be.scan(C.blkOff, N); be.scan(C.alnOff, N);
    be.to_host(&nBlocks, C.blkOff + N, 4); be.to_host(&nAln, C.alnOff + N, 4);
    C.blocks = be.template alloc<Block>((uint64_t)nBlocks + 1); C.hitCnt = be.template alloc<uint32_t>((uint64_t)nBlocks + 1);
    C.alnBlk = be.template alloc<uint32_t>((uint64_t)nAln + 1); be.to_device(C.alnBlk + nAln, &nBlocks, 4);
    for (int f = 0; f < 3; f++) be.memset(C.proc[f], 0, C.sam[f].nLines);
    { const ChunkView C = this->C; be.template launch_occ<st::blocks_fill, 6>(nB, [=] EQP_LAMBDA(uint64_t i) { uint32_t nb, na; fragment_blocks(A, C, bperm[i], true, P.stranded, nb, na); }); }
    be.stage("blocks");
    // ---- a11: interval join (count, scan, fill)
    { const ChunkView C = this->C; be.template launch<st::join_count>(nBlocks, [=] EQP_LAMBDA(uint64_t b) { C.hitCnt[b] = join_block(A, C.blocks[b], nullptr, 0, (uint32_t)b, P.stranded != 0); }); }
    be.scan(C.hitCnt, nBlocks); be.to_host(&nHits, C.hitCnt + nBlocks, 4);
    C.hits = be.template alloc<Hit>((uint64_t)nHits + 1);
    { const ChunkView C = this->C; be.template launch<st::join_fill>(nBlocks, [=] EQP_LAMBDA(uint64_t b) { join_block(A, C.blocks[b], C.hits + C.hitCnt[b], C.hitCnt[b + 1] - C.hitCnt[b], (uint32_t)b, P.stranded != 0); }); }
    be.stage("join");
    // ---- a12-a25
    V.hx = be.template alloc<HitX>((uint64_t)nHits + 1); V.alns = be.template alloc<Aln>((uint64_t)nAln + 1);
    V.dedupTab = be.template alloc<uint32_t>(2 * (uint64_t)nAln + 2); be.memset(V.dedupTab, 0, (2 * (uint64_t)nAln + 2) * 4); V.alnTail = be.template alloc<uint32_t>((uint64_t)nAln + 1);
    V.pack = be.template alloc<SlotPack>((uint64_t)N + 1);
    V.ivAdj = be.template alloc<IvAdj>((uint64_t)nHits + 1);                // a17 scratch, touched for fragments with a positive length difference only
    V.alnG = nullptr;
    if (A.nTrChr > 0) V.alnG = be.template alloc<AlnG>((uint64_t)nAln + 1);
    uint32_t* listLd = be.template alloc<uint32_t>((uint64_t)N + 1); uint32_t* listTr = be.template alloc<uint32_t>((uint64_t)N + 1);
    uint32_t* listCnt = scal + 96; be.memset(listCnt, 0, 8);                // [0] fragments with a positive length difference, [1] with tr- exon hits
    for (int m = 0; m < 3; m++) { V.acc[m] = accOut[m] ? accOut[m] : be.template alloc<unsigned long long>((uint64_t)A.nFeat[m] + 1); be.memset(V.acc[m], 0, ((uint64_t)A.nFeat[m] + 1) * 8); }
    be.stage("alloc-a12");
    // fragments with hits, ordered by descending hit-count class (stable within a class): warps of the thread-per-fragment stages get
    // fragments of similar cost while neighbouring lanes still touch neighbouring memory
    uint32_t* work = be.template alloc<uint32_t>((uint64_t)N + 1); uint32_t* perm = be.template alloc<uint32_t>((uint64_t)N + 1);
    // work key: small fragments (at most 15 alignments and 15 hits) are binned by their exact (alignments, hits) pair, so that the lanes of a warp
    // run the same trip counts in the per-fragment loops; larger ones by 1 + floor(log2(hits)).  Bins keep slot order: neighbouring lanes still touch
    // neighbouring memory.  Keys sort descending: the expensive fragments start first.
    { const ChunkView C = this->C; SlotPack* pack = V.pack;
      be.template launch<st::work_class>(N, [=] EQP_LAMBDA(uint64_t s) {       // slot order: the SlotPack of every fragment + its work key
        SlotPack sp; sp.hit0 = C.hitCnt[C.blkOff[s]]; sp.hit1 = C.hitCnt[C.blkOff[s + 1]]; sp.aln0 = C.alnOff[s]; sp.aln1 = C.alnOff[s + 1];
        sp.a20 = sp.nFinal = sp.a20G = 0; sp.g = sp.hit0 != sp.hit1 ? slot_weight_mode(C, P, (uint32_t)s) : 0u;
        pack[s] = sp;
        uint32_t h = sp.hit1 - sp.hit0, na = sp.aln1 - sp.aln0, key = 0;
#ifdef EQP_VAR_ONE_CLASS
        if (h) key = 1;
        if (0)
#elif defined(EQP_VAR_COARSE_KEYS)
        if (h) { uint32_t c = 0, x = h; while (x) { c++; x >>= 1; } key = c; }
        if (0)
#else
        if (h)
#endif
        { if (h <= 15 && na <= 14) key = (na << 4) | h; else { uint32_t c = 0, x = h; while (x) { c++; x >>= 1; } key = 240 + (c < 15 ? c : 15); } }
        work[s] = key; }); }
    uint32_t nWork = be.order_by_work(work, N, perm);
    { const ChunkView C = this->C; const CountView V = this->V;
      be.stage("order");
      be.template launch<st::alignment_build>(nAln, [=] EQP_LAMBDA(uint64_t a) { alignment_build(A, C, V, P, (uint32_t)a); });
      be.stage("alignment-build");
      // the rare shapes (transcript exon longer than its genome exon; tr- exons) go through work lists built on the device
      be.template launch<st::work_lists>(N, [=] EQP_LAMBDA(uint64_t s) {      // slot order (coalesced): flags from alignment_build into the packs and the lists
        if (C.slotLd[s]) { list_push(listCnt, listLd, (uint32_t)s); V.pack[s].g |= SP_LD; }
        if (C.slotTr[s]) { list_push(listCnt + 1, listTr, (uint32_t)s); V.pack[s].g |= SP_TR; } });
      be.template launch_counted<st::fragment_adjust>(listCnt, N, [=] EQP_LAMBDA(uint64_t i) { fragment_adjust(A, C, V, listLd[i]); });                         // a17 before the dedupe
      be.template launch_occ<st::fragment_weight, 4>(nWork, [=] EQP_LAMBDA(uint64_t i) { fragment_weight(A, C, V, P, perm[i]); });
      if (V.alnG) be.template launch_counted<st::fragment_weight_gene>(listCnt + 1, N, [=] EQP_LAMBDA(uint64_t i) { fragment_weight_gene(C, V, listTr[i]); });      // gene-run keys of tr- fragments
      be.stage("fragment-weight");
      // the three counting runs of one fragment back to back in one thread: runs two and three find the fragment's alignments and
      // hits in cache instead of paying the DRAM latency chain again (the runs were latency-bound as three concurrent kernels).
      // First launch: small per-thread tables; (fragment, run) pairs that overflow them are queued for a second launch with the large tables.
      uint32_t* bigCnt = listCnt + 2; uint32_t* bigList = listLd;        // the a17 list is consumed by now; N + 1 entries hold 3 runs of N/3 fragments at least (overflow beyond is an error)
      be.memset(bigCnt, 0, 4);
      auto tally = [=] EQP_LAMBDA(int m, int r, uint64_t ws) {             // 0 not a weight object, 1 included, 2 excluded (ComputeCounts.java:986-989)
        counter_add(&V.stats[m * 3 + 0], r ? 1u : 0u); counter_add(&V.stats[m * 3 + 1], r == 1 ? 1u : 0u); counter_add(&V.stats[m * 3 + 2], r == 2 ? 1u : 0u);
        counter_add64(&V.stats[10 + m], ws);
      };
      // (Fusing did not pay: alignment_build + weight + count in one thread per fragment ran 217-266 ms against 67 ms for the three kernels at 25M
      // reads, weight + count alone 66-82 ms against 54 ms: profiles/r03d_variants_fused_25Mreads.jsonl, profiles/r03e_variants_25Mreads.jsonl.)
      be.template launch<st::fragment_count>(nWork, [=] EQP_LAMBDA(uint64_t i) {
        uint32_t s = perm[i];
        for (int m = 0; m < 3; m++) if (P.modes & (1 << m)) {
          uint64_t ws; int r = fragment_count<EQP_FS_SMALL, EQP_FT_SMALL>(A, C, V, P, m, s, ws);
          if (r < 0) { uint32_t k = atomic_inc_u32(bigCnt); if (k < N) bigList[k] = s * 4u + (uint32_t)m; else raise(C.err, DERR_CAPACITY, C.base + s); r = 0; ws = 0; }
          tally(m, r, ws);
        }
      });
      be.template launch_counted<st::fragment_count_large>(bigCnt, N, [=] EQP_LAMBDA(uint64_t i) {
        uint32_t s = bigList[i] >> 2; int m = (int)(bigList[i] & 3u);
        uint64_t ws; int r = fragment_count<EQP_FS_CAP, EQP_FT_CAP>(A, C, V, P, m, s, ws);
        tally(m, r, ws);
      }); }
    be.stage("fragment-count x3");
    // ---- results
    be.to_host(R.err, scal, 8);
    if (R.err[0]) return -1;
    be.to_host(R.stats, V.stats, sizeof R.stats);
    for (int m = 0; m < 3; m++) {
      R.fixed[m].resize(A.nFeat[m]); if (A.nFeat[m]) be.to_host(R.fixed[m].data(), V.acc[m], (uint64_t)A.nFeat[m] * 8);
      R.counts[m].resize(A.nFeat[m]);
      { const double scale = 1.0 / (double)(1ULL << EQP_FIXED_SHIFT_D); const unsigned long long* fx = R.fixed[m].data(); double* out = R.counts[m].data();   // exact: power-of-two scale
        for (int i = 0; i < A.nFeat[m]; i++) out[i] = (double)fx[i] * scale; }
    }
    R.junctionActive.resize(A.nAct); if (A.nAct) be.to_host(R.junctionActive.data(), V.junctionActive, A.nAct);
    return 0;
  }
};

}  // namespace eqpb
