// a8-a11: combined record set -> aligned blocks (CIGAR), then the interval join against the bucketed index.
// Reference: SamProcessorBed.processSamRecords (J/sam-source/SamProcessorBed.java:152-204), SamRecord.printBedRecords
// (:484-573), SamRecord.computeBedEntries (:316-475), BedEntry (:76-88); bedtools intersect -wo (CIB:938-940).
#pragma once
#include "eqp_select.h"

namespace eqpb {

struct ChunkView {                 // everything the per-fragment stages read/write for one chunk
  SamView sam[3];
  uint8_t* proc[3];                // per-line scratch
  uint32_t base, nSlots;           // fragment id of slot 0, slot count
  FragFile* ff[3]; FragStat* st[3]; FragStat* stTJ;
  SetDesc* setTJ; SetDesc* setC;   // stage-1 and final record sets
  uint32_t* auxSlot;               // first slot >= s that has a T+J weight entry (UINT32_MAX if none)
  uint32_t* blkOff; uint32_t* alnOff;        // per slot: exclusive offsets (nSlots + 1)
  Block* blocks; uint32_t* hitCnt;           // hitCnt: per block count, then exclusive offsets (nBlocks + 1)
  uint32_t* alnBlk;                          // per alignment: its first block (nAln + 1 entries; blocks of an alignment are contiguous)
  Hit* hits;
  uint8_t* seenRef;                // a27: reference ids seen in combined SAM or junction SAM
  uint8_t* slotTr; uint8_t* slotLd;  // per slot: the fragment has a tr- exon hit / an interval with a positive length difference (set by alignment_build)
  uint32_t* err;
};

// iterate the mapped records of a record set in output order
struct SetCursor { int part; uint32_t line; };
EQP_HD bool set_next(const ChunkView& C, SetDesc sd, uint32_t slot, SetCursor& c, int& file) {
  while (c.part < sd_n(sd)) {
    int f = sd_file(sd, c.part), flt = sd_filter(sd, c.part);
    const FragFile& ff = C.ff[f][slot];
    if (c.line == 0xFFFFFFFFu) c.line = ff.first; else c.line++;
    for (; c.line < ff.end; c.line++) {
      uint8_t rf = C.sam[f].rf[c.line];
      if (!(rf & RF_REC)) continue;
      if (flt && rf_read_index(rf) != flt) continue;
      if (!(rf & RF_MAPPED)) continue;                                  // SamReader.java:197
      file = f; return true;
    }
    c.part++; c.line = 0xFFFFFFFFu;
  }
  return false;
}

// CIGAR -> segments (split at N).  Calls emit(seg, refStartOffset, refLen, readLen, insSoft, del) for every segment.
template <class F>
EQP_HD int cigar_segments(const char* cg, int n, F&& emit, int& sumRead, bool& badOp) {
  int seg = 0, m = 0, ins = 0, del = 0, soft = 0; int sumRef = 0, sumIntron = 0; long num = -1; sumRead = 0; badOp = false;
  for (int i = 0; i <= n; i++) {
    char c = i < n ? cg[i] : 'N';                                       // virtual terminator closes the last segment
    if (i < n && c >= '0' && c <= '9') { num = (num < 0 ? 0 : num) * 10 + (c - '0'); continue; }
    bool op = i == n || (num >= 0 && ((c >= 'A' && c <= 'Z') || c == '='));
    if (op) {
      if (c == 'N') {
        int refLen = m + del, readLen = m + ins + soft;
        emit(seg, sumRef + sumIntron, refLen, readLen, ins + soft, del);
        sumRef += refLen; sumRead += readLen; sumIntron += (i < n ? (int)num : 0);
        seg++; m = ins = del = soft = 0;
      } else if (c == 'M' || c == 'X' || c == '=') m += (int)num;
      else if (c == 'I') ins += (int)num;
      else if (c == 'D') del += (int)num;
      else if (c == 'S') soft += (int)num;
      else if (c != 'H' && c != 'P') badOp = true;
    }
    num = -1;
  }
  return seg;                                                           // number of segments (ciagrFieldNum)
}

// segments of one record: "<n>M" CIGARs were already reduced to n by the decoder (SamView::cigM), anything else is parsed from the text
template <class F>
EQP_HD int record_segments(const SamView& R, uint32_t li, F&& emit, int& sumRead, bool& badOp) {
  uint16_t m = R.cigM[li];
  if (m != 0xFFFFu) { badOp = false; sumRead = m; emit(0, 0, (int)m, (int)m, 0, 0); return 1; }
  return cigar_segments(R.text + R.lineOff[li] + R.cigOff[li], R.cigLen[li], emit, sumRead, badOp);
}

// Pass 1 (count) and pass 2 (fill) over one fragment's combined record set.
// stranded: 0 = no, 1 = forward, 2 = backward: the block strand becomes the fragment strand (SamRecord.java:431-436,859-873)
EQP_HD void fragment_blocks(const AnnoView& A, const ChunkView& C, uint32_t slot, bool fill, int stranded, uint32_t& nBlocksOut, uint32_t& nAlnOut) {
  SetDesc sd = C.setC[slot];
  uint32_t nb = 0, na = 0; uint32_t bo = fill ? C.blkOff[slot] : 0;
  SetCursor ci{0, 0xFFFFFFFFu}; int fi; bool taken = false;             // taken: a record of this fragment was consumed as a mate (only then proc[] is read)
  while (set_next(C, sd, slot, ci, fi)) {
    const SamView& S = C.sam[fi]; uint32_t i = ci.line;
    if (taken && C.proc[fi][i]) continue;
    int mateFile = -1; uint32_t mateLine = 0;
    if (S.rf[i] & RF_HASMATE) {                                          // findMate over the rest of the group
      SetCursor cj = ci; int fj;
      while (set_next(C, sd, slot, cj, fj)) {
        if (taken && C.proc[fj][cj.line]) continue;
        const SamView& T = C.sam[fj];
        // isMate across inputs only needs the decoded fields
        int ia = rec_flag_read_index(S, i), ib = rec_flag_read_index(T, cj.line);
        bool mate = (S.hi && S.hi[i] >= 0) ? (T.hi && T.hi[cj.line] == S.hi[i])      // hit index decides alone (SamRecord.java:658-660)
                  : (ia != ib && ((S.rf[i] >> 5) & 3) == ((T.rf[cj.line] >> 5) & 3) && S.rnext[i] == T.ref[cj.line] && S.pnext[i] == T.pos[cj.line] &&
                     S.ref[i] == T.rnext[cj.line] && S.pos[i] == T.pnext[cj.line]);
        if (mate) { mateFile = fj; mateLine = cj.line; C.proc[fj][cj.line] = 1; taken = true; break; }
      }
      if (mateFile >= 0) {
        const SamView& T = C.sam[mateFile];
        if (!(S.ref[i] == T.rnext[mateLine] && S.rnext[i] == T.ref[mateLine])) raise(C.err, DERR_MATE_REF, C.base + slot);
      }
    }
    int32_t ref = S.ref[i];
    uint32_t v0 = ref >= 0 ? A.refVerOff[ref] : 0, v1 = ref >= 0 ? A.refVerOff[ref + 1] : 0;
    for (uint32_t v = v0; v < v1; v++) {
      int32_t bucket = A.refVerBucket[v];
      if (fill) C.alnBlk[C.alnOff[slot] + na] = bo + nb;                 // first block of this alignment
      for (int side = 0; side < 2; side++) {
        if (side == 1 && mateFile < 0) break;
        const SamView& R = side ? C.sam[mateFile] : S; uint32_t li = side ? mateLine : i;
        int bedStart = R.pos[li] - 1 > 0 ? R.pos[li] - 1 : 0;
        uint8_t ridx = (uint8_t)rec_flag_read_index(R, li), paired = (R.rf[li] & RF_HASMATE) ? 1 : 0, strand = (R.flag[li] & 16) ? '-' : '+';
        if (stranded && rf_read_index(R.rf[li]) != (stranded == 1 ? 1 : 2)) strand = strand == '+' ? '-' : '+';   // getFragmentStrand / getReverseFragmentStrand
        int sumRead; bool bad; uint32_t firstOfRec = nb;
        int nSeg = record_segments(R, li, [&](int seg, int off, int refLen, int readLen, int insSoft, int del) {
          if (readLen <= 0) return;
          if (fill) {                                                    // assembled in registers, stored as one 32-byte struct
            Block b;
            b.bucket = bucket; b.start = bedStart + off; b.end = bedStart + off + refLen; b.L = (uint16_t)readLen; b.I = (uint16_t)insSoft; b.D = (uint16_t)del;
            b.seg = (uint8_t)seg; b.maxSeg = 0; b.aln = na; b.readIdx = ridx; b.paired = paired; b.strand = strand; b.pad = 0; b.slot = slot;
            C.blocks[bo + nb] = b;
          }
          nb++;
        }, sumRead, bad);
        if (fill && nSeg > 1) for (uint32_t k = firstOfRec; k < nb; k++) C.blocks[bo + k].maxSeg = (uint8_t)(nSeg - 1);
        if (bad) raise(C.err, DERR_CIGAR_OP, C.base + slot);
        if (sumRead != R.seqLen[li]) raise(C.err, DERR_CIGAR_LEN, C.base + slot);
      }
      na++;
    }
    if (v0 == v1) {                                                      // reference without B rows: still validates the CIGAR(s)
      for (int side = 0; side < 2; side++) {
        if (side == 1 && mateFile < 0) break;
        const SamView& R = side ? C.sam[mateFile] : S; uint32_t li = side ? mateLine : i;
        int sumRead; bool bad;
        record_segments(R, li, [&](int, int, int, int, int, int) {}, sumRead, bad);
        if (bad) raise(C.err, DERR_CIGAR_OP, C.base + slot);
        if (sumRead != R.seqLen[li]) raise(C.err, DERR_CIGAR_LEN, C.base + slot);
      }
    }
  }
  nBlocksOut = nb; nAlnOut = na;
}

// a11 for one block: rows of the bucket overlapping [start,end) by >= 1 bp.
// sameStrand: `bedtools intersect -s` (CIB:938-940 with BED_STRAND_SPECIFIC_OPTION): only rows on the block's strand.
// The rows are walked from the last candidate down; with `out` the hits are stored in ASCENDING row order (nOut = the count of the first
// pass), which is the order BedRecord.compareTo wants for the records of one block (rows are sorted by start, end, strand, file order),
// so that the per-alignment sort of alignment_build finds them sorted.
EQP_HD uint32_t join_block(const AnnoView& A, const Block& b, Hit* out, uint32_t nOut, uint32_t blockIdx, bool sameStrand) {
  if (b.bucket < 0) return 0;
  const uint32_t wantStrand = b.strand == '+' ? 0u : 1u;
  int32_t lo = (int32_t)A.bktRowOff[b.bucket], hi = (int32_t)A.bktRowOff[b.bucket + 1];
  int32_t l = lo, h = hi;                                                // first row with start >= b.end
  while (l < h) { int32_t m = (l + h) >> 1; if (A.row[m].start < b.end) l = m + 1; else h = m; }
  uint32_t n = 0;
  for (int32_t k = l - 1; k >= lo; k--) {
    const RowA r = A.row[k];
    if (r.maxEnd <= b.start) break;
    int32_t re = r.end;
    if (re > b.start && (!sameStrand || ((r.flags() >> 1) & 3u) == wantStrand)) {
      int32_t ov = (re < b.end ? re : b.end) - (r.start > b.start ? r.start : b.start);
#ifdef EQP_VAR_DESC_HITS
      if (ov >= 1) { if (out) out[n] = Hit{blockIdx, k, ov, 0}; n++; }
#else
      if (ov >= 1) { if (out) out[nOut - 1 - n] = Hit{blockIdx, k, ov, 0}; n++; }
#endif
    }
  }
  return n;
}

}  // namespace eqpb
