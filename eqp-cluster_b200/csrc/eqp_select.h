// a3-a7: per-fragment alignment statistics and best-set selection (per-fragment device functions).
// Reference: SamRecord.findMate/isMate (J/sam-source/SamRecord.java:582-608,650-666),
// SamProcessorFragmentEntry.processSamRecords (:100-217), FragmentEntry (:430-503,595-609,677-679,787-837),
// CombineReadWeightFiles (J/util-source/CombineReadWeightFiles.java:173-248), CombineSamFiles (:208-354,682-954).
#pragma once
#include "eqp_decode.h"

namespace eqpb {

EQP_HD int rec_flag_read_index(const SamView& S, uint32_t i) { uint8_t rf = S.rf[i]; return flag_read_index(S.flag[i], (rf >> 5) & 3, rf_read_index(rf)); }

EQP_HD bool is_mate(const SamView& S, uint32_t a, uint32_t b) {       // a.isMate(b); same fragment is given
  if (S.hi && S.hi[a] >= 0) return S.hi[b] == S.hi[a];                   // hit index decides alone (SamRecord.java:658-660)
  if (rec_flag_read_index(S, a) == rec_flag_read_index(S, b)) return false;
  if (((S.rf[a] >> 5) & 3) != ((S.rf[b] >> 5) & 3)) return false;      // originalQueryName equality (/1 vs /2 names never mate)
  return S.rnext[a] == S.ref[b] && S.pnext[a] == S.pos[b] && S.ref[a] == S.rnext[b] && S.pos[a] == S.pnext[b];
}

// a4 for one fragment of one file.  proc[] is a per-line scratch byte (zeroed by the caller), set for records consumed as mates.
EQP_HD void fragment_stats(const AnnoView& A, const SamView& S, uint32_t first, uint32_t end, bool whitelist, uint8_t* proc,
                           FragStat& out, FragFile& ff, uint32_t* err, uint32_t fragId) {
  FragStat f{}; int ri = 1; bool hasMate = false, any = false, added = false; uint8_t bits = 0;
  bool taken = false;                                                   // a record of this fragment was consumed as a mate: only then proc[] is read
  for (uint32_t i = first; i < end; i++) {
    uint8_t rf = S.rf[i];
    if (!(rf & RF_REC)) continue;
    int idx = rf_read_index(rf);
    bits |= 1 | (idx == 1 ? 8 : 16);
    if (!(rf & RF_MAPPED)) continue;
    bits |= (idx == 1 ? 2 : 4);
    any = true;
    if (taken && proc[i]) continue;
    int32_t ref = S.ref[i];
    bool ok = !whitelist || (ref >= 0 && (A.refFlags[ref] & 1));
    if (rf & RF_HASMATE) {
      uint32_t j = i + 1; bool found = false;
      for (; j < end; j++) { if ((S.rf[j] & RF_MAPPED) && !(taken && proc[j]) && is_mate(S, i, j)) { found = true; break; } }
      if (found) {
        proc[j] = 1; taken = true;                                        // i itself is never visited again (searches only look forward)
        if (ok) { f.edPe += S.nm[i] + S.nm[j]; f.nPe++; hasMate = true; added = true; }
      }                                                                 // else dropped (SamProcessorFragmentEntry.java:192-197)
    } else {
      if (ok) {                                                          // FragmentEntry.add(SamRecord) :430-477
        hasMate = false; added = true;
        if (idx != ri) ri = 3;
        if (idx == 1) { f.nSr1++; f.edSr1 += S.nm[i]; } else { f.nSr2++; f.edSr2 += S.nm[i]; }
      }
    }
  }
  if (any && !added) raise(err, DERR_NULL_WEIGHT, fragId);              // weight line would be named "null": reference aborts
  f.flags = (uint16_t)((any ? FS_EXISTS : 0) | (hasMate ? FS_HASMATE : 0) | (ri << 2));
  out = f; ff.bits = bits;
}

// ---- FragmentEntry arithmetic (fp64 compares exactly as the reference)
EQP_HD int fe_n(const FragStat& f) { int sr = f.nSr1 > f.nSr2 ? f.nSr1 : f.nSr2; return f.nPe + sr; }
EQP_HD int fe_sum(const FragStat& f) { return f.edPe + f.edSr1 + f.edSr2; }
EQP_HD double fe_ed(const FragStat& f) { return fe_sum(f) * 1.0 / fe_n(f); }
EQP_HD bool fe_mate(const FragStat& f) { return f.flags & FS_HASMATE; }
EQP_HD int fe_mult(const FragStat& f) { return (fe_mate(f) || fs_read_index(f) > 2) ? 2 : 1; }
EQP_HD int fe_cmp_mult(const FragStat& a, const FragStat& b) {        // FragmentEntry.compareMultiplicity :787-807
  if (fe_mate(a)) { if (!fe_mate(b)) return 1; } else if (fe_mate(b)) return -1;
  if (fe_mult(a) > 1) { if (fe_mult(b) <= 1) return 1; } else if (fe_mult(b) > 1) return -1;
  return 0;
}
EQP_HD int fe_cmp_ed(const FragStat& a, const FragStat& b, double slack) {   // compareEditDistance :815-837
  int rv = fe_cmp_mult(a, b);
  if (rv != 0) return rv;
  if (fe_mate(a)) slack = 2 * slack;
  if (fe_ed(b) <= fe_ed(a) && fe_n(b) > 1.1 * fe_n(a)) return -1;
  if (fe_ed(b) < fe_ed(a) - slack) return -1;
  return 1;
}

// ---- record-set descriptors: up to 4 parts, each (file, read-index filter) — 4 bits per part, count in bits 16-18
struct SetDesc { uint32_t v; };
EQP_HD int sd_n(SetDesc s) { return (s.v >> 16) & 7; }
EQP_HD int sd_file(SetDesc s, int k) { return (s.v >> (4 * k)) & 3; }
EQP_HD int sd_filter(SetDesc s, int k) { return (s.v >> (4 * k + 2)) & 3; }
EQP_HD SetDesc sd_single(int file) { return SetDesc{(uint32_t)file | (1u << 16)}; }
EQP_HD SetDesc sd_append(SetDesc a, SetDesc b, int filterA, int filterB) {   // printIdx(a, fA) then printIdx(b, fB)
  SetDesc r{0}; int n = 0;
  for (int side = 0; side < 2; side++) {
    SetDesc s = side ? b : a; int f = side ? filterB : filterA;
    for (int k = 0; k < sd_n(s); k++) {
      int of = sd_filter(s, k), nf;
      if (of == 0) nf = f; else if (f == 0 || of == f) nf = of; else continue;
      if (n < 4) { r.v |= (uint32_t)(sd_file(s, k) | (nf << 2)) << (4 * n); n++; }
    }
  }
  r.v |= (uint32_t)n << 16; return r;
}

struct Side {                     // what CombineSamFiles knows about one input for one fragment
  bool present; int mappedIdx;    // bit0: mapped record with read index 1, bit1: read index 2
  FragStat w; SetDesc set;
};
EQP_HD Side side_of_file(int file, const FragFile& ff, const FragStat& st) {
  Side s; s.present = ff.bits & 1; s.mappedIdx = (ff.bits >> 1) & 3; s.w = st; s.set = sd_single(file); return s;
}
EQP_HD Side side_of_set(SetDesc set, const FragFile* ff, const FragStat& w) {   // the printed output of a previous stage
  Side s; s.set = set; s.present = sd_n(set) > 0; s.mappedIdx = 0; s.w = w;
  for (int k = 0; k < sd_n(set); k++) {
    int m = (ff[sd_file(set, k)].bits >> 1) & 3, f = sd_filter(set, k);
    s.mappedIdx |= f == 0 ? m : (m & f);
  }
  return s;
}

// CombineSamFiles for a fragment present in both inputs (main loop body :717-832 + printSamRecords :290-354)
EQP_HD SetDesc combine_both(const Side& a, const Side& b, double slack, uint32_t* err, uint32_t fragId) {
  bool m1 = a.mappedIdx != 0, m2 = b.mappedIdx != 0;
  if (!m1 && !m2) return a.set;
  if (!m2) return a.set;
  if (!m1) return b.set;
  if (!(a.w.flags & FS_EXISTS) || !(b.w.flags & FS_EXISTS)) { raise(err, DERR_INCONSISTENT, fragId); return a.set; }
  int cm = fe_cmp_mult(a.w, b.w);
  if (cm == 1) return a.set;
  if (cm == -1) return b.set;
  if (fe_ed(a.w) < 0 || fe_ed(b.w) < 0) { raise(err, DERR_INCONSISTENT, fragId); return a.set; }
  const Side& x = fe_cmp_ed(a.w, b.w, slack) == 1 ? a : b; const Side& y = (&x == &a) ? b : a;
  if (fe_mate(x.w)) return x.set;
  if (fe_mate(y.w)) return y.set;
  if (x.mappedIdx == 3) return x.set;
  if (x.mappedIdx & 1) return sd_append(x.set, y.set, 1, 2);
  return sd_append(y.set, x.set, 1, 2);                                // x covers read 2 only: printIdx(y,1); printIdx(x,2)
}

// One CombineSamFiles run for slot s, including the drop quirks for fragments missing from one input:
// first1/first2 = first present slot of each input, last1 = last present slot of input 1 (UINT32_MAX / 0 when empty).
EQP_HD SetDesc combine_stage(const Side& a, const Side& b, uint32_t s, uint32_t first1, uint32_t first2, uint32_t last1, bool any1, bool any2,
                             double slack, uint32_t* err, uint32_t fragId) {
  bool p1 = a.present, p2 = b.present;
  if (any1 && any2) {                                                  // the first iteration advances BOTH readers (:839-843)
    if (first2 < first1 && s == first1) p1 = false;
    if (first1 < first2 && s == first2) p2 = false;
  }
  if (!any1 || !any2) {                                                // while-loop never runs; tail keeps input 1 only (:922-954)
    return (p1 && !any2) ? a.set : SetDesc{0};
  }
  if (p1 && p2) return combine_both(a, b, slack, err, fragId);
  if (p1) return a.set;                                                // realign loop or tail of input 1
  if (p2) return s < last1 ? b.set : SetDesc{0};                       // input-2-only groups after input 1 ends are dropped
  return SetDesc{0};
}

// CombineReadWeightFiles -e for one fragment (T weights vs J weights, no aux)
EQP_HD FragStat combine_weights_ed(const FragStat& t, const FragStat& j, double slack) {
  bool ht = t.flags & FS_EXISTS, hj = j.flags & FS_EXISTS;
  if (!ht) return j;
  if (!hj) return t;
  return fe_cmp_ed(t, j, slack) >= 0 ? t : j;
}

}  // namespace eqpb
