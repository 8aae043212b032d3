// Plain-old-data views shared by the CUDA pipeline and the host-side annotation builder.
// Layouts are described in DESIGN.md ("Data layout in HBM").
#pragma once
#include <stddef.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define EQP_HD __host__ __device__ __forceinline__
#ifdef EQP_VAR_INLINE_KEYS
#define EQP_HD_CALL __host__ __device__ __forceinline__
#else
#define EQP_HD_CALL __host__ __device__ __noinline__      // real calls: keeps the callers' register allocation small
#endif
#else
#define EQP_HD inline
#define EQP_HD_CALL inline
#endif

namespace eqpb {

enum { FILE_T = 0, FILE_J = 1, FILE_G = 2 };
enum { MODE_GENE = 0, MODE_EXON = 1, MODE_JUNCTION = 2 };

// device error word: first error wins (atomicCAS), detail = fragment id or record index
enum {
  DERR_NONE = 0,
  DERR_QNAME = 1,          // QNAME is not Fnnnnnnnnn[/1|/2]  (SamReader.java:157 would renumber)
  DERR_FIELDS = 2,         // too few / too many SAM fields (SamRecord.java:251-257)
  DERR_SEQ_QUAL = 3,       // SEQ/QUAL length mismatch (SamRecord.java:246-249)
  DERR_READ_INDEX = 4,     // read index in flag differs from name (SamRecord.java:182-186)
  DERR_ORDER = 5,          // SAM records out of order (CombineSamFiles.java:191-193)
  DERR_NULL_WEIGHT = 6,    // fragment with only non-whitelisted genome hits: reference prints "null" and aborts
  DERR_CIGAR_OP = 7,       // unknown CIGAR operation (SamRecord.java:419-421)
  DERR_CIGAR_LEN = 8,      // read length != M+I+S (SamRecord.java:463-466)
  DERR_MATE_REF = 9,       // "Differing referenceName mate found" (SamRecord.java:567-571)
  DERR_TR_LENDIFF = 10,    // unsupported corner: one alignment has both a tr- exon record and a record with a length difference
                           // (its field indices, hence its adapted coordinates, then differ between the junction run and the others)
  DERR_ACC_OVERFLOW = 11,  // a fixed-point accumulator passed 2^62 (one feature took more than 2^24 reads' worth of weight)
  DERR_CAPACITY = 12,      // per-fragment table overflow
  DERR_GENOME_COORD = 13,  // BedRecord.java:669-673 / :683-687 self-checks
  DERR_INCONSISTENT = 14,  // GenomicAlignment.java consistency throws (:123-148, :245-300)
  DERR_BAD_JUNCTION_ID = 15,
  DERR_STRANDED = 16,      // strand-specific mode: direction is neither forward (1) nor backward (2)
  DERR_INT = 17            // integer field does not parse
};

// reference-name dictionary slot: key = (name_hash(name) & ~0xFFFF) | length, 0 = empty.  One 16-byte load yields everything lookup_ref needs.
struct alignas(16) DictEntry { uint64_t key; int32_t val; uint32_t off; };

// one row of the B file / one distinct exon, 16 bytes each: a hit needs all fields of its row and of its exon, one load apiece
struct alignas(16) RowA { int32_t start, end, maxEnd; uint32_t exonFlags;      // exonFlags = exon << 3 | flags: bit0 genomic row (bucket name == exon chromosome), bits1-2 row strand (0 '+', 1 '-', 2 other)
  EQP_HD int32_t exon() const { return (int32_t)(exonFlags >> 3); } EQP_HD uint32_t flags() const { return exonFlags & 7u; } };
struct alignas(16) ExonA { int32_t chr, start, end; uint32_t strand; };         // chr <= -2: tr- exon of transcript -2 - chr; strand: the character

// Where the features of a distinct exon are, for all three runs in one 16-byte load (the counting runs look up every hit's exon in each run)
struct alignas(16) ExonFeat {
  uint32_t g, e;                   // gene / exon run: the feature id itself when the exon has exactly one (the usual case), else the offset of its list in feat[run]
  uint32_t jOff;                   // junction run: offset of its id-sorted list in feat[MODE_JUNCTION]
  uint32_t n;                      // list lengths: bits 0-7 gene run, 8-15 exon run, 16-31 junction run
};

struct AnnoView {
  // reference-name dictionary (open addressing, power-of-two capacity)
  const DictEntry* dict; uint32_t dictMask;
  const char* strPool; const uint32_t* refStrOff; const uint16_t* refStrLen;
  const uint8_t* refFlags;          // bit0 in genome.fa.fai, bit1 malformed junction-like id
  const uint32_t* refVerOff; const int32_t* refVerBucket;   // SAM name -> B buckets (a8 version expansion)
  const uint32_t* refActOff; const int32_t* refAct;         // SAM name -> junction features it activates (a27)
  int32_t nRef;
  // B index: buckets -> rows sorted by start
  const uint32_t* bktRowOff; const uint8_t* bktFlags;       // bit0 junction, bit1 .inv-strand
  const int32_t* bktRefId;          // identity of the bucket's reference id after BedRecord.java:154-176 (.inv-strand and version stripped)
  int32_t nBucket;
  const RowA* row;                  // rows of a bucket sorted by start; maxEnd = running maximum of end (the join's search key)
  // distinct exons
  const ExonA* ex;
  int32_t nExon; int32_t nTrChr;    // nTrChr: number of distinct tr- chromosomes (0: the annotation has no tr- exons)
  // exon -> feature lists per mode; junction entries carry bit31 = "row also in exon_junction_transcript.map"
  const ExonFeat* exf; const uint32_t* feat[3]; int32_t nFeat[3];
  // 3-column junction map (Exon.java:107-125): a row is switched on by its column-2 id, counted under its column-3 id.
  // featAct is parallel to feat[MODE_JUNCTION] (null: the activation index is the feature id); nAct = number of activation ids
  const uint32_t* featAct; int32_t nAct;
};

struct SamView {                   // one decoded SAM input (structure of arrays, one entry per text line)
  const char* text; uint64_t len;
  uint64_t* lineOff;               // nLines + 1 entries
  uint32_t nLines;
  uint32_t* frag; int32_t* ref; int32_t* pos; int32_t* rnext; int32_t* pnext;
  uint16_t* flag; int16_t* nm; uint16_t* cigOff; uint16_t* cigLen; uint16_t* seqLen;
  uint16_t* cigM;                  // n when the CIGAR is exactly "<n>M" (the a10 stage then never touches the text), else 0xFFFF
  uint8_t* rf;                     // bit0 record (not header/empty), bit1 mapped, bit2 hasMate, bits3-4 readIndex
  int32_t* hi;                     // HI:i: hit index (-1: none), SamRecord.java:233-240; null unless the input comes from a genomic aligner
                                   // (compute-genomic-counts.sh) — bowtie2 writes no HI tag and the default flow rejects one
};
enum { RF_REC = 1, RF_MAPPED = 2, RF_HASMATE = 4 };
EQP_HD int rf_read_index(uint8_t rf) { return (rf >> 3) & 3; }

struct FragStat {                  // FragmentEntry of one fragment in one file (a4), 20 bytes
  uint16_t nPe, nSr1, nSr2; uint16_t flags;        // flags: bit0 entry exists, bit1 hasMate(last add), bits2-3 readIndex state (1,2,3)
  int32_t edPe, edSr1, edSr2;                      // NM sums (a missing NM tag contributes -1, SamRecord.java:1063-1065)
};
enum { FS_EXISTS = 1, FS_HASMATE = 2 };
EQP_HD int fs_read_index(const FragStat& f) { return (f.flags >> 2) & 3; }

struct FragFile {                  // record range + presence bits of one fragment in one file
  uint32_t first, end;             // line range [first, end)
  uint8_t bits;                    // bit0 present, bit1 mapped idx1, bit2 mapped idx2, bit3 has idx1 rec, bit4 has idx2 rec, bit5 mapped idx0
};

struct alignas(16) Block {         // one aligned block of the combined BED (a10), 32 bytes: two 16-byte loads / stores
  int32_t bucket; int32_t start, end;              // 0-based half-open, in the bucket's coordinates
  uint16_t L, I, D; uint8_t seg, maxSeg;           // -L -I -D -F<seg>-<maxSeg>
  uint32_t aln;                                    // alignment number within the fragment (A-counter - 1)
  uint8_t readIdx, paired, strand, pad;            // strand '+'/'-'
  uint32_t slot;                                   // fragment slot the block belongs to
};

struct alignas(16) Hit { uint32_t block; int32_t row; int32_t overlap; uint32_t pad; };   // a11 output, padded to one 16-byte load / store

struct alignas(8) HitX {            // per hit, after the per-alignment sort/dedupe (a12-a16), 24 bytes = three 8-byte words
  int32_t gs, ge;                  // genomic interval, 1-based closed
  int32_t chr;                     // chromosome index of the exon
  int8_t ivField, ivFieldJ;        // field index of the (kept) interval this record belongs to; ivFieldJ: the same in the junction run, where
                                   // tr- exon records do not take an index (-1, GenomicAlignment.java:233-234)
  uint16_t ivRank;                 // rank of that interval among the distinct intervals of its half
  int32_t exon;                    // distinct-exon id of the hit's annotation row
  uint8_t flags;                   // bit0 valid, bit1 genomic row, bits2-4 conforming in gene/exon/junction run, bit5 interval strand '-'
  uint8_t half; uint16_t grp;      // half: 1 or 2 (read index; after a19 the half of the REPRESENTATIVE it merges into); grp: representative's index within the fragment.
                                   // Fragments with a tr- exon hit are never tagged: half stays the read index and grp holds the interval's ordinal
                                   // among the non-tr intervals of its half (0xFFFF for a tr- interval), see eqp_accumulate.h
};

struct alignas(16) Aln {            // one WeightObjectAlignment (a15-a19), 48 bytes
  uint32_t hit0, split, hit1;      // hits [hit0,split) read 1, [split,hit1) read 2
  uint32_t repr;                   // global index of the first alignment of the fragment with the same key
  uint64_t key;
  int32_t start1, start2;          // GenomicAlignment.start per half
  uint16_t nIv1, nIv2; int16_t maxField1, maxField2;
  uint8_t flags;                   // bit0 paired, bit1 has half 1, bit2 has half 2, bits3-4 SR read index, bit5 halves swapped relative to repr
  uint8_t tr;                      // per half h (shift 3*(h-1)): bit0 has a tr- interval, bit1 all intervals are tr-, i.e. its string is the literal "" outside the gene run
  uint8_t pad[2];
  uint32_t next;                   // next alignment of the fragment with the same key (duplicate chain, file order), UINT32_MAX = end
};

// Everything the per-fragment stages (weight, count) need to know about a fragment slot, and their per-fragment outputs, in ONE 32-byte
// sector: those kernels walk the fragments in work order, so every separate per-slot array they touch costs its own DRAM sector per fragment
// (ncu: fragment_weight moved 1.8 KB per fragment, profiles/r03d).  Built in slot order after the join (coalesced).
struct alignas(32) SlotPack {
  uint32_t hit0, hit1;             // the fragment's hit range
  uint32_t aln0, aln1;             // its alignment range
  uint32_t a20, nFinal;            // fragment_weight: number of genomic alignments (0 = not in the intersection), numAlignments of the final.wgt entry
  uint32_t a20G;                   // fragment_weight_gene: a20 under the gene run's keys (fragments with tr- exon hits)
  uint32_t g;                      // bits 0-23 numAlignments of the genome SAM weight entry, bits 24-25 how it combines with a20 (SP_*), bit 27 tr- hit, bit 28 positive length difference
};
enum { SP_A20 = 0, SP_GENOME = 1, SP_MAX = 2, SP_TR = 1u << 27, SP_LD = 1u << 28 };

// second dedupe of the alignments of a fragment with tr- exon hits: keys of the gene run, where tr- intervals are printed (GenomicInterval.java:711-713)
struct AlnG { uint64_t key; uint32_t repr, next; uint8_t flags; uint8_t pad[7]; };   // flags bit5: halves swapped relative to repr

}  // namespace eqpb
