// device_model.cuh — device-side read / reference model and the per-(read, locus) CIGAR cursor.
//
// B200-native restatement of the reference's PileupElement
// (src/main/scala/org/hammerlab/guacamole/pileup/PileupElement.scala, "PE.scala" below):
// instead of a chain of immutable cursor objects advanced locus by locus, the element of
// read r at locus x is recomputed from the packed CIGAR in registers — reads are
// columnar in HBM and a locus never needs the previous locus's objects.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/guac_b200.h"

namespace guac {

// Device view of an uploaded guac_read_batch (all pointers in device memory).
struct ReadsView {
  int64_t n_reads;
  const int64_t* start;
  const int64_t* end;         // derived: start + padded reference length (MappedRead.scala:75)
  const int64_t* runmax_key;  // derived: running max of (contig_id << 40 | end)
  const uint32_t* cigar_off;
  const uint32_t* cigar;
  const uint64_t* seq_off;
  const uint8_t* bases2;
  const uint8_t* nmask;
  int64_t n_exc;
  const uint64_t* exc_pos;
  const uint8_t* exc_base;
  const uint8_t* qual;
  const uint8_t* mapq;
  const uint8_t* flags;
  const uint8_t* sample_id;
  // derived at upload, or null: per read (start low word, span | sample << 16 | RF_* flags << 24, seq_off low, high)
  // (read_headers_pack_kernel)
  const uint4* hdr;
};

// One contig window in device memory: ASCII upper-case, 0 = missing.
struct RefView {
  const uint8_t* bases;
  int64_t offset;
  int64_t len;
  const uint32_t* bad;  // bit g: loci [offset + 16 g, +16) hold a base that is not A C G T (nullptr where unused)
  const uint32_t* ref2; // 2-bit copy of the window, 16 loci per word (stream form; nullptr where unused)
  __device__ __forceinline__ uint8_t at(int64_t locus) const {
    int64_t i = locus - offset;
    return (i >= 0 && i < len) ? __ldg(bases + i) : (uint8_t)0;
  }
};

// Tile of consecutive loci of one range, with the span of reads that can overlap it.
struct Tile {
  int64_t locus0;    // first locus
  int32_t n_loci;    // <= TILE_LOCI
  int32_t contig_id;
  int64_t slot0;     // dense output slot of locus0
  int64_t read_lo;   // reads [read_lo, read_hi) are the only ones that can overlap the tile
  int64_t read_hi;
  int32_t ref_index; // index into the RefView table
  int32_t _pad;
  uint64_t seq_base; // sequence offset of read_lo's first base, and of read_hi's (tiles built by build_tiles_kernel only)
  uint64_t seq_end;
};

enum : int { OPC_M = 0, OPC_I = 1, OPC_D = 2, OPC_N = 3, OPC_S = 4, OPC_H = 5, OPC_P = 6, OPC_EQ = 7, OPC_X = 8 };
// htsjdk CigarOperator.consumesReadBases / consumesReferenceBases as bit masks over op codes
__device__ __forceinline__ bool consumes_read(int op) { return (0x193 >> op) & 1; }  // M I S = X
__device__ __forceinline__ bool consumes_ref(int op) { return (0x18D >> op) & 1; }   // M D N = X

// device-side error codes (first error wins; reported through guac_last_error)
enum : int {
  DERR_NONE = 0,
  DERR_REF_MISSING = 1,       // "Position c:l missing from reference" (ReferenceBroadcast.scala:76-81)
  DERR_INSERTION_ELEMENT = 2, // InvalidCigarElementException (PE.scala:92)
  DERR_DELETION_PRECEDED = 3, // AssertionError "Found deletion preceded by cigar operator" (PE.scala:100-104)
  DERR_PADDING = 4,           // AssertionError on P (PE.scala:114-115)
  DERR_ALLELE_TABLE_FULL = 5, // library limit: > 32 distinct variant alleles at one locus of one sample
  DERR_EMPTY_MAXBY = 6,       // SomaticStandardCaller.scala:188 maxBy on an empty map
  DERR_CURSOR = 7,            // cursor ran past the last CIGAR element
  DERR_MISMATCH_HIST = 8      // library limit: median mismatch count beyond histogram
};

struct DeviceError {
  int code;
  int contig_id;
  long long locus;
  long long read_index;
};

__device__ __forceinline__ void report_error(DeviceError* e, int code, int contig, int64_t locus, int64_t read) {
  if (atomicCAS(&e->code, 0, code) == 0) {
    e->contig_id = contig;
    e->locus = locus;
    e->read_index = read;
  }
}

// Sequenced base i (batch-global index) as upper-case ASCII.
__device__ __forceinline__ uint8_t read_base(const ReadsView& R, uint64_t i) {
  uint8_t m = __ldg(R.nmask + (i >> 3));
  if (!((m >> (i & 7)) & 1)) {
    uint8_t b = __ldg(R.bases2 + (i >> 2));
    return (uint8_t)("ACGT"[(b >> ((i & 3) * 2)) & 3]);
  }
  // rare: binary search the exception list
  int64_t lo = 0, hi = R.n_exc;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (__ldg(R.exc_pos + mid) < i)
      lo = mid + 1;
    else
      hi = mid;
  }
  return __ldg(R.exc_base + lo);
}

// The PileupElement of one read at one locus (PE.scala:24-31) plus its Alignment class
// (pileup/Alignment.scala:14-75) and qualityScore (PE.scala:149-154).
struct Elem {
  int32_t read_position;
  int32_t cigar_element_index;
  int64_t cigar_element_locus;
  int32_t index_within_cigar_element;
  int32_t indel_length;
  uint8_t kind;
  uint8_t base;     // Match/Mismatch: sequenced base
  uint8_t quality;  // qualityScore
  uint8_t ref_base;
  // isFinalCigarBase and the operator of the following cigar element (-1 if none); only meaningful for it
  bool followed_by_ref_skip;  // final base of its element and the next element consumes reference but not read (D, N)
};

// PileupElement.apply(read, locus, contig) = cursor at read.start . advanceToLocus(locus)
// (PE.scala:250-259, 209-235, 165-188) followed by `alignment` (PE.scala:59-117).
// Requires start <= locus < end.  Returns a DERR_* code.
__device__ inline int element_at(const ReadsView& R, int64_t r, int64_t locus, const RefView& ref, Elem& e) {
  const int64_t start = __ldg(R.start + r);
  const uint32_t c0 = __ldg(R.cigar_off + r), c1 = __ldg(R.cigar_off + r + 1);
  const uint64_t s0 = __ldg(R.seq_off + r);
  const int32_t read_len = (int32_t)(__ldg(R.seq_off + r + 1) - s0);
  int64_t cl = start;
  int32_t rp = 0;
  uint32_t ci = c0;
  int op = 0, len = 0, idx = 0;
  bool found = false;
  for (; ci < c1; ++ci) {
    uint32_t w = __ldg(R.cigar + ci);
    op = w & 15;
    len = (int)(w >> 4);
    bool cref = consumes_ref(op), crd = consumes_read(op);
    if (cref && locus < cl + len) {  // currentCigarElementContainsLocus (PE.scala:195-197)
      idx = (int)(locus - cl);
      if (crd) rp += idx;
      found = true;
      break;
    }
    if (locus == 0 && op == OPC_I) {  // insertion at the start of a contig (PE.scala:227-232)
      idx = 0;
      found = true;
      break;
    }
    if (crd) rp += len;
    if (cref) cl += len;
  }
  if (!found) return DERR_CURSOR;
  e.read_position = rp;
  e.cigar_element_index = (int32_t)(ci - c0);
  e.cigar_element_locus = cl;
  e.index_within_cigar_element = idx;
  e.indel_length = 0;
  e.base = 0;
  e.ref_base = ref.at(locus);
  if (e.ref_base == 0) return DERR_REF_MISSING;  // PE.scala:36 `val referenceBase` is eager

  const bool is_final = idx == len - 1;  // isFinalCigarBase (PE.scala:57)
  int nop = op, nlen = len;
  if (is_final) {
    if (ci + 1 < c1) {
      uint32_t w = __ldg(R.cigar + ci + 1);
      nop = w & 15;
      nlen = (int)(w >> 4);
    } else {
      nop = -1;
    }
  }
  const uint8_t mapq = __ldg(R.mapq + r);
  e.followed_by_ref_skip = is_final && (nop == OPC_D || nop == OPC_N);  // ReadSubsequence.scala:108-112
  if ((op == OPC_M || op == OPC_EQ) && nop == OPC_I) {  // PE.scala:82-83
    e.kind = GUAC_ALN_INSERTION;
    e.indel_length = nlen;
  } else if (op == OPC_I && nop >= 0 && cl == 0) {  // PE.scala:88-89
    e.kind = GUAC_ALN_INSERTION;
    e.indel_length = len;
  } else if (op == OPC_I) {
    return DERR_INSERTION_ELEMENT;
  } else if ((op == OPC_M || op == OPC_EQ || op == OPC_X) && nop == OPC_D) {  // PE.scala:94-97
    e.kind = GUAC_ALN_DELETION;
    e.indel_length = nlen;
    // contigSequence.slice(locus, locus + len + 1) must exist
    for (int k = 1; k <= nlen; ++k)
      if (ref.at(locus + k) == 0) return DERR_REF_MISSING;
    e.quality = __ldg(R.qual + s0 + rp);
    return DERR_NONE;
  } else if (op == OPC_D) {  // PE.scala:98-99
    e.kind = GUAC_ALN_MID_DELETION;
    e.quality = mapq;
    return DERR_NONE;
  } else if (nop == OPC_D) {
    return DERR_DELETION_PRECEDED;
  } else if (op == OPC_M || op == OPC_EQ || op == OPC_X) {  // PE.scala:105-112
    e.base = read_base(R, s0 + rp);
    e.kind = e.base == e.ref_base ? GUAC_ALN_MATCH : GUAC_ALN_MISMATCH;
    e.quality = __ldg(R.qual + s0 + rp);
    return DERR_NONE;
  } else if (op == OPC_S || op == OPC_N || op == OPC_H) {  // PE.scala:113
    e.kind = GUAC_ALN_CLIPPED;
    e.quality = mapq;
    return DERR_NONE;
  } else {
    return DERR_PADDING;
  }
  // Insertion: qualityScore = min over anchor + inserted qualities (PE.scala:153);
  // the Scala `view(from, until)` clamps to the read length.
  int until = min(rp + e.indel_length + 1, read_len);
  uint8_t q = 255;
  for (int k = rp; k < until; ++k) q = min(q, __ldg(R.qual + s0 + k));
  e.quality = q;
  return DERR_NONE;
}

// 64-bit identity of a *variant* allele at a fixed locus (variants/Allele.scala:7-17):
//   Mismatch  (ref base, sequenced base)            -> sequenced base
//   Deletion  (ref[x .. x+len], ref[x])             -> len
//   MidDeletion (ref base, "")                      -> constant
//   Insertion (seq[0], seq[0 .. len])               -> the sequenced bases, exact for <= 27
//              standard bases, otherwise a 60-bit FNV-1a hash of the ASCII bases.
__device__ inline uint64_t allele_key(const ReadsView& R, int64_t r, const Elem& e) {
  switch (e.kind) {
    case GUAC_ALN_MISMATCH: return (1ull << 60) | e.base;
    case GUAC_ALN_DELETION: return (3ull << 60) | (uint32_t)e.indel_length;
    case GUAC_ALN_MID_DELETION: return (4ull << 60);
    case GUAC_ALN_INSERTION: {
      const uint64_t s0 = __ldg(R.seq_off + r);
      const int32_t read_len = (int32_t)(__ldg(R.seq_off + r + 1) - s0);
      int until = min(e.read_position + e.indel_length + 1, read_len);
      int n = until - e.read_position;
      bool exact = n <= 27;
      uint64_t packed = 0, h = 1469598103934665603ull;
      for (int k = 0; k < n; ++k) {
        uint8_t b = read_base(R, s0 + e.read_position + k);
        int code = b == 'A' ? 0 : b == 'C' ? 1 : b == 'G' ? 2 : b == 'T' ? 3 : -1;
        if (code < 0) exact = false;
        if (k < 27) packed |= (uint64_t)(code & 3) << (2 * k);
        h = (h ^ b) * 1099511628211ull;
      }
      if (exact) return (2ull << 60) | ((uint64_t)n << 54) | packed;
      return (5ull << 60) | (h & ((1ull << 60) - 1));
    }
    default: return 0;
  }
}

}  // namespace guac
