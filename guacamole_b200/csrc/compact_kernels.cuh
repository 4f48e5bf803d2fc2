// compact_kernels.cuh — expansion / compaction between guac_read_batch_compact (the PCIe wire form) and the
// columnar batch the kernels read.  All of it is streaming byte work: one pass per column.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace guac {
namespace compact {

// per-read contig id from the runs (binary search over a handful of runs)
__global__ void contig_from_runs_kernel(int64_t n, const int32_t* __restrict__ run_contig, const int64_t* __restrict__ run_first,
                                        int n_runs, int32_t* __restrict__ contig_id) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  int lo = 0, hi = n_runs - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (run_first[mid] <= r) lo = mid; else hi = mid - 1;
  }
  contig_id[r] = run_contig[lo];
}

// narrow start column: start = base of the read's contig run + a 32-bit offset
__global__ void starts_from_rel_kernel(int64_t n, const int64_t* __restrict__ run_base, const int64_t* __restrict__ run_first, int n_runs,
                                       const uint32_t* __restrict__ rel, int64_t* __restrict__ start) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  int lo = 0, hi = n_runs - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (run_first[mid] <= r) lo = mid; else hi = mid - 1;
  }
  start[r] = run_base[lo] + (int64_t)rel[r];
}

// every read has `len` bases: offsets without a scan
__global__ void uniform_offsets_kernel(int64_t n_plus_1, uint64_t len, uint64_t* __restrict__ seq_off) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n_plus_1) seq_off[i] = (uint64_t)i * len;
}

// narrow operator counts: a read with count 0 has one (implicit) operator
__global__ void effective_ops_kernel(int64_t n, const uint8_t* __restrict__ ops8, uint8_t* __restrict__ eff) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < n) eff[r] = ops8[r] ? ops8[r] : (uint8_t)1;
}
// ... and the full operator array: the implicit operator is `read length` M, the others are copied from the wire
__global__ void cigar_from_ops8_kernel(int64_t n, const uint8_t* __restrict__ ops8, const uint32_t* __restrict__ cigar_off,
                                       const int64_t* __restrict__ wire_off, const uint32_t* __restrict__ wire,
                                       const uint64_t* __restrict__ seq_off, uint32_t* __restrict__ cigar) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  const uint32_t o = cigar_off[r];
  const int k = ops8[r];
  if (k == 0) {
    cigar[o] = (uint32_t)(seq_off[r + 1] - seq_off[r]) << 4;  // operator code 0 = M
  } else {
    const int64_t w = wire_off[r];
    for (int j = 0; j < k; ++j) cigar[o + j] = wire[w + j];
  }
}

__global__ void narrow_offsets_kernel(int64_t n, const int64_t* __restrict__ in, uint32_t* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (uint32_t)in[i];
}

// not-ACGT mask from the exception list (the mask buffer is zeroed first)
__global__ void nmask_from_exceptions_kernel(int64_t n_exc, const uint64_t* __restrict__ exc_pos, uint8_t* __restrict__ nmask) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_exc) return;
  const uint64_t p = exc_pos[i];
  unsigned int* w = reinterpret_cast<unsigned int*>(nmask) + (p >> 5);
  atomicOr(w, 1u << (p & 31));
}

struct QualTable {
  uint8_t q[16];
};

// BITS-bit codes (2 or 4) -> one byte per base; each thread expands 16 bases
template <int BITS>
__global__ void expand_qual_kernel(uint64_t n_bases, const uint8_t* __restrict__ packed, QualTable T, uint8_t* __restrict__ qual) {
  __shared__ uint8_t s_q[16];
  if (threadIdx.x < 16) s_q[threadIdx.x] = T.q[threadIdx.x];
  __syncthreads();
  const uint64_t i = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 16;
  if (i >= n_bases) return;
  constexpr uint32_t MASK = (1u << BITS) - 1;
  // 16 bases = 64 bits (4-bit codes) or 32 bits (2-bit codes); buffers are padded to 16 bytes
  unsigned long long w;
  if (BITS == 4) w = *reinterpret_cast<const unsigned long long*>(packed + (i >> 1));
  else w = *reinterpret_cast<const uint32_t*>(packed + (i >> 2));
  uint32_t o[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const uint32_t src = (uint32_t)(w >> (4 * BITS * k));
    o[k] = (uint32_t)s_q[src & MASK] | ((uint32_t)s_q[(src >> BITS) & MASK] << 8) | ((uint32_t)s_q[(src >> (2 * BITS)) & MASK] << 16) |
           ((uint32_t)s_q[(src >> (3 * BITS)) & MASK] << 24);
  }
  if (i + 16 <= n_bases) {
    *reinterpret_cast<uint4*>(qual + i) = make_uint4(o[0], o[1], o[2], o[3]);
  } else {
    for (uint64_t k = i; k < n_bases; ++k) qual[k] = (uint8_t)(o[(k - i) >> 2] >> (8 * ((k - i) & 3)));
  }
}

// ---- the other direction (export) ----
__global__ void lengths_kernel(int64_t n, const uint64_t* __restrict__ seq_off, const uint32_t* __restrict__ cigar_off,
                               uint16_t* __restrict__ read_len, uint16_t* __restrict__ n_ops, int* __restrict__ too_long) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  const uint64_t len = seq_off[r + 1] - seq_off[r];
  const uint32_t ops = cigar_off[r + 1] - cigar_off[r];
  if (len > 65535 || ops > 65535) *too_long = 1;
  read_len[r] = (uint16_t)len;
  n_ops[r] = (uint16_t)ops;
}

// narrow columns of the export: 32-bit start offsets within the contig run, 8-bit operator counts with the implicit
// single-M form; flags bit 0: an offset beyond 32 bits, bit 1: a read with more than 255 operators, bit 2: reads of
// different lengths
__global__ void narrow_columns_kernel(int64_t n, const int64_t* __restrict__ start, const int64_t* __restrict__ run_first, int n_runs,
                                      const uint64_t* __restrict__ seq_off, const uint32_t* __restrict__ cigar_off,
                                      const uint32_t* __restrict__ cigar, uint32_t* __restrict__ rel, uint8_t* __restrict__ ops8,
                                      int* __restrict__ flags) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  int lo = 0, hi = n_runs - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (run_first[mid] <= r) lo = mid; else hi = mid - 1;
  }
  const int64_t d = start[r] - start[run_first[lo]];
  if (d < 0 || d > 0xFFFFFFFFll) atomicOr(flags, 1);
  rel[r] = (uint32_t)d;
  const uint64_t len = seq_off[r + 1] - seq_off[r];
  const uint32_t c0 = cigar_off[r], ops = cigar_off[r + 1] - c0;
  if (ops > 255) atomicOr(flags, 2);
  if (len != seq_off[1] - seq_off[0]) atomicOr(flags, 4);
  const bool implicit = ops == 1 && cigar[c0] == ((uint32_t)len << 4) && len < (1u << 28);
  ops8[r] = implicit ? (uint8_t)0 : (uint8_t)ops;
}
__global__ void wire_cigar_kernel(int64_t n, const uint8_t* __restrict__ ops8, const uint32_t* __restrict__ cigar_off,
                                  const uint32_t* __restrict__ cigar, const int64_t* __restrict__ wire_off, uint32_t* __restrict__ wire) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  const int k = ops8[r];
  const uint32_t o = cigar_off[r];
  const int64_t w = wire_off[r];
  for (int j = 0; j < k; ++j) wire[w + j] = cigar[o + j];
}

__global__ void qual_histogram_kernel(uint64_t n_bases, const uint8_t* __restrict__ qual, unsigned int* __restrict__ present) {
  __shared__ unsigned int s_present[256];
  s_present[threadIdx.x] = 0;
  __syncthreads();
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_bases; i += (uint64_t)gridDim.x * blockDim.x)
    s_present[qual[i]] = 1;  // benign race: every writer stores 1
  __syncthreads();
  if (s_present[threadIdx.x]) present[threadIdx.x] = 1;
}

struct QualCodes {
  uint8_t code[256];
};

template <int BITS>
__global__ void pack_qual_kernel(uint64_t n_bases, const uint8_t* __restrict__ qual, QualCodes C, uint8_t* __restrict__ packed) {
  __shared__ uint8_t s_code[256];
  s_code[threadIdx.x] = C.code[threadIdx.x];  // 256 threads per block
  __syncthreads();
  constexpr int PER = 8 / BITS;  // bases per output byte, lowest bits = lowest base index
  const uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;  // output byte
  if (PER * b >= n_bases) return;
  uint32_t v = 0;
#pragma unroll
  for (int k = 0; k < PER; ++k)
    if (PER * b + k < n_bases) v |= (uint32_t)s_code[qual[PER * b + k]] << (BITS * k);
  packed[b] = (uint8_t)v;
}

// start-sortedness within each contig run, sample ids in range (SlidingWindow.scala:36-37)
__global__ void validate_kernel(int64_t n, const int64_t* __restrict__ start, const int32_t* __restrict__ contig_id,
                                const uint8_t* __restrict__ sample_id, int n_samples, int* __restrict__ bad) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  const int64_t s = start[r];
  if (s < 0 || s >= ((int64_t)1 << 39)) atomicOr(bad, 4);
  if (sample_id[r] >= n_samples) atomicOr(bad, 2);
  if (r > 0 && contig_id[r] == contig_id[r - 1] && s < start[r - 1]) atomicOr(bad, 1);
}


// ---- reference-relative bases (guac_read_batch_compact::mm_count / mm_words) ------------------------------------------

// 32 bits of a 2-bit stream starting at bit `bit` (the word after the last one of a buffer is padding, never garbage-sensitive:
// callers mask what they use)
__device__ __forceinline__ uint32_t bits32_at(const uint32_t* __restrict__ words, uint64_t bit) {
  const uint32_t* w = words + (bit >> 5);
  return __funnelshift_r(__ldg(w), __ldg(w + 1), (int)(bit & 31));
}

// Copies n 2-bit symbols from (src, symbol src_sym) to (dst, symbol dst_sym).  dst is zeroed beforehand; the words this run
// shares with its neighbours are merged with atomicOr, the words it owns are stored.
__device__ __forceinline__ void copy_symbols(uint32_t* __restrict__ dst, uint64_t dst_sym, const uint32_t* __restrict__ src, uint64_t src_sym, int n) {
  uint64_t dbit = 2 * dst_sym, sbit = 2 * src_sym;
  int left = 2 * n;  // bits
  if (left <= 0) return;
  const int head = (int)(dbit & 31);
  if (head) {  // up to the next word boundary of dst
    const int take = min(left, 32 - head);
    const uint32_t v = bits32_at(src, sbit) & (take == 32 ? 0xFFFFFFFFu : ((1u << take) - 1u));
    atomicOr(dst + (dbit >> 5), v << head);
    dbit += take;
    sbit += take;
    left -= take;
  }
  for (; left >= 32; left -= 32, dbit += 32, sbit += 32) dst[dbit >> 5] = bits32_at(src, sbit);
  if (left > 0) atomicOr(dst + (dbit >> 5), bits32_at(src, sbit) & ((1u << left) - 1u));
}

// per read: how many of its bases travel explicitly, and how many difference words it has
__global__ void refrel_lengths_kernel(int64_t n, const uint8_t* __restrict__ mm_count, const uint64_t* __restrict__ seq_off,
                                      uint16_t* __restrict__ expl_len, uint8_t* __restrict__ mm_n, int* __restrict__ bad) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  const uint64_t len = seq_off[r + 1] - seq_off[r];
  const bool expl = mm_count[r] == 255;
  expl_len[r] = expl ? (uint16_t)len : (uint16_t)0;
  mm_n[r] = expl ? (uint8_t)0 : mm_count[r];
  if (!expl && len >= 16384) atomicOr(bad, 8);  // positions are 14 bits
}

// The full 2-bit base column from the wire form, one thread per read of one contig: explicit reads from the explicit
// stream, the others from the contig's 2-bit reference window with their differences patched in.
__global__ void refrel_expand_kernel(int64_t read_lo, int64_t read_hi, const uint8_t* __restrict__ mm_count,
                                     const uint16_t* __restrict__ mm_words, const int64_t* __restrict__ mm_off,
                                     const uint32_t* __restrict__ expl, const int64_t* __restrict__ expl_off,
                                     const int64_t* __restrict__ start, const uint64_t* __restrict__ seq_off,
                                     const uint32_t* __restrict__ ref2, int64_t ref_offset, int64_t ref_len, uint32_t* __restrict__ bases2,
                                     int* __restrict__ bad) {
  const int64_t r = read_lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= read_hi) return;
  const uint64_t s0 = seq_off[r];
  const int len = (int)(seq_off[r + 1] - s0);
  if (mm_count[r] == 255) {
    copy_symbols(bases2, s0, expl, (uint64_t)expl_off[r], len);
    return;
  }
  const int64_t rel = start[r] - ref_offset;
  if (ref2 == nullptr || rel < 0 || rel + len > ref_len) {  // the sender promised a read inside the window
    atomicOr(bad, 16);
    return;
  }
  copy_symbols(bases2, s0, ref2, (uint64_t)rel, len);
  const int64_t m0 = mm_off[r];
  for (int k = 0; k < (int)mm_count[r]; ++k) {
    const uint32_t w = mm_words[m0 + k];
    const uint32_t pos = w >> 2, x = w & 3u;
    if ((int)pos >= len || x == 0) {
      atomicOr(bad, 32);
      continue;
    }
    const uint64_t bit = 2 * (s0 + pos);
    atomicXor(bases2 + (bit >> 5), x << (bit & 31));
  }
}

// Export side.  Which reads can travel as differences: one aligned operator over A C G T bases (RF_FAST, bit 0 of the flags
// byte of the packed header), a complete slot of the read pass (every record of such a read is a Mismatch), wholly on
// A C G T reference inside the window, shorter than 16384 bases.
__global__ void refrel_classify_kernel(int64_t read_lo, int64_t read_hi, const uint4* __restrict__ hdr, const uint8_t* __restrict__ ev_count,
                                       int ev_cap, const int64_t* __restrict__ start, const int64_t* __restrict__ end,
                                       const uint64_t* __restrict__ seq_off, int64_t ref_offset, int64_t ref_len,
                                       const uint32_t* __restrict__ ref_bad, uint8_t* __restrict__ mm_count, uint16_t* __restrict__ expl_len,
                                       uint8_t* __restrict__ mm_n) {
  const int64_t r = read_lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= read_hi) return;
  const uint64_t len = seq_off[r + 1] - seq_off[r];
  const int64_t s = start[r], e = end[r];
  const int cnt = ev_count ? (int)ev_count[r] : 255;
  bool ok = ((hdr[r].y >> 24) & 1u) && cnt <= ev_cap && len < 16384 && e > s && s >= ref_offset && e <= ref_offset + ref_len;
  if (ok && ref_bad != nullptr)
    for (int64_t g = (s - ref_offset) >> 4; ok && g <= (e - 1 - ref_offset) >> 4; ++g) ok = !((__ldg(ref_bad + (g >> 5)) >> (g & 31)) & 1u);
  mm_count[r] = ok ? (uint8_t)cnt : (uint8_t)255;
  expl_len[r] = ok ? (uint16_t)0 : (uint16_t)len;
  mm_n[r] = ok ? (uint8_t)cnt : (uint8_t)0;
}

// ... their difference words, from the read pass's records (x = offset | kind << 16 | read base code << 19)
__global__ void refrel_words_kernel(int64_t read_lo, int64_t read_hi, const uint8_t* __restrict__ mm_count, const int64_t* __restrict__ mm_off,
                                    const uint2* __restrict__ events, int ev_cap, const int64_t* __restrict__ start,
                                    const uint32_t* __restrict__ ref2, int64_t ref_offset, uint16_t* __restrict__ mm_words) {
  const int64_t r = read_lo + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= read_hi || mm_count[r] == 255) return;
  const int64_t rel = start[r] - ref_offset, m0 = mm_off[r];
  for (int k = 0; k < (int)mm_count[r]; ++k) {
    const uint32_t x = events[r * ev_cap + k].x;
    const uint32_t pos = x & 0xFFFFu, code = (x >> 19) & 3u;
    const uint64_t bit = 2 * (uint64_t)(rel + pos);
    const uint32_t ref_code = (__ldg(ref2 + (bit >> 5)) >> (bit & 31)) & 3u;
    mm_words[m0 + k] = (uint16_t)((pos << 2) | (ref_code ^ code));
  }
}

// ... and the explicit reads' bases, concatenated
__global__ void refrel_explicit_kernel(int64_t n, const uint8_t* __restrict__ mm_count, const int64_t* __restrict__ expl_off,
                                       const uint64_t* __restrict__ seq_off, const uint32_t* __restrict__ bases2, uint32_t* __restrict__ expl) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n || mm_count[r] != 255) return;
  copy_symbols(expl, (uint64_t)expl_off[r], bases2, seq_off[r], (int)(seq_off[r + 1] - seq_off[r]));
}

}  // namespace compact
}  // namespace guac
