// synth_model.h — arithmetic of the synthetic-read generator (twin of guacamole_b200/synth.py,
// SURVEY.md §8(d)), shared by the CUDA kernels (synth_kernels.cuh) and the oracle's C twin.
// Fills HBM with the seeded synthetic reads and reference for the full-size benchmark; every
// value is a pure function of (seed, read index, base index), bit-identical to the numpy
// and C twins, so the CPU oracle can be run on any window of the same data.
#pragma once
#include <stdint.h>

#include "../../include/guac_b200.h"

#ifndef __CUDACC__
#ifndef __host__
#define __host__
#define __device__
#endif
#ifndef __forceinline__
#define __forceinline__ inline
#endif
#endif

namespace guac {
namespace synth {

constexpr int READ_LEN = 100;
constexpr int MARGIN = 120;
enum : uint64_t { S_REF = 1, S_START = 2, S_READ = 3, S_BASE = 4, S_QUAL = 5, S_ERR = 6, S_SITE = 7, S_INS = 8, S_DENS = 9, S_TAIL = 10 };

__host__ __device__ __forceinline__ uint64_t mix(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ uint64_t stream_key(uint64_t seed, uint64_t stream) { return mix(seed ^ (stream << 32)); }
__host__ __device__ __forceinline__ uint64_t rnd_k(uint64_t key, uint64_t a, uint64_t b = 0) {
  return mix(mix(key ^ (a * 0x9E3779B97F4A7C15ull)) ^ (b * 0xC2B2AE3D27D4EB4Full));
}

struct Params {
  uint64_t seed;
  int64_t contig_len;
  int64_t n_reads;     // total reads of the batch
  int32_t n_samples;
  int32_t pattern_len;
  uint8_t pattern[64];   // sample id of read j = pattern[j % pattern_len]
  uint8_t is_tumor[16];  // per sample
  // 0: the workload of SURVEY.md 8(d) (stratified starts, qualities 40/30/20/10, 3 % multi-operator reads).
  // 1: the "realistic" variant: read density modulated in 256-locus blocks (depth varies like a Poisson pileup's),
  //    15 % of reads end in a run of 5..30 quality-2 bases (10 % errors there), 10 % multi-operator reads, 1 % of reads
  //    with a 15 % error rate over their whole length (10..20 mismatches).
  int32_t profile;
  int32_t _pad;
};

// The per-read header shared by all kernels (also the layout decisions: CIGAR kind, lengths).
struct ReadHeader {
  int64_t start;
  int sample, hap, kind /*0 plain 1 ins 2 del 3 clip*/, a, b, s, n_ops;
  int tail, noisy;  // profile 1: bases of the quality-2 tail (0: none); the whole read is error-prone
  uint8_t mapq, flags;
};

__host__ __device__ __forceinline__ ReadHeader read_header(const Params& P, int64_t j) {
  ReadHeader h;
  const uint64_t ju = (uint64_t)j;
  h.sample = P.pattern[j % P.pattern_len];
  const int64_t span = P.contig_len - MARGIN;
  const int64_t u16 = (int64_t)(rnd_k(stream_key(P.seed, S_START), ju) & 0xFFFF);
  // floor((j * 65536 + u16) * span / (N * 65536)) without overflow: j * span = q * N + rem
  const int64_t js = j * span, q = js / P.n_reads, rem = js - q * P.n_reads;
  h.start = q + (rem * 65536 + u16 * span) / (P.n_reads * 65536);
  if (P.profile == 1) {
    // density modulation, monotone in the start (so the batch stays sorted): inside every 256-locus block the first half
    // receives 128 + m of the block's 256 "slots" of reads, m in [-30, 30] from the block's hash
    const int64_t blk = h.start >> 8;
    if (((blk + 1) << 8) <= span) {
      const int64_t t = h.start & 255;
      const int64_t cut = 128 + (int64_t)(rnd_k(stream_key(P.seed, S_DENS), (uint64_t)blk) % 61) - 30;
      h.start = (blk << 8) + (t < cut ? t * 128 / cut : 128 + (t - cut) * 128 / (256 - cut));
    }
  }
  const uint64_t hr = rnd_k(stream_key(P.seed, S_READ), ju);
  h.flags = (uint8_t)((hr & 1 ? GUAC_READ_POSITIVE_STRAND : 0) | GUAC_READ_PAIRED);
  h.hap = (int)((hr >> 1) & 1);
  const int mq = (int)((hr >> 8) % 100);
  h.mapq = mq < 97 ? 60 : (mq < 99 ? 20 : 0);
  const int ck = (int)((hr >> 16) % 100);
  h.b = (int)((hr >> 24) % 10) + 1;
  h.kind = ck == 97 ? 1 : ck == 98 ? 2 : ck == 99 ? 3 : 0;
  h.tail = h.noisy = 0;
  if (P.profile == 1) {
    h.kind = ck < 90 ? 0 : ck < 94 ? 1 : ck < 98 ? 2 : 3;  // 4 % insertions, 4 % deletions, 2 % soft clips
    h.noisy = ((hr >> 48) % 100) == 0;
    const uint64_t ht = rnd_k(stream_key(P.seed, S_TAIL), ju);
    if (((ht >> 32) % 100) < 15) h.tail = 5 + (int)(ht % 26);
  }
  h.a = h.kind == 1 ? 5 + (int)((hr >> 32) % (uint64_t)(91 - h.b)) : 5 + (int)((hr >> 32) % 91);
  h.s = (int)((hr >> 40) % 20) + 1;
  h.n_ops = (h.kind == 1 || h.kind == 2) ? 3 : (h.kind == 3 ? 2 : 1);
  return h;
}

// reference position of base k of the read, or -1 for inserted / clipped bases
__host__ __device__ __forceinline__ int64_t base_position(const ReadHeader& h, int k) {
  switch (h.kind) {
    case 1: return k < h.a ? h.start + k : (k < h.a + h.b ? -1 : h.start + k - h.b);
    case 2: return k < h.a ? h.start + k : h.start + k + h.b;
    case 3: return k < h.s ? -1 : h.start + k - h.s;
    default: return h.start + k;
  }
}

// (2-bit base code, quality) of base k of read j
__host__ __device__ __forceinline__ void base_and_qual(const Params& P, const ReadHeader& h, int64_t j, int k, int& code,
                                                       uint8_t& qual) {
  const int64_t pos = base_position(h, k);
  const uint64_t gi = (uint64_t)j * READ_LEN + (uint64_t)k;
  if (pos >= 0) {
    const int ref_code = (int)(rnd_k(stream_key(P.seed, S_REF), (uint64_t)pos) & 3);
    code = ref_code;
    const uint64_t hs = rnd_k(stream_key(P.seed, S_SITE), (uint64_t)pos);
    if (hs % 1000 == 0) {  // germline site: 2/3 het, 1/3 hom
      const bool hom = ((hs >> 20) % 3) == 2;
      if (hom || h.hap == 1) code = (ref_code + (int)((hs >> 24) % 3) + 1) & 3;
    }
    if (((hs >> 32) % 10000) == 0 && P.is_tumor[h.sample]) {  // somatic site
      const int vsel = (int)((hs >> 48) % 3);
      const int64_t vthr = vsel == 0 ? 6554 : (vsel == 1 ? 16384 : 32768);
      const int64_t draw = (int64_t)(rnd_k(stream_key(P.seed, S_SITE), (uint64_t)j, (uint64_t)pos + 1) & 0xFFFF);
      if (draw < vthr) code = (ref_code + (int)((hs >> 52) % 3) + 1) & 3;
    }
  } else {
    code = (int)(rnd_k(stream_key(P.seed, S_INS), gi) & 3);
  }
  const int hq = (int)(rnd_k(stream_key(P.seed, S_QUAL), gi) % 100);
  qual = hq < 70 ? 40 : (hq < 90 ? 30 : (hq < 98 ? 20 : 10));
  const uint64_t he = rnd_k(stream_key(P.seed, S_ERR), gi);
  int64_t thr = qual == 40 ? 1678 : (qual == 30 ? 16777 : (qual == 20 ? 167772 : 1677722));
  if (k >= READ_LEN - h.tail) {  // the quality-2 tail: 10 % errors
    qual = 2;
    thr = 1677722;
  }
  if (h.noisy) thr = thr > 2516582 ? thr : 2516582;  // 15 % errors
  if ((int64_t)(he & 0xFFFFFF) < thr) code = (code + (int)((he >> 24) % 3) + 1) & 3;
}

__host__ __device__ __forceinline__ void write_cigar(const ReadHeader& h, uint32_t* c) {
  switch (h.kind) {
    case 1:
      c[0] = ((uint32_t)h.a << 4) | 0;
      c[1] = ((uint32_t)h.b << 4) | 1;
      c[2] = ((uint32_t)(READ_LEN - h.a - h.b) << 4) | 0;
      break;
    case 2:
      c[0] = ((uint32_t)h.a << 4) | 0;
      c[1] = ((uint32_t)h.b << 4) | 2;
      c[2] = ((uint32_t)(READ_LEN - h.a) << 4) | 0;
      break;
    case 3:
      c[0] = ((uint32_t)h.s << 4) | 4;
      c[1] = ((uint32_t)(READ_LEN - h.s) << 4) | 0;
      break;
    default: c[0] = ((uint32_t)READ_LEN << 4) | 0;
  }
}


}  // namespace synth
}  // namespace guac
