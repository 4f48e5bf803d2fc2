// synth_kernels.cuh — CUDA twin of guacamole_b200/synth.py: fills HBM with the seeded synthetic
// reads and reference for the full-size benchmark (arithmetic in synth_model.h).
#pragma once
#include <cuda_runtime.h>

#include "synth_model.h"

namespace guac {
namespace synth {

__global__ void reference_kernel(uint64_t seed, int64_t n, uint8_t* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (uint8_t)("ACGT"[rnd_k(stream_key(seed, S_REF), (uint64_t)i) & 3]);
}

__global__ void read_headers_kernel(Params P, int64_t* __restrict__ start, int32_t* __restrict__ contig_id, int32_t contig,
                                    uint8_t* __restrict__ mapq, uint8_t* __restrict__ flags, uint8_t* __restrict__ sample,
                                    uint64_t* __restrict__ seq_off, uint8_t* __restrict__ n_ops) {
  int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j > P.n_reads) return;
  seq_off[j] = (uint64_t)j * READ_LEN;
  if (j == P.n_reads) return;
  ReadHeader h = read_header(P, j);
  start[j] = h.start;
  contig_id[j] = contig;
  mapq[j] = h.mapq;
  flags[j] = h.flags;
  sample[j] = (uint8_t)h.sample;
  n_ops[j] = (uint8_t)h.n_ops;
}

__global__ void read_cigars_kernel(Params P, const int64_t* __restrict__ cigar_off64, uint32_t* __restrict__ cigar_off,
                                   uint32_t* __restrict__ cigar) {
  int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (j > P.n_reads) return;
  cigar_off[j] = (uint32_t)cigar_off64[j];
  if (j == P.n_reads) return;
  ReadHeader h = read_header(P, j);
  write_cigar(h, cigar + cigar_off64[j]);
}

// one thread per 4 bases: one byte of bases2, one 32-bit store of qualities
__global__ void read_bases_kernel(Params P, uint8_t* __restrict__ bases2, uint8_t* __restrict__ qual) {
  int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // group of 4 bases
  const int64_t n_groups = P.n_reads * (READ_LEN / 4);
  if (g >= n_groups) return;
  const int64_t j = g / (READ_LEN / 4);
  const int k0 = (int)(g - j * (READ_LEN / 4)) * 4;
  const ReadHeader h = read_header(P, j);
  uint32_t q4 = 0;
  uint8_t b = 0;
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    int code;
    uint8_t q;
    base_and_qual(P, h, j, k0 + t, code, q);
    b |= (uint8_t)(code << (2 * t));
    q4 |= (uint32_t)q << (8 * t);
  }
  bases2[g] = b;
  reinterpret_cast<uint32_t*>(qual)[g] = q4;
}

}  // namespace synth
}  // namespace guac
