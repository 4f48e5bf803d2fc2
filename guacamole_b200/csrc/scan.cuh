// scan.cuh — hand-written device-wide scans (three-phase: block reduce, scan of block
// totals, block scan + carry).  Warp-level shuffles inside a block; used for the running
// max of read ends, for compacting visited loci and for per-locus element offsets.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace guac {

struct OpSum {
  template <class T>
  __device__ __forceinline__ T operator()(T a, T b) const { return a + b; }
  template <class T>
  __device__ __forceinline__ static T identity() { return T(0); }
};
struct OpMax {
  template <class T>
  __device__ __forceinline__ T operator()(T a, T b) const { return a > b ? a : b; }
  template <class T>
  __device__ __forceinline__ static T identity() { return T(INT64_MIN); }
};

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

template <class T, class Op>
__device__ __forceinline__ T warp_inclusive(T v, Op op) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    T o = __shfl_up_sync(0xffffffffu, v, d);
    if ((threadIdx.x & 31) >= d) v = op(o, v);
  }
  return v;
}

// Block-wide inclusive scan of one value per thread; returns the inclusive value and the
// block total through `total`.
template <class T, class Op>
__device__ __forceinline__ T block_inclusive(T v, Op op, T* smem /*[33]*/, T& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  T inc = warp_inclusive(v, op);
  if (lane == 31) smem[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    T w = lane < (SCAN_THREADS / 32) ? smem[lane] : Op::template identity<T>();
    T ws = warp_inclusive(w, op);
    smem[lane] = ws;
  }
  __syncthreads();
  if (warp > 0) inc = op(smem[warp - 1], inc);
  total = smem[SCAN_THREADS / 32 - 1];
  __syncthreads();
  return inc;
}

// phase 1: per-tile totals
template <class TIn, class T, class Op>
__global__ void scan_reduce_kernel(const TIn* __restrict__ in, T* __restrict__ tile_totals, int64_t n, Op op) {
  __shared__ T smem[33];
  const int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
  T acc = Op::template identity<T>();
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    int64_t i = base + (int64_t)k * SCAN_THREADS + threadIdx.x;
    if (i < n) acc = op(acc, (T)in[i]);
  }
  T total;
  block_inclusive(acc, op, smem, total);
  if (threadIdx.x == 0) tile_totals[blockIdx.x] = total;
}

// phase 3: scan within the tile, seeded by the exclusive prefix of tile totals.
// EXCLUSIVE: out[i] = op over in[0..i); else inclusive.
template <bool EXCLUSIVE, class TIn, class T, class Op>
__global__ void scan_apply_kernel(const TIn* __restrict__ in, T* __restrict__ out, const T* __restrict__ tile_prefix,
                                  int64_t n, Op op) {
  __shared__ T smem[33];
  const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
  T v[SCAN_ITEMS];
  T acc = Op::template identity<T>();
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    int64_t i = base + k;
    v[k] = i < n ? (T)in[i] : Op::template identity<T>();
    acc = op(acc, v[k]);
  }
  T total;
  T inc = block_inclusive(acc, op, smem, total);
  // exclusive prefix of this thread = inclusive of previous thread
  T prev = __shfl_up_sync(0xffffffffu, inc, 1);
  __shared__ T warp_last[SCAN_THREADS / 32];
  if ((threadIdx.x & 31) == 31) warp_last[threadIdx.x >> 5] = inc;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) prev = threadIdx.x == 0 ? Op::template identity<T>() : warp_last[(threadIdx.x >> 5) - 1];
  T carry = op(tile_prefix[blockIdx.x], prev);  // tile_prefix[0] is the identity
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    int64_t i = base + k;
    if (EXCLUSIVE) {
      if (i < n) out[i] = carry;
      carry = op(carry, v[k]);
    } else {
      carry = op(carry, v[k]);
      if (i < n) out[i] = carry;
    }
  }
}

// single-block exclusive scan of up to a few thousand tile totals (loops if more)
template <class T, class Op>
__global__ void scan_totals_kernel(T* __restrict__ totals, int64_t n, Op op) {
  __shared__ T smem[33];
  T carry = Op::template identity<T>();
  for (int64_t base = 0; base < n; base += SCAN_THREADS) {
    int64_t i = base + threadIdx.x;
    T v = i < n ? totals[i] : Op::template identity<T>();
    T total;
    T inc = block_inclusive(v, op, smem, total);
    T prev = __shfl_up_sync(0xffffffffu, inc, 1);
    __shared__ T warp_last[SCAN_THREADS / 32];
    if ((threadIdx.x & 31) == 31) warp_last[threadIdx.x >> 5] = inc;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) prev = threadIdx.x == 0 ? Op::template identity<T>() : warp_last[(threadIdx.x >> 5) - 1];
    if (i < n) totals[i] = op(carry, prev);
    carry = op(carry, total);
    __syncthreads();
  }
  if (threadIdx.x == 0) totals[n] = carry;  // grand total in the extra slot
}

}  // namespace guac
