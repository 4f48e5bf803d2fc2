// partition_kernels.cuh — coverage-aware ("capped") loci partitioning on the device:
//   coverage_kernel   CoverageRDD.traversalCoverage (readsets/rdd/CoverageRDD.scala:120-143) over
//                     ContigCoverageIterator (readsets/iterator/ContigCoverageIterator.scala:28-117): per eligible locus,
//                     Coverage(depth, starts) — a read counts towards `depth` at every eligible locus of
//                     [start - halfWindow, end + halfWindow) and towards `starts` at the first of them.  One thread per
//                     read: a difference array for the depths (prefix-summed afterwards), a counter for the starts.
//   take_loci_kernel  TakeLociIterator (loci/set/TakeLociIterator.scala:25-196): greedy loci sets of at most
//                     maxRegionsPerPartition regions, loci deeper than that dropped.  The walk is sequential by nature
//                     (one thread), but between two decisions it jumps: the regions a run of loci adds are a difference
//                     of prefix sums of `starts`, so the end of a run is a binary search.
// Reference file names are relative to src/main/scala/org/hammerlab/guacamole/.
#pragma once
#include "kernels.cuh"

namespace guac {

// slot of the first eligible locus >= `locus` on the contig of ranges [r0, r1) (sorted by start); n_slots if none
__device__ inline int64_t first_slot_at_or_after(const RangeDev* __restrict__ ranges, int r0, int r1, int64_t locus, int* range_out) {
  int lo = r0, hi = r1;
  while (lo < hi) {  // first range with end > locus
    const int mid = (lo + hi) >> 1;
    if (ranges[mid].end > locus) hi = mid; else lo = mid + 1;
  }
  *range_out = lo;
  if (lo >= r1) return -1;
  return ranges[lo].slot0 + max((long long)0, (long long)(locus - ranges[lo].start));
}

__global__ void coverage_kernel(ReadsView R, const int32_t* __restrict__ contig_id, const RangeDev* __restrict__ ranges, int n_ranges,
                                int half_window, int* __restrict__ depth_diff /* [n_slots + 1] */, int* __restrict__ starts /* [n_slots] */) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R.n_reads) return;
  const int c = contig_id[r];
  // the ranges of this read's contig are contiguous in the (sorted) table
  int a = 0, b = n_ranges;
  while (a < b) {
    const int m = (a + b) >> 1;
    if (ranges[m].contig_id < c) a = m + 1; else b = m;
  }
  const int r0 = a;
  b = n_ranges;
  while (a < b) {
    const int m = (a + b) >> 1;
    if (ranges[m].contig_id <= c) a = m + 1; else b = m;
  }
  const int r1 = a;
  if (r0 == r1) return;
  const int64_t lo = max((long long)0, (long long)(R.start[r] - half_window)), hi = R.end[r] + half_window;
  int ri;
  const int64_t first = first_slot_at_or_after(ranges, r0, r1, lo, &ri);
  if (first < 0 || ranges[ri].start >= hi) return;  // overlaps no eligible locus
  atomicAdd(starts + first, 1);
  for (; ri < r1 && ranges[ri].start < hi; ++ri) {
    const int64_t s = max((long long)lo, (long long)ranges[ri].start), e = min((long long)hi, (long long)ranges[ri].end);
    atomicAdd(depth_diff + ranges[ri].slot0 + (s - ranges[ri].start), 1);
    atomicAdd(depth_diff + ranges[ri].slot0 + (e - ranges[ri].start), -1);
  }
}

// (Slots are dense over the eligible ranges.  A read's -1 lands at most on the first slot of the next range, i.e. before
// any of that range's own +1 is read by the prefix sum: no correction is needed at range boundaries.)

__global__ void widen_kernel(const int32_t* __restrict__ in, int64_t n, int64_t* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = in[i];
}

struct TakeLociOut {
  int32_t partition, contig_id;
  int64_t start, end;
};

// One locus of the walk: slots with depth 0 are not part of the input (CoverageRDD emits covered loci only).
struct TakeLociArgs {
  const int64_t* depth;       // [n] inclusive prefix sums of the difference array = depth per slot
  const int* starts;          // [n]
  const int64_t* starts_incl; // [n] inclusive prefix sums of starts
  const int64_t* locus;       // [n] locus of the slot (explicit-coverage form), or nullptr: from the ranges
  const int32_t* contig;      // [n] contig of the slot (explicit form), or nullptr
  const RangeDev* ranges;     // range form: slots are the dense loci of these ranges
  int n_ranges;
  int64_t n;
  int max_regions;
  int trim_ranges;
  TakeLociOut* out;
  int max_out;
  int* n_out;
  int* n_partitions;
};

__global__ void take_loci_kernel(TakeLociArgs A) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  int range_i = 0;
  auto locus_of = [&](int64_t i) -> int64_t {
    if (A.locus) return A.locus[i];
    while (range_i + 1 < A.n_ranges && A.ranges[range_i + 1].slot0 <= i) ++range_i;
    while (range_i > 0 && A.ranges[range_i].slot0 > i) --range_i;
    return A.ranges[range_i].start + (i - A.ranges[range_i].slot0);
  };
  auto contig_of = [&](int64_t i) -> int {
    if (A.contig) return A.contig[i];
    locus_of(i);
    return A.ranges[range_i].contig_id;
  };
  // end (exclusive slot) of the contig that slot i lies on
  auto contig_end = [&](int64_t i) -> int64_t {
    const int c = contig_of(i);
    if (A.contig) {
      int64_t lo = i, hi = A.n;
      while (lo < hi) {
        const int64_t m = (lo + hi) >> 1;
        if (A.contig[m] == c) lo = m + 1; else hi = m;
      }
      return lo;
    }
    int k = range_i;
    while (k + 1 < A.n_ranges && A.ranges[k + 1].contig_id == c) ++k;
    return k + 1 < A.n_ranges ? A.ranges[k + 1].slot0 : A.n;
  };
  auto emit = [&](int part, int contig, int64_t s, int64_t e) {
    const int k = (*A.n_out)++;
    if (k < A.max_out) {
      A.out[k].partition = part;
      A.out[k].contig_id = contig;
      A.out[k].start = s;
      A.out[k].end = e;
    }
  };
  int64_t p = 0;  // next slot to look at
  int part = 0;
  bool part_has_loci = false;
  int cur_total = 0;  // regions of the loci set being built
  while (p < A.n) {
    // skip slots that are not in the input (no coverage)
    if (A.depth[p] == 0) {
      ++p;
      continue;
    }
    // takeLoci on the contig of p (:108-196)
    const int contig = contig_of(p);
    const int64_t cend = contig_end(p);
    const int budget = A.max_regions - cur_total;
    int cur = 0;
    bool took = false, interval_open = false;
    int64_t iv_start = 0, iv_end = 0;
    bool returned = false;  // takeLoci returned because a locus did not fit (it stays at the head)
    while (p < cend) {
      const int64_t dp = A.depth[p];
      if (dp == 0) {
        ++p;
        continue;
      }
      const long long fresh = cur == 0 ? dp : (long long)A.starts[p];
      if (cur + fresh > budget) {
        if (interval_open) {
          emit(part, contig, iv_start, iv_end);
          interval_open = false;
        }
        if (fresh > A.max_regions) {  // deeper than any loci set can hold: dropped (:163-168)
          ++p;
          continue;
        }
        returned = true;
        break;
      }
      cur += (int)fresh;
      took = true;
      const int64_t x = locus_of(p);
      if (interval_open && (!A.trim_ranges || x == iv_end)) {
        iv_end = x + 1;
      } else {
        if (interval_open) emit(part, contig, iv_start, iv_end);
        iv_start = x;
        iv_end = x + 1;
        interval_open = true;
      }
      ++p;
      // jump: the loci after this one add their `starts`; find how far they fit (prefix sums), and take them in one go
      if (!A.trim_ranges && p < cend) {
        const int64_t base = A.starts_incl[p - 1];
        const int64_t room = (int64_t)budget - cur;
        int64_t lo = p, hi = cend;  // first slot q in [p, cend) with starts_incl[q] - base > room
        while (lo < hi) {
          const int64_t m = (lo + hi) >> 1;
          if (A.starts_incl[m] - base > room) hi = m; else lo = m + 1;
        }
        if (lo > p) {  // slots [p, lo) fit; the last covered one among them ends the interval so far
          int64_t last = lo - 1;
          while (last >= p && A.depth[last] == 0) --last;
          if (last >= p) {
            cur += (int)(A.starts_incl[last] - base);
            iv_end = locus_of(last) + 1;
          }
          p = lo;
        }
      }
    }
    if (interval_open) emit(part, contig, iv_start, iv_end);
    if (took) {
      cur_total += cur;
      part_has_loci = true;
    }
    if (returned && !took) {
      // (:60-77) takeLoci gave nothing although the contig has loci left: the head locus does not fit in what remains of
      // this loci set — the set is finished and the locus opens the next one.  (After a takeLoci that DID take loci the
      // reference calls it once more on the same contig, with the remaining budget and the head locus counted at its full
      // depth — which may drop a too-deep locus and go on: that is simply the next turn of this loop.)
      if (part_has_loci) {
        ++part;
        part_has_loci = false;
        cur_total = 0;
      } else {
        ++p;  // (cannot happen: an empty set takes any locus of depth <= max; guards against a stall)
      }
    }
    // contig exhausted (p == cend): the loop goes on with the next contig in the same loci set
  }
  if (part_has_loci) ++part;
  *A.n_partitions = part;
}

}  // namespace guac
