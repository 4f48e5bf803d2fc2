// kernels.cuh — per-locus kernels, generic ("gather") formulation: one warp per locus, lanes
// stride over the reads that can overlap the locus's tile.  These serve the element-materialising
// operator (arbitrary host closures), the counting closure, and the first fused somatic-standard
// path.  The streaming tile kernel of tile_kernels.cuh replaces them on the benchmarked path.
#pragma once
#include "device_model.cuh"

namespace guac {

constexpr int TILE_LOCI = 64;
constexpr int WARPS_PER_CTA = 8;
constexpr int CTA_THREADS = WARPS_PER_CTA * 32;

__constant__ double c_success_probability[256];  // ADAM PhredUtils.phredToSuccessProbability table
__constant__ float c_odds_minus_one[256];         // p / (1 - p) - 1 of the same table, rounded up (LOD upper bound of the tile kernel)

// One range of loci as the planner hands it to the device.
struct RangeDev {
  int64_t start, end;
  int64_t slot0;     // dense slot of `start`
  int64_t tile0;     // index of the range's first tile
  int64_t read_clo;  // reads of this contig are [read_clo, read_chi)
  int64_t read_chi;
  int32_t contig_id;
  int32_t ref_index;
};

// One thread per tile: locate the tile's range, then binary-search the span of reads that can
// overlap it — start < tile end (reads are sorted by start) and running-max end > tile start.
__global__ void build_tiles_kernel(const RangeDev* __restrict__ ranges, int n_ranges, int64_t n_tiles, int tile_loci,
                                   ReadsView R, Tile* __restrict__ tiles, const RefView* __restrict__ refs = nullptr) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_tiles) return;
  int lo = 0, hi = n_ranges - 1;
  while (lo < hi) {  // last range with tile0 <= t
    int mid = (lo + hi + 1) >> 1;
    if (ranges[mid].tile0 <= t) lo = mid; else hi = mid - 1;
  }
  const RangeDev rg = ranges[lo];
  Tile tl;
  int64_t k = t - rg.tile0;
  tl.locus0 = rg.start + k * tile_loci;
  tl.n_loci = (int32_t)min((long long)tile_loci, (long long)(rg.end - tl.locus0));
  tl.contig_id = rg.contig_id;
  tl.slot0 = rg.slot0 + k * tile_loci;
  tl.ref_index = rg.ref_index;
  // _pad = 1: some locus of the tile has no A C G T reference base (outside the window, or flagged in the window's
  // one-bit-per-16-loci map): the locus pass of the stream form hands such a tile to the generic kernel without looking
  tl._pad = 0;
  if (refs != nullptr) {
    const RefView ref = refs[rg.ref_index];
    if (ref.bad != nullptr) {
      const int64_t rel0 = tl.locus0 - ref.offset;
      bool bad = rel0 < 0 || rel0 + tl.n_loci > ref.len;
      if (!bad) {
        const int64_t g0 = rel0 >> 4, g1 = (rel0 + tl.n_loci - 1) >> 4;
        for (int64_t w = g0 >> 5; w <= (g1 >> 5) && !bad; ++w) {
          uint32_t bits = __ldg(ref.bad + w);
          if (w == (g0 >> 5)) bits &= 0xFFFFFFFFu << (g0 & 31);
          if (w == (g1 >> 5)) bits &= 0xFFFFFFFFu >> (31 - (g1 & 31));
          bad = bits != 0;
        }
      }
      tl._pad = bad ? 1 : 0;
    }
  }
  const int64_t tile_end = tl.locus0 + tl.n_loci;
  int64_t a = rg.read_clo, b = rg.read_chi;
  while (a < b) {  // first read with start >= tile_end
    int64_t m = (a + b) >> 1;
    if (__ldg(R.start + m) < tile_end) a = m + 1; else b = m;
  }
  tl.read_hi = a;
  const int64_t key = ((int64_t)rg.contig_id << 40) | tl.locus0;
  a = rg.read_clo;
  b = tl.read_hi;
  while (a < b) {  // first read whose running-max end exceeds locus0
    int64_t m = (a + b) >> 1;
    if (__ldg(R.runmax_key + m) > key) b = m; else a = m + 1;
  }
  tl.read_lo = a;
  tl.seq_base = tl.read_hi > tl.read_lo ? __ldg(R.seq_off + tl.read_lo) : 0;
  tl.seq_end = tl.read_hi > tl.read_lo ? __ldg(R.seq_off + tl.read_hi) : 0;
  tiles[t] = tl;
}

// Offsets of a caller-built batch (guac_reads_upload): non-decreasing and inside the uploaded streams, and the
// CIGAR's sequenced length equal to the read's base count — a batch that breaks either would make the kernels
// index outside the read's own words / bases.  bad: bit 0 offsets, bit 1 lengths.
__global__ void validate_offsets_kernel(int64_t n, const uint32_t* __restrict__ cigar_off, const uint64_t* __restrict__ seq_off,
                                        uint32_t n_cigar, uint64_t n_bases, int* __restrict__ bad) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  if (cigar_off[r + 1] < cigar_off[r] || cigar_off[r + 1] > n_cigar || seq_off[r + 1] < seq_off[r] || seq_off[r + 1] > n_bases)
    atomicOr(bad, 1);
}
__global__ void validate_lengths_kernel(int64_t n, const uint32_t* __restrict__ cigar_off, const uint32_t* __restrict__ cigar,
                                        const uint64_t* __restrict__ seq_off, int* __restrict__ bad) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  uint64_t len = 0;
  for (uint32_t c = cigar_off[r]; c < cigar_off[r + 1]; ++c) {
    const uint32_t w = cigar[c];
    if (consumes_read(w & 15)) len += w >> 4;
  }
  if (len != seq_off[r + 1] - seq_off[r]) atomicOr(bad, 2);
}

// read.end = start + getPaddedReferenceLength (MappedRead.scala:75): M D N = X and P count.
__global__ void read_ends_kernel(int64_t n, const int64_t* __restrict__ start, const int32_t* __restrict__ contig_id,
                                 const uint32_t* __restrict__ cigar_off, const uint32_t* __restrict__ cigar,
                                 int64_t* __restrict__ end, int64_t* __restrict__ key) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  int64_t len = 0;
  for (uint32_t c = cigar_off[r]; c < cigar_off[r + 1]; ++c) {
    uint32_t w = cigar[c];
    int op = w & 15;
    if (consumes_ref(op) || op == OPC_P) len += w >> 4;
  }
  int64_t e = start[r] + len;
  end[r] = e;
  key[r] = ((int64_t)contig_id[r] << 40) | e;
}

// ASCII reference -> stays ASCII (1 B/locus); nothing to convert.  Present so that the
// reference upload is a single kernel-free memcpy.

// ---- counting closure: Pileup.depth / referenceDepth / positiveDepth + alignment classes ------
__global__ void __launch_bounds__(CTA_THREADS)
pileup_counts_kernel(ReadsView R, const RefView* __restrict__ refs, const Tile* __restrict__ tiles, int n_samples,
                     uint8_t* __restrict__ visited, guac_locus_counts* __restrict__ counts, DeviceError* err) {
  const Tile tile = tiles[blockIdx.x];
  const RefView ref = refs[tile.ref_index];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int li = warp; li < tile.n_loci; li += WARPS_PER_CTA) {
    const int64_t x = tile.locus0 + li, slot = tile.slot0 + li;
    bool any = false;
    for (int64_t base = tile.read_lo; base < tile.read_hi; base += 32) {
      const int64_t r = base + lane;
      bool ov = false;
      if (r < tile.read_hi) ov = __ldg(R.start + r) <= x && x < __ldg(R.end + r);
      int smp = ov ? (int)__ldg(R.sample_id + r) : 0;
      if (ov && smp >= n_samples) ov = false;
      if (ov) {
        Elem e;
        int ec = element_at(R, r, x, ref, e);
        if (ec != DERR_NONE) {
          report_error(err, ec, tile.contig_id, x, r);
        } else {
          uint32_t* c = reinterpret_cast<uint32_t*>(counts + slot * n_samples + smp);
          atomicAdd(c + 0, 1u);
          if (e.kind == GUAC_ALN_MATCH) atomicAdd(c + 1, 1u);
          if (__ldg(R.flags + r) & GUAC_READ_POSITIVE_STRAND) atomicAdd(c + 2, 1u);
          if (e.kind != GUAC_ALN_MATCH) atomicAdd(c + 2 + e.kind, 1u);  // mismatches.. clipped
        }
      }
      any = any || __any_sync(0xffffffffu, ov);
    }
    if (lane == 0) visited[slot] = any ? 1 : 0;
  }
}

// ---- element materialisation: per-locus totals, then records in read order ------------------
__global__ void __launch_bounds__(CTA_THREADS)
depth_total_kernel(ReadsView R, const Tile* __restrict__ tiles, int n_samples, uint8_t* __restrict__ visited,
                   int32_t* __restrict__ totals) {
  const Tile tile = tiles[blockIdx.x];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int li = warp; li < tile.n_loci; li += WARPS_PER_CTA) {
    const int64_t x = tile.locus0 + li, slot = tile.slot0 + li;
    int n = 0;
    for (int64_t base = tile.read_lo; base < tile.read_hi; base += 32) {
      const int64_t r = base + lane;
      bool ov = false;
      if (r < tile.read_hi) ov = __ldg(R.start + r) <= x && x < __ldg(R.end + r) && __ldg(R.sample_id + r) < n_samples;
      n += __popc(__ballot_sync(0xffffffffu, ov));
    }
    if (lane == 0) {
      totals[slot] = n;
      visited[slot] = n > 0;
    }
  }
}

// elements of visited locus i go to out[elem_off[vis_idx[slot]] ...), ordered by read index
__global__ void __launch_bounds__(CTA_THREADS)
write_elements_kernel(ReadsView R, const RefView* __restrict__ refs, const Tile* __restrict__ tiles, int n_samples,
                      int skip_empty, const int64_t* __restrict__ elem_off /*per slot, exclusive*/,
                      guac_element* __restrict__ out, DeviceError* err) {
  const Tile tile = tiles[blockIdx.x];
  const RefView ref = refs[tile.ref_index];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int li = warp; li < tile.n_loci; li += WARPS_PER_CTA) {
    const int64_t x = tile.locus0 + li, slot = tile.slot0 + li;
    int64_t o = elem_off[slot];
    for (int64_t base = tile.read_lo; base < tile.read_hi; base += 32) {
      const int64_t r = base + lane;
      bool ov = false;
      if (r < tile.read_hi) ov = __ldg(R.start + r) <= x && x < __ldg(R.end + r) && __ldg(R.sample_id + r) < n_samples;
      unsigned m = __ballot_sync(0xffffffffu, ov);
      if (ov) {
        Elem e;
        int ec = element_at(R, r, x, ref, e);
        if (ec != DERR_NONE) report_error(err, ec, tile.contig_id, x, r);
        guac_element g;
        g.read_index = r;
        g.read_position = e.read_position;
        g.cigar_element_index = e.cigar_element_index;
        g.cigar_element_locus = e.cigar_element_locus;
        g.index_within_cigar_element = e.index_within_cigar_element;
        g.kind = e.kind;
        g.base = e.base;
        g.quality = e.quality;
        g.sample_id = __ldg(R.sample_id + r);
        g.indel_length = e.indel_length;
        g._pad = 0;
        out[o + __popc(m & ((1u << lane) - 1))] = g;
      }
      o += __popc(m);
    }
  }
}

// slot -> (contig, locus) for visited loci, at their compacted index
__global__ void compact_loci_kernel(const Tile* __restrict__ tiles, const uint8_t* __restrict__ visited,
                                    const int64_t* __restrict__ vis_idx, int32_t* __restrict__ out_contig,
                                    int64_t* __restrict__ out_locus) {
  const Tile tile = tiles[blockIdx.x];
  for (int li = threadIdx.x; li < tile.n_loci; li += blockDim.x) {
    int64_t slot = tile.slot0 + li;
    if (visited[slot]) {
      int64_t d = vis_idx[slot];
      out_contig[d] = tile.contig_id;
      out_locus[d] = tile.locus0 + li;
    }
  }
}

template <class T>
__global__ void gather_rows_kernel(const T* __restrict__ in, const uint8_t* __restrict__ visited,
                                   const int64_t* __restrict__ vis_idx, int64_t n_slots, int row, T* __restrict__ out) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_slots * row) return;
  int64_t slot = i / row;
  if (visited[slot]) out[vis_idx[slot] * row + (i - slot * row)] = in[i];
}

// per-slot element offsets for the compacted (visited-only or all) loci: elem_cnt -> exclusive scan done by caller.
__global__ void gather_offsets_kernel(const int64_t* __restrict__ elem_off, const uint8_t* __restrict__ visited,
                                      const int64_t* __restrict__ vis_idx, int64_t n_slots, int64_t* __restrict__ out) {
  int64_t slot = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= n_slots) return;
  if (visited[slot]) out[vis_idx[slot]] = elem_off[slot];
}

// ---- somatic-standard (SomaticStandardCaller.scala:155-253), generic gather formulation ------
struct SSParamsDev {
  int normal_sample, tumor_sample, n_samples;
  int normal_odds, tumor_odds;
  int max_read_depth, max_normal_alt;
  int min_alt_reads;          // germline / joint-prefilter modes of the tile kernel
  unsigned int sample_mask;   // joint-prefilter mode: samples whose reads can start an allele (the discovery samples)
  double min_vaf;  // float widened to double (SomaticStandardCaller.scala:97,189)
};

struct SSCallDev {
  int64_t slot;
  int64_t locus;
  int64_t read_lo, read_hi;  // span of the locus's tile
  uint64_t key;              // allele_key of the called allele
  int64_t rep_read;          // first supporting read (lowest index)
  int32_t contig_id, ref_index;
  double tumor_ref_lp, tumor_alt_lp, normal_ref_lp, normal_alt_lp;
  // filled by ss_evidence_kernel
  uint32_t ref_off, ref_len, alt_off, alt_len;
  guac_allele_evidence tumor_ev, normal_ev;
};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
  return v;
}

// breeze logNormalize over two values (likelihood/Likelihood.scala:136-143)
__device__ __forceinline__ void log_normalize2(double a, double b, double& na, double& nb) {
  double mx = fmax(a, b);
  double sm;
  if (isinf(mx)) {
    sm = mx;
  } else {
    double acc = 0.0;
    acc += exp(a - mx);
    acc += exp(b - mx);
    sm = mx + log(acc);
  }
  na = a - sm;
  nb = b - sm;
}


// First index in [lo, hi) where the monotone predicate turns true (hi if it never does), by the whole warp: every round
// the lanes probe 32 evenly spaced indices and one ballot narrows the bracket 32-fold.
template <class Pred>
__device__ __forceinline__ int64_t warp_partition_point(int64_t lo, int64_t hi, int lane, Pred pred) {
  while (lo < hi) {
    const int64_t step = (hi - lo + 31) >> 5;
    const int64_t i = lo + lane * step;
    const unsigned m = __ballot_sync(0xffffffffu, i >= hi || pred(i));
    if (m & 1u) return lo;
    if (m == 0) {
      lo += 31 * step + 1;
    } else {
      const int L = __ffs(m) - 1;
      hi = min(hi, lo + L * step);
      lo += (L - 1) * step + 1;
    }
  }
  return lo;
}

// Read r at locus x for the per-locus kernels: overlap, sample, and (when `want`) the pileup element.  With the packed
// headers of the batch (ReadsView::hdr) one 16-byte load answers the first two, and a read of one aligned operator over
// standard bases (bit 0 of the flags byte, RF_FAST) has its element at read position x - start: no cursor.
struct ReadAtLocus {
  bool overlaps;
  int sample;
  int error;  // DERR_* of the element (DERR_NONE when not wanted)
};
template <class Want>
__device__ __forceinline__ ReadAtLocus read_at_locus(const ReadsView& R, int64_t r, int64_t x, const RefView& ref, Elem& e, Want want) {
  ReadAtLocus out;
  out.error = DERR_NONE;
  if (R.hdr == nullptr) {
    out.overlaps = __ldg(R.start + r) <= x && x < __ldg(R.end + r);
    out.sample = out.overlaps ? (int)__ldg(R.sample_id + r) : -1;
    if (out.overlaps && want(out.sample)) out.error = element_at(R, r, x, ref, e);
    return out;
  }
  const uint4 h = __ldg(R.hdr + r);
  const int span = (int)(h.y & 0xFFFF);
  const int off = (int)((uint32_t)x - h.x);  // exact: reads and loci of a call lie within 2^31 of each other on a contig
  out.overlaps = span != 0xFFFF ? (off >= 0 && off < span) : (__ldg(R.start + r) <= x && x < __ldg(R.end + r));
  out.sample = out.overlaps ? (int)((h.y >> 16) & 0xFF) : -1;
  if (!(out.overlaps && want(out.sample))) return out;
  const uint8_t rb = ref.at(x);
  if (((h.y >> 24) & 1u) && rb != 0) {
    const uint64_t gi = ((uint64_t)h.z | ((uint64_t)h.w << 32)) + (uint64_t)off;
    e.read_position = off;
    e.indel_length = 0;
    e.ref_base = rb;
    e.base = (uint8_t)"ACGT"[(__ldg(R.bases2 + (gi >> 2)) >> ((gi & 3) * 2)) & 3];
    e.kind = e.base == rb ? GUAC_ALN_MATCH : GUAC_ALN_MISMATCH;
    e.quality = __ldg(R.qual + gi);
    e.followed_by_ref_skip = false;
  } else {
    out.error = element_at(R, r, x, ref, e);
  }
  return out;
}

__global__ void __launch_bounds__(CTA_THREADS, 3)
somatic_standard_kernel(ReadsView R, const RefView* __restrict__ refs, const Tile* __restrict__ tiles,
                        const int* __restrict__ tile_list /* nullptr: identity */, int n_tiles,
                        int tile_per_warp /* 0: the CTA's warps share tile blockIdx.x; 1: one (single-locus) tile per warp */,
                        SSParamsDev P, uint8_t* __restrict__ visited, guac_locus_summary* __restrict__ summary,
                        SSCallDev* __restrict__ calls, int* __restrict__ n_calls, int max_calls,
                        unsigned long long* __restrict__ n_visited /* optional */, DeviceError* err,
                        const int* __restrict__ n_tiles_dev = nullptr /* optional: the tile count lives on the device (capped by n_tiles) */) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (n_tiles_dev) n_tiles = min(n_tiles, __ldg(n_tiles_dev));
  // grid-stride over the tiles: with a device-side count the launch is sized before the count is known
  for (int tsel = tile_per_warp ? (int)blockIdx.x * WARPS_PER_CTA + warp : (int)blockIdx.x; tsel < n_tiles;  // no block-wide barrier below
       tsel += (int)gridDim.x * (tile_per_warp ? WARPS_PER_CTA : 1)) {
  Tile tile = tiles[tile_list ? tile_list[tsel] : tsel];
  const RefView ref = refs[tile.ref_index];
  const unsigned FULL = 0xffffffffu;
  if (tile_per_warp && tile.n_loci == 1) {
    // a single-locus tile arrives with the read span of the tile it came from: narrow it to the reads that can overlap
    // the locus (start <= x, running-max end > x), two 32-ary searches by the whole warp
    const int64_t x = tile.locus0, key = ((int64_t)tile.contig_id << 40) | x;
    tile.read_hi = warp_partition_point(tile.read_lo, tile.read_hi, lane, [&](int64_t i) { return __ldg(R.start + i) > x; });
    tile.read_lo = warp_partition_point(tile.read_lo, tile.read_hi, lane, [&](int64_t i) { return __ldg(R.runmax_key + i) > key; });
  }
  for (int li = tile_per_warp ? 0 : warp; li < tile.n_loci; li += tile_per_warp ? 1 : WARPS_PER_CTA) {
    const int64_t x = tile.locus0 + li, slot = tile.slot0 + li;
    int depthN = 0, depthT = 0, refN = 0, refT = 0;
    bool any = false;
    // distinct variant alleles of the tumor pileup, one table entry per lane
    uint64_t tkey = 0;
    int tcnt = 0;
    int64_t tfirst = 0;
    int ntab = 0;
    for (int64_t base = tile.read_lo; base < tile.read_hi; base += 32) {
      const int64_t r = base + lane;
      bool ov = false;
      int smp = -1;
      Elem e;
      e.kind = GUAC_ALN_CLIPPED;
      if (r < tile.read_hi) {
        const ReadAtLocus at = read_at_locus(R, r, x, ref, e, [&](int s) { return s == P.normal_sample || s == P.tumor_sample; });
        ov = at.overlaps;
        smp = at.sample;
        if (at.error != DERR_NONE) report_error(err, at.error, tile.contig_id, x, r);
      }
      if (ov && smp >= P.n_samples) ov = false;
      any = any || __any_sync(FULL, ov);
      const bool mine = ov && (smp == P.normal_sample || smp == P.tumor_sample);
      const bool isN = mine && smp == P.normal_sample, isT = mine && smp == P.tumor_sample;
      depthN += __popc(__ballot_sync(FULL, isN));
      depthT += __popc(__ballot_sync(FULL, isT));
      refN += __popc(__ballot_sync(FULL, isN && e.kind == GUAC_ALN_MATCH));
      refT += __popc(__ballot_sync(FULL, isT && e.kind == GUAC_ALN_MATCH));
      const bool var = isT && e.kind != GUAC_ALN_MATCH && e.kind != GUAC_ALN_CLIPPED;  // allele.isVariant
      uint64_t key = var ? allele_key(R, r, e) : 0;
      unsigned vm = __ballot_sync(FULL, var);
      while (vm) {
        const int L = __ffs(vm) - 1;
        const uint64_t k = __shfl_sync(FULL, key, L);
        const unsigned same = __ballot_sync(FULL, var && key == k);
        const int c = __popc(same);
        const unsigned hit = __ballot_sync(FULL, lane < ntab && tkey == k);
        if (hit) {
          if (lane == __ffs(hit) - 1) tcnt += c;
        } else if (ntab < 32) {
          if (lane == ntab) {
            tkey = k;
            tcnt = c;
            tfirst = base + L;
          }
          ++ntab;
        } else if (lane == 0) {
          report_error(err, DERR_ALLELE_TABLE_FULL, tile.contig_id, x, base + L);
        }
        vm &= ~same;
      }
    }
    // most frequent variant allele; ties -> lowest first supporting read (documented canonical order)
    int bcnt = lane < ntab ? tcnt : -1;
    int64_t bfirst = lane < ntab ? tfirst : INT64_MAX;
    uint64_t bkey = tkey;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
      int oc = __shfl_xor_sync(FULL, bcnt, d);
      int64_t of = __shfl_xor_sync(FULL, bfirst, d);
      uint64_t ok = __shfl_xor_sync(FULL, bkey, d);
      if (oc > bcnt || (oc == bcnt && of < bfirst)) {
        bcnt = oc;
        bfirst = of;
        bkey = ok;
      }
    }
    // early outs (SomaticStandardCaller.scala:164-172)
    const bool candidate = depthT > 0 && depthN > 0 && depthT <= P.max_read_depth && depthN <= P.max_read_depth &&
                           refT != depthT && (depthN - refN) <= P.max_normal_alt;
    if (lane == 0) {
      visited[slot] = any ? 1 : 0;
      if (any && n_visited) atomicAdd(n_visited, 1ull);
    }
    if (candidate && ntab == 0) {
      if (lane == 0) report_error(err, DERR_EMPTY_MAXBY, tile.contig_id, x, -1);
    }
    const bool evaluate = candidate && ntab > 0;
    int normalAlt = 0;
    bool called = false;
    if (evaluate) {
      const double f = fmax(P.min_vaf, (double)bcnt / (double)depthT);  // :189
      double aTR = 0, aTA = 0, aNR = 0, aNA = 0;
      for (int64_t base = tile.read_lo; base < tile.read_hi; base += 32) {
        const int64_t r = base + lane;
        bool ov = false;
        int smp = -1;
        Elem e;
        int ec = DERR_NONE;
        if (r < tile.read_hi) {
          const ReadAtLocus at = read_at_locus(R, r, x, ref, e, [&](int s) { return s == P.normal_sample || s == P.tumor_sample; });
          ov = at.overlaps;
          smp = at.sample;
          ec = at.error;
        }
        const bool mine = ov && (smp == P.normal_sample || smp == P.tumor_sample);
        bool isAlt = false;
        if (mine) {
          if (ec == DERR_NONE) {
            const double p = c_success_probability[e.quality];  // probabilityCorrectIgnoringAlignment
            const bool isRef = e.kind == GUAC_ALN_MATCH;
            isAlt = e.kind != GUAC_ALN_MATCH && e.kind != GUAC_ALN_CLIPPED && allele_key(R, r, e) == bkey;
            const double pr = isRef ? p : 1 - p;  // Likelihood.scala:199-204
            const double pa = isAlt ? p : 1 - p;
            if (smp == P.tumor_sample) {
              aTR += log(pr * 1.0);
              aTA += log(pr * (1.0 - f) + pa * f);
            } else {
              aNR += log(pr * 1.0);
              aNA += log(pr * 0.5 + pa * 0.5);
            }
          }
        }
        normalAlt += __popc(__ballot_sync(FULL, mine && smp == P.normal_sample && isAlt));
      }
      aTR = warp_sum(aTR);
      aTA = warp_sum(aTA);
      aNR = warp_sum(aNR);
      aNA = warp_sum(aNA);
      double tR, tA, nR, nA;
      log_normalize2(aTR, aTA, tR, tA);
      log_normalize2(aNR, aNA, nR, nA);
      const double tumorAltLOD = tA - tR, normalRefLOD = nR - nA;
      // :231-233; MidDeletion has empty alt bases
      called = tumorAltLOD > P.tumor_odds && normalRefLOD > P.normal_odds && (bkey >> 60) != 4;
      if (called && lane == 0) {
        int idx = atomicAdd(n_calls, 1);
        if (idx < max_calls) {
          SSCallDev c;
          c.slot = slot;
          c.locus = x;
          c.read_lo = tile.read_lo;
          c.read_hi = tile.read_hi;
          c.key = bkey;
          c.rep_read = bfirst;
          c.contig_id = tile.contig_id;
          c.ref_index = tile.ref_index;
          c.tumor_ref_lp = tR;
          c.tumor_alt_lp = tA;
          c.normal_ref_lp = nR;
          c.normal_alt_lp = nA;
          c.ref_off = c.ref_len = c.alt_off = c.alt_len = 0;
          calls[idx] = c;
        }
      }
    }
    if (summary != nullptr && lane == 0) {
      const uint16_t fl = (evaluate ? 1 : 0) | (called ? 2 : 0);
      guac_locus_summary sn, st;
      sn.depth = (uint16_t)min(depthN, 65535);
      sn.reference_depth = (uint16_t)min(refN, 65535);
      sn.top_alt_count = (uint16_t)min(evaluate ? normalAlt : 0, 65535);
      sn.flags = fl;
      st.depth = (uint16_t)min(depthT, 65535);
      st.reference_depth = (uint16_t)min(refT, 65535);
      st.top_alt_count = (uint16_t)min(evaluate ? bcnt : 0, 65535);
      st.flags = fl;
      summary[slot * 2 + 0] = sn;
      summary[slot * 2 + 1] = st;
    }
  }
  }
}

// MappedRead.countOfMismatches (MappedRead.scala:53-64): Mismatch elements over the whole read.
__device__ inline int count_of_mismatches(const ReadsView& R, int64_t r, const RefView& ref, DeviceError* err, int contig) {
  const int64_t s = __ldg(R.start + r), en = __ldg(R.end + r);
  int n = 0;
  const uint32_t c0 = __ldg(R.cigar_off + r), c1 = __ldg(R.cigar_off + r + 1);
  if (c1 - c0 == 1) {  // a single aligned operator: every element is a Match or a Mismatch
    const uint32_t w = __ldg(R.cigar + c0);
    const int op = w & 15;
    if (op == OPC_M || op == OPC_EQ || op == OPC_X) {
      const uint64_t s0 = __ldg(R.seq_off + r);
      for (int64_t x = s; x < en; ++x) {
        const uint8_t rb = ref.at(x);
        if (rb == 0) {
          report_error(err, DERR_REF_MISSING, contig, x, r);
          return n;
        }
        n += read_base(R, s0 + (uint64_t)(x - s)) != rb;
      }
      return n;
    }
  }
  // Several operators out of {M,=,X,I,D,S,H}: walk them.  Only bases of aligned operators can be Mismatch elements, and
  // the last base of an operator that anchors a following insertion (M/= only) or deletion is an Insertion / Deletion
  // element whatever its base (PE.scala:59-117) — the same rule the tile kernel applies.
  {
    bool plain = !(s == 0 && (__ldg(R.cigar + c0) & 15) == OPC_I);  // PE.scala:88-89: leave to the exact cursor
    for (uint32_t ci = c0; plain && ci < c1; ++ci) {
      const uint32_t w = __ldg(R.cigar + ci);
      const int op = w & 15;
      plain = (w >> 4) != 0 && (op == OPC_M || op == OPC_EQ || op == OPC_X || op == OPC_I || op == OPC_D || op == OPC_S || op == OPC_H);
    }
    if (plain) {
      const uint64_t s0 = __ldg(R.seq_off + r);
      int64_t cl = s;
      uint64_t rp = 0;
      for (uint32_t ci = c0; ci < c1; ++ci) {
        const uint32_t w = __ldg(R.cigar + ci);
        const int op = w & 15, len = (int)(w >> 4);
        if (op == OPC_M || op == OPC_EQ || op == OPC_X) {
          const int nop = ci + 1 < c1 ? (int)(__ldg(R.cigar + ci + 1) & 15) : -1;
          const int plain_len = len - ((nop == OPC_D || (nop == OPC_I && op != OPC_X)) ? 1 : 0);
          for (int k = 0; k < plain_len; ++k) {
            const uint8_t rb = ref.at(cl + k);
            if (rb == 0) {
              report_error(err, DERR_REF_MISSING, contig, cl + k, r);
              return n;
            }
            n += read_base(R, s0 + rp + (uint64_t)k) != rb;
          }
          cl += len;
          rp += (uint64_t)len;
        } else if (op == OPC_D) {
          cl += len;
        } else if (op == OPC_I || op == OPC_S) {
          rp += (uint64_t)len;
        }
      }
      return n;
    }
  }
  for (int64_t x = s; x < en; ++x) {
    Elem e;
    int ec = element_at(R, r, x, ref, e);
    if (ec != DERR_NONE) {
      report_error(err, ec, contig, x, r);
      return n;
    }
    n += e.kind == GUAC_ALN_MISMATCH;
  }
  return n;
}

// Position of every call in (range order, locus) order = the number of calls with a smaller slot (slots are distinct: one
// call per locus).  Quadratic, for the few thousand calls of a partition; the host sorts instead beyond SS_RANK_MAX.
constexpr int SS_RANK_MAX = 32768;
constexpr int SS_RANK_SPLIT = 8;  // blockIdx.y: the slots are compared in this many slices, partial counts added up
__global__ void __launch_bounds__(256) ss_call_rank_kernel(const SSCallDev* __restrict__ calls, int n, int* __restrict__ rank /* zeroed */) {
  __shared__ int64_t s_slot[256];
  const int i = blockIdx.x * 256 + threadIdx.x;
  const int64_t mine = i < n ? calls[i].slot : 0;
  const int per = ((n + SS_RANK_SPLIT - 1) / SS_RANK_SPLIT + 255) & ~255;  // slice length, whole tiles of 256
  const int j0 = blockIdx.y * per, j1 = min(n, j0 + per);
  int below = 0;
  for (int base = j0; base < j1; base += 256) {
    __syncthreads();
    s_slot[threadIdx.x] = base + threadIdx.x < j1 ? calls[base + threadIdx.x].slot : INT64_MAX;
    __syncthreads();
#pragma unroll 8
    for (int k = 0; k < 256; ++k) below += s_slot[k] < mine;
  }
  if (i < n && below) atomicAdd(rank + i, below);
}

// The same order for any number of records, linear in the number of loci: a bitmap over the call's slots, the records'
// slots set in it, an exclusive scan of the words' population counts; the position of a record is the count of set bits
// below its own.  `records` + i * stride + slot_offset holds the (distinct) int64 slot of record i.
__global__ void slot_bitmap_set_kernel(const unsigned char* __restrict__ records, size_t stride, size_t slot_offset, int n,
                                       unsigned int* __restrict__ bitmap) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int64_t slot = *reinterpret_cast<const int64_t*>(records + (size_t)i * stride + slot_offset);
  atomicOr(bitmap + (slot >> 5), 1u << (slot & 31));
}
__global__ void slot_bitmap_popc_kernel(const unsigned int* __restrict__ bitmap, int64_t n_words, uint8_t* __restrict__ counts) {
  const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (w < n_words) counts[w] = (uint8_t)__popc(bitmap[w]);
}
__global__ void slot_bitmap_rank_kernel(const unsigned char* __restrict__ records, size_t stride, size_t slot_offset, int n,
                                        const unsigned int* __restrict__ bitmap, const int64_t* __restrict__ word_prefix,
                                        int* __restrict__ rank) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int64_t slot = *reinterpret_cast<const int64_t*>(records + (size_t)i * stride + slot_offset);
  rank[i] = (int)word_prefix[slot >> 5] + __popc(bitmap[slot >> 5] & ((1u << (slot & 31)) - 1u));
}

constexpr int MM_BINS = 1024;

// lower/upper middle of a histogram with n entries -> (a, b): values at ranks (n-1)/2 and n/2
__device__ inline void hist_middle(const unsigned* h, int bins, int n, int& a, int& b) {
  int ra = (n - 1) / 2, rb = n / 2;
  int acc = 0;
  a = b = 0;
  bool gotA = false;
  for (int v = 0; v < bins; ++v) {
    acc += (int)h[v];
    if (!gotA && acc > ra) {
      a = v;
      gotA = true;
    }
    if (acc > rb) {
      b = v;
      return;
    }
  }
}

// AlleleEvidence.apply (variants/AlleleEvidence.scala:42-85) for both evidences of one call:
// tumor with the called allele, normal with the reference allele.  One CTA per call.
// Also writes the allele strings (variants/Allele.scala) into the arena.
__global__ void __launch_bounds__(128)
ss_evidence_kernel(ReadsView R, const RefView* __restrict__ refs, SSParamsDev P, SSCallDev* __restrict__ calls,
                   char* __restrict__ arena, unsigned int* __restrict__ arena_used, unsigned int arena_cap, DeviceError* err,
                   const uint4* __restrict__ hdr /* the three below: null unless the read pass of this call left its events */,
                   const uint2* __restrict__ ev_by_read, const uint8_t* __restrict__ ev_count, int ev_cap,
                   const int* __restrict__ rank /* position of each call in the output */, guac_ss_call* __restrict__ out) {
  // [0] tumor with the called allele, [1] normal with the reference allele
  __shared__ unsigned h_mapq[2][256], h_bq[2][256], h_mm[2][MM_BINS];
  __shared__ unsigned s_depth[2], s_pos[2], s_adepth[2], s_apos[2];
  __shared__ int s_list[1024];  // reads of either sample that overlap the locus (read offset << 1 | which)
  __shared__ int s_mm_read[512], s_mm_cnt[512];  // reads whose mismatches are to be counted, and the counts
  __shared__ unsigned char s_mm_slow[512];       // 1: not countable from the read pass's events
  __shared__ int s_n, s_nmm;
  SSCallDev& c = calls[blockIdx.x];
  const RefView ref = refs[c.ref_index];
  const int64_t x = c.locus;
  for (int i = threadIdx.x; i < 512; i += blockDim.x) (&h_mapq[0][0])[i] = (&h_bq[0][0])[i] = 0;
  for (int i = threadIdx.x; i < 2 * MM_BINS; i += blockDim.x) (&h_mm[0][0])[i] = 0;
  if (threadIdx.x < 2) s_depth[threadIdx.x] = s_pos[threadIdx.x] = s_adepth[threadIdx.x] = s_apos[threadIdx.x] = 0;
  if (threadIdx.x == 0) s_n = s_nmm = 0;
  __syncthreads();
  // reads are sorted by start: nothing at or after the first read that starts beyond the locus can overlap it
  int64_t hi = c.read_hi;
  {
    int64_t a = c.read_lo, b = c.read_hi;
    while (a < b) {
      const int64_t m = (a + b) >> 1;
      if (__ldg(R.start + m) <= x) a = m + 1; else b = m;
    }
    hi = a;
  }
  auto visit = [&](int64_t r, int which) {
    Elem e;
    if (element_at(R, r, x, ref, e) != DERR_NONE) return;
    const bool pos = __ldg(R.flags + r) & GUAC_READ_POSITIVE_STRAND;
    atomicAdd(&s_depth[which], 1u);
    if (pos) atomicAdd(&s_pos[which], 1u);
    bool isAllele;
    if (which == 0)
      isAllele = e.kind != GUAC_ALN_MATCH && e.kind != GUAC_ALN_CLIPPED && allele_key(R, r, e) == c.key;
    else
      isAllele = e.kind == GUAC_ALN_MATCH;
    if (isAllele) {
      atomicAdd(&s_adepth[which], 1u);
      if (pos) atomicAdd(&s_apos[which], 1u);
      atomicAdd(&h_mapq[which][__ldg(R.mapq + r)], 1u);
      atomicAdd(&h_bq[which][e.quality], 1u);
      // countOfMismatches walks the whole read: queued, and shared out over all threads below
      const int j = atomicAdd(&s_nmm, 1);
      if (j < 512) {
        s_mm_read[j] = (int)((r - c.read_lo) << 1) | which;
        s_mm_cnt[j] = 0;
      } else {  // deeper than the queue: in place
        int mm = count_of_mismatches(R, r, ref, err, c.contig_id);
        if (mm >= MM_BINS) {
          report_error(err, DERR_MISMATCH_HIST, c.contig_id, x, r);
          mm = MM_BINS - 1;
        }
        atomicAdd(&h_mm[which][mm], 1u);
      }
    }
  };
  // pass 1: which reads overlap the locus (cheap, coalesced); pass 2: one thread per overlapping read
  for (int64_t r = c.read_lo + threadIdx.x; r < hi; r += blockDim.x) {
    if (x >= __ldg(R.end + r)) continue;
    const int smp = __ldg(R.sample_id + r);
    if (smp != P.tumor_sample && smp != P.normal_sample) continue;
    const int which = smp == P.tumor_sample ? 0 : 1;
    const int i = atomicAdd(&s_n, 1);
    if (i < 1024) s_list[i] = (int)((r - c.read_lo) << 1) | which; else visit(r, which);  // deeper than the list: in place
  }
  __syncthreads();
  const int n_list = min(s_n, 1024);
  for (int i = threadIdx.x; i < n_list; i += blockDim.x) visit(c.read_lo + (s_list[i] >> 1), s_list[i] & 1);
  __syncthreads();
  // MappedRead.countOfMismatches (MappedRead.scala:53-64) of the queued reads: 16 loci of a single-operator read per
  // work item, so a call's ~45 reads x 100 bases keep all 128 threads busy instead of 45 threads for 100 iterations
  const int n_mm = min(s_nmm, 512);
  if (ev_count != nullptr) {
    // The read pass of this call (ss_read_events_kernel) already compared every such read with the reference: the
    // Mismatch records of its slot are the count, provided the slot is complete (no overflow), the read lies inside the
    // reference window and none of its loci has a reference base outside A C G T (those the walk below must see).
    for (int j = threadIdx.x; j < n_mm; j += blockDim.x) {
      const int64_t r = c.read_lo + (s_mm_read[j] >> 1);
      const int cnt = __ldg(ev_count + r);
      const int64_t s = __ldg(R.start + r), en = __ldg(R.end + r);
      bool ok = ((__ldg(hdr + r).y >> 24) & 4 /* RF_MEDIUM */) && cnt <= ev_cap /* a complete slot */ && s >= ref.offset && en <= ref.offset + ref.len && en > s;
      if (ok && ref.bad != nullptr)
        for (int64_t g = (s - ref.offset) >> 4; ok && g <= (en - 1 - ref.offset) >> 4; ++g) ok = !((__ldg(ref.bad + (g >> 5)) >> (g & 31)) & 1u);
      int n = -1;
      if (ok) {
        n = 0;
        for (int k = 0; k < cnt; ++k) n += ((__ldg(&ev_by_read[r * ev_cap + k].x) >> 16) & 7) == GUAC_ALN_MISMATCH;
      }
      s_mm_cnt[j] = max(n, 0);
      s_mm_slow[j] = n < 0;
    }
    __syncthreads();
  }
  for (int item = threadIdx.x; item < n_mm * 16; item += blockDim.x) {
    const int j = item >> 4, chunk = item & 15;
    if (ev_count != nullptr && !s_mm_slow[j]) continue;  // counted from the events
    const int64_t r = c.read_lo + (s_mm_read[j] >> 1);
    const uint32_t c0 = __ldg(R.cigar_off + r), c1 = __ldg(R.cigar_off + r + 1);
    const int64_t s = __ldg(R.start + r), en = __ldg(R.end + r);
    const int op = c1 - c0 == 1 ? (int)(__ldg(R.cigar + c0) & 15) : -1;
    if ((op == OPC_M || op == OPC_EQ || op == OPC_X) && en - s <= 256) {
      const uint64_t s0 = __ldg(R.seq_off + r);
      int n = 0;
      for (int64_t l = s + 16 * chunk; l < min(en, s + 16 * chunk + 16); ++l) {
        const uint8_t rb = ref.at(l);
        if (rb == 0) {
          report_error(err, DERR_REF_MISSING, c.contig_id, l, r);
          break;
        }
        n += read_base(R, s0 + (uint64_t)(l - s)) != rb;
      }
      if (n) atomicAdd(&s_mm_cnt[j], n);
    } else if (chunk == 0) {  // several operators, or a very long read: one thread takes the whole read
      atomicAdd(&s_mm_cnt[j], count_of_mismatches(R, r, ref, err, c.contig_id));
    }
  }
  __syncthreads();
  for (int j = threadIdx.x; j < n_mm; j += blockDim.x) {
    int mm = s_mm_cnt[j];
    if (mm >= MM_BINS) {
      report_error(err, DERR_MISMATCH_HIST, c.contig_id, x, c.read_lo + (s_mm_read[j] >> 1));
      mm = MM_BINS - 1;
    }
    atomicAdd(&h_mm[s_mm_read[j] & 1][mm], 1u);
  }
  __syncthreads();
  if (threadIdx.x < 2) {
    const int which = threadIdx.x;
    guac_allele_evidence ev;
    const double lp = which == 0 ? c.tumor_alt_lp : c.normal_ref_lp;
    ev.probability = exp(-lp);  // SomaticStandardCaller.scala:237-238 (sic)
    ev.read_depth = (int)s_depth[which];
    ev.allele_read_depth = (int)s_adepth[which];
    ev.forward_depth = (int)s_pos[which];
    ev.allele_forward_depth = (int)s_apos[which];
    const int n = (int)s_adepth[which];
    if (n == 0) {
      const double nan = __longlong_as_double(0x7ff8000000000000ll);
      ev.mean_mapping_quality = ev.median_mapping_quality = ev.mean_base_quality = ev.median_base_quality =
          ev.median_mismatches_per_read = nan;
    } else {
      double sm = 0, sq = 0;
      for (int v = 0; v < 256; ++v) {
        sm += (double)v * h_mapq[which][v];
        sq += (double)v * h_bq[which][v];
      }
      ev.mean_mapping_quality = sm / n;
      ev.mean_base_quality = sq / n;
      int a, b;
      hist_middle(h_mapq[which], 256, n, a, b);
      ev.median_mapping_quality = (n & 1) ? (double)a : ((double)a + (double)b) / 2;
      hist_middle(h_bq[which], 256, n, a, b);
      ev.median_base_quality = (n & 1) ? (double)a : ((double)a + (double)b) / 2;
      hist_middle(h_mm[which], MM_BINS, n, a, b);
      ev.median_mismatches_per_read = (n & 1) ? (double)a : (double)((a + b) / 2);  // DenseVector[Int] median
    }
    if (which == 0) c.tumor_ev = ev; else c.normal_ev = ev;
  }
  // allele strings: lengths, a slice of the arena (claimed with one atomic; the host grows the arena and re-runs
  // this kernel in the rare case the slices do not fit), then the bases
  if (threadIdx.x == 0) {
    const int kind = (int)(c.key >> 60);
    Elem e;
    bool have_elem = false;
    uint32_t ref_len = 1, alt_len = 1;
    if (kind == 3) {
      ref_len = (uint32_t)(c.key & 0xffffffffu) + 1;
    } else if (kind != 1) {
      have_elem = element_at(R, c.rep_read, x, ref, e) == DERR_NONE;
      if (have_elem) {
        const int32_t read_len = (int32_t)(__ldg(R.seq_off + c.rep_read + 1) - __ldg(R.seq_off + c.rep_read));
        alt_len = (uint32_t)(min(e.read_position + e.indel_length + 1, read_len) - e.read_position);
      }
    }
    const unsigned int off = atomicAdd(arena_used, ref_len + alt_len);
    c.ref_len = ref_len;
    c.alt_len = alt_len;
    c.ref_off = off;
    c.alt_off = off + ref_len;
    if (off + ref_len + alt_len <= arena_cap) {
      char* rp = arena + c.ref_off;
      char* ap = arena + c.alt_off;
      const uint8_t rb = ref.at(x);
      if (kind == 1) {  // Mismatch
        rp[0] = (char)rb;
        ap[0] = (char)(c.key & 0xff);
      } else if (kind == 3) {  // Deletion: ref = ref[x .. x+len], alt = ref[x]
        for (uint32_t k = 0; k < ref_len; ++k) rp[k] = (char)ref.at(x + k);
        ap[0] = (char)rb;
      } else if (kind == 2 || kind == 5) {  // Insertion: ref = seq[0], alt = seq
        if (have_elem) {
          const uint64_t s0 = __ldg(R.seq_off + c.rep_read);
          for (uint32_t k = 0; k < alt_len; ++k) ap[k] = (char)read_base(R, s0 + e.read_position + k);
          rp[0] = ap[0];
        }
      }
    }
  }
  __syncthreads();  // both evidences and the string offsets are in the record
  if (threadIdx.x == 0) {  // the finished record, at its place in (range order, locus) order
    guac_ss_call g = {};
    g.contig_id = c.contig_id;
    g.start = c.locus;
    g.ref_off = c.ref_off;
    g.ref_len = c.ref_len;
    g.alt_off = c.alt_off;
    g.alt_len = c.alt_len;
    g.somatic_log_odds = c.tumor_alt_lp - c.tumor_ref_lp;
    g.tumor_ref_log_prob = c.tumor_ref_lp;
    g.tumor_alt_log_prob = c.tumor_alt_lp;
    g.normal_ref_log_prob = c.normal_ref_lp;
    g.normal_alt_log_prob = c.normal_alt_lp;
    g.tumor_variant_evidence = c.tumor_ev;
    g.normal_reference_evidence = c.normal_ev;
    out[rank[blockIdx.x]] = g;
  }
}

}  // namespace guac
