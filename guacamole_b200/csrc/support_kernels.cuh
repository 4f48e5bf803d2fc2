// support_kernels.cuh — the two small consumers of the pileup operator besides the callers:
//   allele_counts_kernel   VariantSupport.Caller.pileupsToAlleleCounts (commands/VariantSupport.scala:100-112): per
//                          visited locus and sample, the distinct alleles of the pileup's elements with their counts
//                          (pileup.elements.groupBy(_.allele)); one warp per locus, the table one entry per lane.
//   vaf_loci_kernel        VariantLocus.apply + the filter of VAFHistogram.variantLociFromReads and the bins of
//                          generateVAFHistograms (commands/VAFHistogram.scala:37-58,194-245) over the counting
//                          closure's output: Float arithmetic as the Scala does it.
// Reference file names are relative to src/main/scala/org/hammerlab/guacamole/.
#pragma once
#include "kernels.cuh"

namespace guac {

struct AlleleCountDev {
  int64_t slot, locus, rep_read;  // rep_read: first element (lowest read index) with this allele
  uint64_t key;
  int32_t contig_id, ref_index, sample, count;
  int32_t kind, indel_len, rep_rp;
  uint8_t base, ref_base, _pad[2];
};

// Allele identity of ANY element (variants/Allele.scala:7-17 over pileup/Alignment.scala:14-75): the variant kinds as
// allele_key(); Match is (base, base), Clipped is ("", "").
__device__ inline uint64_t element_allele_key(const ReadsView& R, int64_t r, const Elem& e) {
  if (e.kind == GUAC_ALN_MATCH) return (6ull << 60) | e.base;
  if (e.kind == GUAC_ALN_CLIPPED) return 7ull << 60;
  return allele_key(R, r, e);
}

__global__ void __launch_bounds__(CTA_THREADS)
allele_counts_kernel(ReadsView R, const RefView* __restrict__ refs, const Tile* __restrict__ tiles, int n_samples,
                     uint8_t* __restrict__ visited, AlleleCountDev* __restrict__ out, int* __restrict__ n_out, int max_out,
                     DeviceError* err) {
  const Tile tile = tiles[blockIdx.x];
  const RefView ref = refs[tile.ref_index];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned FULL = 0xffffffffu;
  for (int li = warp; li < tile.n_loci; li += WARPS_PER_CTA) {
    const int64_t x = tile.locus0 + li, slot = tile.slot0 + li;
    // table: one (sample, allele) per lane
    uint64_t tkey = 0;
    int tsmp = 0, tcnt = 0, tkind = 0, tindel = 0, trp = 0, ntab = 0;
    int64_t tfirst = 0;
    uint8_t tbase = 0, trefb = 0;
    bool any = false;
    for (int64_t base = tile.read_lo; base < tile.read_hi; base += 32) {
      const int64_t r = base + lane;
      bool mine = r < tile.read_hi && __ldg(R.start + r) <= x && x < __ldg(R.end + r);
      const int smp = mine ? (int)__ldg(R.sample_id + r) : 0;
      if (mine && smp >= n_samples) mine = false;
      Elem e;
      e.kind = GUAC_ALN_CLIPPED;
      uint64_t key = 0;
      if (mine) {
        const int ec = element_at(R, r, x, ref, e);
        if (ec != DERR_NONE) {
          report_error(err, ec, tile.contig_id, x, r);
          mine = false;
        } else {
          key = element_allele_key(R, r, e);
        }
      }
      any = any || __any_sync(FULL, mine);
      unsigned vm = __ballot_sync(FULL, mine);
      while (vm) {  // groupBy(_.allele) within the sample
        const int L = __ffs(vm) - 1;
        const uint64_t k = __shfl_sync(FULL, key, L);
        const int s = __shfl_sync(FULL, smp, L);
        const unsigned same = __ballot_sync(FULL, mine && key == k && smp == s);
        const int c = __popc(same);
        const unsigned hit = __ballot_sync(FULL, lane < ntab && tkey == k && tsmp == s);
        const int kind = __shfl_sync(FULL, (int)e.kind, L), indel = __shfl_sync(FULL, e.indel_length, L);
        const int rp = __shfl_sync(FULL, e.read_position, L), b = __shfl_sync(FULL, (int)e.base, L), rb = __shfl_sync(FULL, (int)e.ref_base, L);
        if (hit) {
          if (lane == __ffs(hit) - 1) tcnt += c;
        } else if (ntab < 32) {
          if (lane == ntab) {
            tkey = k;
            tsmp = s;
            tcnt = c;
            tfirst = base + L;
            tkind = kind;
            tindel = indel;
            trp = rp;
            tbase = (uint8_t)b;
            trefb = (uint8_t)rb;
          }
          ++ntab;
        } else if (lane == 0) {
          report_error(err, DERR_ALLELE_TABLE_FULL, tile.contig_id, x, base + L);
        }
        vm &= ~same;
      }
    }
    if (lane == 0) visited[slot] = any ? 1 : 0;
    if (ntab == 0) continue;
    int o0 = 0;
    if (lane == 0) o0 = atomicAdd(n_out, ntab);
    o0 = __shfl_sync(FULL, o0, 0);
    if (lane < ntab && o0 + lane < max_out) {
      AlleleCountDev a;
      a.slot = slot;
      a.locus = x;
      a.rep_read = tfirst;
      a.key = tkey;
      a.contig_id = tile.contig_id;
      a.ref_index = tile.ref_index;
      a.sample = tsmp;
      a.count = tcnt;
      a.kind = tkind;
      a.indel_len = tindel;
      a.rep_rp = trp;
      a.base = tbase;
      a.ref_base = trefb;
      a._pad[0] = a._pad[1] = 0;
      out[o0 + lane] = a;
    }
  }
}

struct VafParams {
  int n_samples, bins, min_read_depth, min_vaf_percent;
};

// flags[slot * S + s] = 1 where sample s has a VariantLocus at the slot's locus that passes the filters; the histogram and
// the per-sample totals are added on the way.  (pileup.depth - pileup.referenceDepth).toFloat / pileup.depth is a Float
// division; `>= minVariantAlleleFrequency / 100.0` compares it as a Double; (vaf * 100).toInt is a Float product.
__global__ void vaf_loci_kernel(const guac_locus_counts* __restrict__ counts, const uint8_t* __restrict__ visited, int64_t n_slots,
                                VafParams P, uint8_t* __restrict__ flags, float* __restrict__ vaf_out,
                                unsigned long long* __restrict__ hist /* [S][101] */, unsigned long long* __restrict__ totals /* [S] */) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_slots * P.n_samples) return;
  const int64_t slot = i / P.n_samples;
  const int s = (int)(i - slot * P.n_samples);
  uint8_t f = 0;
  if (visited[slot]) {
    const guac_locus_counts c = counts[i];
    const int depth = (int)c.depth, ref_depth = (int)c.reference_depth;
    if (depth >= P.min_read_depth && ref_depth != depth) {
      const float vaf = __fdiv_rn((float)(depth - ref_depth), (float)depth);
      if ((double)vaf >= (double)P.min_vaf_percent / 100.0) {
        const int percent = (int)__fmul_rn(vaf, 100.0f);
        const int bin = percent - (percent % (100 / P.bins));
        f = 1;
        vaf_out[i] = vaf;
        atomicAdd(hist + (size_t)s * 101 + bin, 1ull);
        atomicAdd(totals + s, 1ull);
      }
    }
  }
  flags[i] = f;
}

__global__ void vaf_records_kernel(const uint8_t* __restrict__ flags, const int64_t* __restrict__ offset, const float* __restrict__ vaf,
                                   const Tile* __restrict__ tiles, int64_t n_tiles, int tile_loci, int n_samples,
                                   guac_variant_locus* __restrict__ out) {
  const int64_t t = blockIdx.x;
  if (t >= n_tiles) return;
  const Tile tile = tiles[t];
  for (int k = threadIdx.x; k < tile.n_loci * n_samples; k += blockDim.x) {
    const int li = k / n_samples, s = k - li * n_samples;
    const int64_t i = (tile.slot0 + li) * n_samples + s;
    if (!flags[i]) continue;
    guac_variant_locus v;
    v.sample_id = s;
    v.contig_id = tile.contig_id;
    v.locus = tile.locus0 + li;
    v.variant_allele_frequency = vaf[i];
    v._pad = 0;
    out[offset[i]] = v;
  }
}

}  // namespace guac
