// germline_kernels.cuh — the germline toy caller's per-locus kernel and the generic AlleleEvidence kernel.
//
//   germline_locus_kernel   one warp per candidate locus (depth - referenceDepth >= minAltReads, found by the
//                           streaming tile kernel in germline mode): Pileup.distinctAlleles
//                           (pileup/Pileup.scala:32), Likelihood.probabilitiesOfAllPossibleGenotypesFromPileup
//                           (likelihood/Likelihood.scala:31-63,99-151), callPileupVariant
//                           (commands/GermlineAssemblyCaller.scala:233-261) and, on request,
//                           AssemblyUtils.isActiveRegion (assembly/AssemblyUtils.scala:28-39).
//   allele_evidence_kernel  AlleleEvidence.apply (variants/AlleleEvidence.scala:42-85) for (locus, sample, allele) jobs.
// Reference paths are relative to src/main/scala/org/hammerlab/guacamole/.
#pragma once
#include "joint_kernels.cuh"

namespace guac {

constexpr int GL_MAX_ALLELES = 8;                                  // distinct alleles with standard alt bases per locus
constexpr int GL_MAX_GENOTYPES = GL_MAX_ALLELES * (GL_MAX_ALLELES + 1) / 2;
constexpr int GL_ELEM_CACHE = 1024;                                // elements staged per warp

struct GLParamsDev {
  int sample, min_alt_reads, respect_active_regions;
  float min_area_vaf;
};

struct GLAlleleDev {  // enough to rebuild the allele's strings
  uint64_t key;
  int64_t rep_read;
  int32_t rep_rp, indel_len;
  uint8_t kind, base, _p0, _p1;
  int32_t _pad;
};

struct GLCallDev {
  int64_t slot, locus, read_lo, read_hi;
  int32_t contig_id, ref_index, n_genotypes, n_called;
  double log_probability;
  GLAlleleDev called[2];
};

// allele identity including the non-variant ones: Match -> 6 << 60, Clipped -> 7 << 60
__device__ inline uint64_t any_allele_key(const ReadsView& R, int64_t r, const Elem& e) {
  if (e.kind == GUAC_ALN_MATCH) return 6ull << 60;
  if (e.kind == GUAC_ALN_CLIPPED) return 7ull << 60;
  return allele_key(R, r, e);
}

struct AlleleStrings {  // character access to Allele(refBases, altBases) of variants/Allele.scala:7
  const ReadsView* R;
  const RefView* ref;
  int64_t x;
  GLAlleleDev a;
  __device__ int ref_len() const {
    return a.kind == GUAC_ALN_CLIPPED ? 0 : a.kind == GUAC_ALN_DELETION ? a.indel_len + 1 : 1;
  }
  __device__ int alt_len() const {
    return a.kind == GUAC_ALN_CLIPPED || a.kind == GUAC_ALN_MID_DELETION ? 0 : a.kind == GUAC_ALN_INSERTION ? a.indel_len + 1 : 1;
  }
  __device__ uint8_t ins_base(int k) const { return read_base(*R, __ldg(R->seq_off + a.rep_read) + a.rep_rp + k); }
  __device__ uint8_t ref_char(int k) const {
    if (a.kind == GUAC_ALN_INSERTION) return ins_base(0);  // Insertion.referenceBases = sequencedBases.head (Alignment.scala:28)
    return ref->at(x + k);
  }
  __device__ uint8_t alt_char(int k) const {
    switch (a.kind) {
      case GUAC_ALN_MATCH: return ref->at(x);
      case GUAC_ALN_MISMATCH: return a.base;
      case GUAC_ALN_INSERTION: return ins_base(k);
      default: return ref->at(x);  // Deletion.sequencedBases = referenceBases.head (Alignment.scala:62)
    }
  }
  __device__ bool alt_all_standard() const {
    for (int k = 0; k < alt_len(); ++k) {
      const uint8_t b = alt_char(k);
      if (!(b == 'A' || b == 'C' || b == 'G' || b == 'T')) return false;
    }
    return true;
  }
};

// Allele.compare: BasesOrdering on refBases, then on altBases (variants/Allele.scala:12-17, util/Bases.scala:27-31)
__device__ inline int compare_alleles(const AlleleStrings& p, const AlleleStrings& q) {
  const int pr = p.ref_len(), qr = q.ref_len();
  for (int k = 0; k < min(pr, qr); ++k) {
    const int d = (int)p.ref_char(k) - (int)q.ref_char(k);
    if (d) return d;
  }
  if (pr != qr) return pr - qr;
  const int pa = p.alt_len(), qa = q.alt_len();
  for (int k = 0; k < min(pa, qa); ++k) {
    const int d = (int)p.alt_char(k) - (int)q.alt_char(k);
    if (d) return d;
  }
  return pa - qa;
}

__global__ void __launch_bounds__(CTA_THREADS)
germline_locus_kernel(ReadsView R, const RefView* __restrict__ refs, const Tile* __restrict__ heavy, int n_heavy, GLParamsDev P,
                      GLCallDev* __restrict__ calls, int* __restrict__ n_calls, int max_calls, long long* __restrict__ active,
                      int* __restrict__ n_active, int max_active, DeviceError* err) {
  __shared__ uint16_t s_elem[WARPS_PER_CTA][GL_ELEM_CACHE];  // allele index (high byte, 0xFF = not a listed allele) | quality
  __shared__ double s_ll[WARPS_PER_CTA][GL_MAX_GENOTYPES];
  __shared__ GLAlleleDev s_alleles[WARPS_PER_CTA][32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tsel = (int)blockIdx.x * WARPS_PER_CTA + warp;
  if (tsel >= n_heavy) return;
  const unsigned FULL = 0xffffffffu;
  Tile tile = heavy[tsel];
  const RefView ref = refs[tile.ref_index];
  const int64_t x = tile.locus0;
  {
    // the single-locus tile arrives with the read span of the tile it came from: narrow it to the reads that can overlap
    // the locus (start <= x, running-max end > x); every pass below, and the evidence kernel after it, walks this span
    const int64_t key = ((int64_t)tile.contig_id << 40) | x;
    tile.read_hi = warp_partition_point(tile.read_lo, tile.read_hi, lane, [&](int64_t i) { return __ldg(R.start + i) > x; });
    tile.read_lo = warp_partition_point(tile.read_lo, tile.read_hi, lane, [&](int64_t i) { return __ldg(R.runmax_key + i) > key; });
  }

  // pass 1: distinct alleles of the pileup (one table entry per lane), depth, active-region votes
  uint64_t tkey = 0;
  int64_t tfirst = 0;
  int trp = 0, tlen = 0, ntab = 0, depth = 0, votes = 0;
  uint8_t tkind = 0, tbase = 0;
  for (int64_t base = tile.read_lo; base < tile.read_hi; base += 32) {
    const int64_t r = base + lane;
    bool mine = r < tile.read_hi && __ldg(R.start + r) <= x && x < __ldg(R.end + r) && (int)__ldg(R.sample_id + r) == P.sample;
    Elem e;
    e.kind = GUAC_ALN_CLIPPED;
    if (mine) {
      const int ec = element_at(R, r, x, ref, e);
      if (ec != DERR_NONE) {
        report_error(err, ec, tile.contig_id, x, r);
        mine = false;
      }
    }
    depth += __popc(__ballot_sync(FULL, mine));
    if (P.respect_active_regions) {  // reads with > 1 mismatch or > 1 cigar element (AssemblyUtils.scala:33-36)
      bool vote = false;
      if (mine) vote = (__ldg(R.cigar_off + r + 1) - __ldg(R.cigar_off + r)) > 1 || count_of_mismatches(R, r, ref, err, tile.contig_id) > 1;
      votes += __popc(__ballot_sync(FULL, vote));
    }
    const uint64_t key = mine ? any_allele_key(R, r, e) : 0;
    unsigned vm = __ballot_sync(FULL, mine);
    while (vm) {
      const int L = __ffs(vm) - 1;
      const uint64_t k = __shfl_sync(FULL, key, L);
      const unsigned same = __ballot_sync(FULL, mine && key == k);
      const unsigned hit = __ballot_sync(FULL, lane < ntab && tkey == k);
      if (!hit) {
        if (ntab < 32) {
          const int kind = __shfl_sync(FULL, (int)e.kind, L), bs = __shfl_sync(FULL, (int)e.base, L);
          const int rp = __shfl_sync(FULL, e.read_position, L), il = __shfl_sync(FULL, e.indel_length, L);
          if (lane == ntab) {
            tkey = k;
            tfirst = base + L;
            trp = rp;
            tlen = il;
            tkind = (uint8_t)kind;
            tbase = (uint8_t)bs;
          }
          ++ntab;
        } else if (lane == 0) {
          report_error(err, DERR_ALLELE_TABLE_FULL, tile.contig_id, x, base + L);
        }
      }
      vm &= ~same;
    }
  }
  if (P.respect_active_regions && (float)votes / (float)depth > P.min_area_vaf) {
    if (lane == 0) {
      const int i = atomicAdd(n_active, 1);
      if (i < max_active) {
        active[2 * i] = tile.contig_id;
        active[2 * i + 1] = x;
      }
    }
    return;
  }
  // alleles whose alt bases are all standard, sorted by (ref string, alt string) (Likelihood.scala:37)
  GLAlleleDev mine_a;
  mine_a.key = tkey;
  mine_a.rep_read = tfirst;
  mine_a.rep_rp = trp;
  mine_a.indel_len = tlen;
  mine_a.kind = tkind;
  mine_a.base = tbase;
  mine_a._p0 = mine_a._p1 = 0;
  mine_a._pad = 0;
  s_alleles[warp][lane] = mine_a;
  __syncwarp();
  AlleleStrings me{&R, &ref, x, mine_a};
  const bool keep = lane < ntab && me.alt_all_standard();
  int rank = 0;
  if (keep)
    for (int j = 0; j < ntab; ++j) {
      if (j == lane) continue;
      AlleleStrings other{&R, &ref, x, s_alleles[warp][j]};
      if (other.alt_all_standard() && compare_alleles(other, me) < 0) ++rank;
    }
  const unsigned kept = __ballot_sync(FULL, keep);
  const int A = __popc(kept);
  if (A == 0) return;  // genotypeProbabilities.isEmpty (:240)
  if (A > GL_MAX_ALLELES) {
    if (lane == 0) report_error(err, DERR_ALLELE_TABLE_FULL, tile.contig_id, x, -1);
    return;
  }
  __syncwarp();
  if (keep) s_alleles[warp][32 - 1 - rank] = mine_a;  // park sorted entries at the top end, then move down
  __syncwarp();
  GLAlleleDev sorted_a = lane < A ? s_alleles[warp][32 - 1 - lane] : mine_a;
  __syncwarp();
  if (lane < A) s_alleles[warp][lane] = sorted_a;
  __syncwarp();

  // pass 2: stage (allele index, quality) per element
  int n_el = 0;
  for (int64_t base = tile.read_lo; base < tile.read_hi; base += 32) {
    const int64_t r = base + lane;
    const bool mine = r < tile.read_hi && __ldg(R.start + r) <= x && x < __ldg(R.end + r) && (int)__ldg(R.sample_id + r) == P.sample;
    Elem e;
    int ai = 0xFF, q = 0;
    if (mine && element_at(R, r, x, ref, e) == DERR_NONE) {
      const uint64_t key = any_allele_key(R, r, e);
      for (int j = 0; j < A; ++j)
        if (s_alleles[warp][j].key == key) ai = j;
      q = e.quality;
    }
    const unsigned m = __ballot_sync(FULL, mine);
    if (mine) {
      const int pos = n_el + __popc(m & ((1u << lane) - 1));
      if (pos < GL_ELEM_CACHE) s_elem[warp][pos] = (uint16_t)((ai << 8) | q);
    }
    n_el += __popc(m);
  }
  if (n_el > GL_ELEM_CACHE) {
    if (lane == 0) report_error(err, DERR_ALLELE_TABLE_FULL, tile.contig_id, x, -2);
    return;
  }
  __syncwarp();
  // genotype log-likelihoods: sum_e log(sum_{a in g} f_a P(e | a)) (Likelihood.scala:124-137), uniform prior
  int g = 0;
  for (int i = 0; i < A; ++i)
    for (int j = i; j < A; ++j, ++g) {
      double acc = 0;
      for (int k = lane; k < n_el; k += 32) {
        const int ai = s_elem[warp][k] >> 8, q = s_elem[warp][k] & 0xFF;
        const double p = c_success_probability[q];
        const double pi = ai == i ? p : 1 - p;
        if (i == j) {
          acc += log(pi * 1.0);
        } else {
          const double pj = ai == j ? p : 1 - p;
          acc += log(pi * 0.5 + pj * 0.5);
        }
      }
      acc = warp_sum(acc);
      if (lane == 0) s_ll[warp][g] = acc + log(1.0);
    }
  __syncwarp();
  if (lane != 0) return;
  const int G = g;
  // logNormalize (breeze softmax) and maxBy (first of equal maxima)
  double mx = -INFINITY;
  for (int k = 0; k < G; ++k) mx = fmax(mx, s_ll[warp][k]);
  double sm;
  if (isinf(mx)) {
    sm = mx;
  } else {
    double acc = 0;
    for (int k = 0; k < G; ++k) acc += exp(s_ll[warp][k] - mx);
    sm = mx + log(acc);
  }
  int best = 0;
  double best_v = s_ll[warp][0] - sm;
  for (int k = 1; k < G; ++k) {
    const double v = s_ll[warp][k] - sm;
    if (v > best_v) {
      best = k;
      best_v = v;
    }
  }
  int bi = 0, bj = 0;
  g = 0;
  for (int i = 0; i < A; ++i)
    for (int j = i; j < A; ++j, ++g)
      if (g == best) {
        bi = i;
        bj = j;
      }
  GLCallDev c;
  c.slot = tile.slot0;
  c.locus = x;
  c.read_lo = tile.read_lo;
  c.read_hi = tile.read_hi;
  c.contig_id = tile.contig_id;
  c.ref_index = tile.ref_index;
  c.n_genotypes = G;
  c.n_called = 0;
  c.log_probability = best_v;
  const int idxs[2] = {bi, bj};
  for (int t = 0; t < (bi == bj ? 1 : 2); ++t) {  // getNonReferenceAlleles with non-empty alt bases (:249-252)
    const GLAlleleDev a = s_alleles[warp][idxs[t]];
    const bool variant = a.kind != GUAC_ALN_MATCH && a.kind != GUAC_ALN_CLIPPED;
    if (variant && a.kind != GUAC_ALN_MID_DELETION) c.called[c.n_called++] = a;
  }
  if (c.n_called == 0) return;
  const int idx = atomicAdd(n_calls, 1);
  if (idx < max_calls) calls[idx] = c;
}

// Dense half for tiles the streaming kernel handed back (capacity overflow): one warp per locus,
// depth / referenceDepth by the exact cursor, candidates appended to the same heavy list.
__global__ void __launch_bounds__(CTA_THREADS)
germline_dense_fallback_kernel(ReadsView R, const RefView* __restrict__ refs, const Tile* __restrict__ tiles,
                               const int* __restrict__ tile_list, GLParamsDev P, uint8_t* __restrict__ visited,
                               guac_locus_summary* __restrict__ summary, Tile* __restrict__ heavy, int* __restrict__ n_heavy,
                               int max_heavy, unsigned long long* __restrict__ n_visited, DeviceError* err) {
  const Tile tile = tiles[tile_list[blockIdx.x]];
  const RefView ref = refs[tile.ref_index];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned FULL = 0xffffffffu;
  for (int li = warp; li < tile.n_loci; li += WARPS_PER_CTA) {
    const int64_t x = tile.locus0 + li, slot = tile.slot0 + li;
    int depth = 0, matches = 0;
    for (int64_t base = tile.read_lo; base < tile.read_hi; base += 32) {
      const int64_t r = base + lane;
      bool mine = r < tile.read_hi && __ldg(R.start + r) <= x && x < __ldg(R.end + r) && (int)__ldg(R.sample_id + r) == P.sample;
      Elem e;
      e.kind = GUAC_ALN_CLIPPED;
      if (mine) {
        const int ec = element_at(R, r, x, ref, e);
        if (ec != DERR_NONE) {
          report_error(err, ec, tile.contig_id, x, r);
          mine = false;
        }
      }
      depth += __popc(__ballot_sync(FULL, mine));
      matches += __popc(__ballot_sync(FULL, mine && e.kind == GUAC_ALN_MATCH));
    }
    if (lane != 0) continue;
    visited[slot] = depth > 0;
    if (depth > 0) atomicAdd(n_visited, 1ull);
    const bool candidate = depth > 0 && depth - matches >= P.min_alt_reads;
    if (candidate) {
      const int h = atomicAdd(n_heavy, 1);
      if (h < max_heavy) {
        Tile t = tile;
        t.locus0 = x;
        t.n_loci = 1;
        t.slot0 = slot;
        heavy[h] = t;
      }
    }
    guac_locus_summary z, s;
    z.depth = z.reference_depth = z.top_alt_count = z.flags = 0;
    s.depth = (uint16_t)min(depth, 65535);
    s.reference_depth = (uint16_t)min(matches, 65535);
    s.top_alt_count = 0;
    s.flags = candidate ? 1 : 0;
    summary[slot * 2 + 0] = z;
    summary[slot * 2 + 1] = s;
  }
}

// ---- generic AlleleEvidence ------------------------------------------------------------------------
struct EvidenceJob {
  int64_t locus, read_lo, read_hi;
  uint64_t key;  // any_allele_key of the allele
  int32_t contig_id, ref_index, sample, _pad;
  double probability;
};

__global__ void __launch_bounds__(128)
allele_evidence_kernel(ReadsView R, const RefView* __restrict__ refs, const EvidenceJob* __restrict__ jobs,
                       guac_allele_evidence* __restrict__ out, DeviceError* err) {
  __shared__ unsigned h_mapq[256], h_bq[256], h_mm[MM_BINS];
  __shared__ unsigned s_depth, s_pos, s_adepth, s_apos;
  const EvidenceJob job = jobs[blockIdx.x];
  const RefView ref = refs[job.ref_index];
  const int64_t x = job.locus;
  for (int i = threadIdx.x; i < 256; i += blockDim.x) h_mapq[i] = h_bq[i] = 0;
  for (int i = threadIdx.x; i < MM_BINS; i += blockDim.x) h_mm[i] = 0;
  if (threadIdx.x == 0) s_depth = s_pos = s_adepth = s_apos = 0;
  __syncthreads();
  for (int64_t r = job.read_lo + threadIdx.x; r < job.read_hi; r += blockDim.x) {
    if (!(__ldg(R.start + r) <= x && x < __ldg(R.end + r)) || (int)__ldg(R.sample_id + r) != job.sample) continue;
    Elem e;
    if (element_at(R, r, x, ref, e) != DERR_NONE) continue;
    const bool pos = __ldg(R.flags + r) & GUAC_READ_POSITIVE_STRAND;
    atomicAdd(&s_depth, 1u);
    if (pos) atomicAdd(&s_pos, 1u);
    if (any_allele_key(R, r, e) != job.key) continue;
    atomicAdd(&s_adepth, 1u);
    if (pos) atomicAdd(&s_apos, 1u);
    atomicAdd(&h_mapq[__ldg(R.mapq + r)], 1u);
    atomicAdd(&h_bq[e.quality], 1u);
    int mm = count_of_mismatches(R, r, ref, err, job.contig_id);
    if (mm >= MM_BINS) {
      report_error(err, DERR_MISMATCH_HIST, job.contig_id, x, r);
      mm = MM_BINS - 1;
    }
    atomicAdd(&h_mm[mm], 1u);
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  guac_allele_evidence ev;
  ev.probability = job.probability;
  ev.read_depth = (int)s_depth;
  ev.allele_read_depth = (int)s_adepth;
  ev.forward_depth = (int)s_pos;
  ev.allele_forward_depth = (int)s_apos;
  const int n = (int)s_adepth;
  if (n == 0) {
    const double nan = __longlong_as_double(0x7ff8000000000000ll);
    ev.mean_mapping_quality = ev.median_mapping_quality = ev.mean_base_quality = ev.median_base_quality =
        ev.median_mismatches_per_read = nan;
  } else {
    double sm = 0, sq = 0;
    for (int v = 0; v < 256; ++v) {
      sm += (double)v * h_mapq[v];
      sq += (double)v * h_bq[v];
    }
    ev.mean_mapping_quality = sm / n;
    ev.mean_base_quality = sq / n;
    int a, b;
    hist_middle(h_mapq, 256, n, a, b);
    ev.median_mapping_quality = (n & 1) ? (double)a : ((double)a + (double)b) / 2;
    hist_middle(h_bq, 256, n, a, b);
    ev.median_base_quality = (n & 1) ? (double)a : ((double)a + (double)b) / 2;
    hist_middle(h_mm, MM_BINS, n, a, b);
    ev.median_mismatches_per_read = (n & 1) ? (double)a : (double)((a + b) / 2);
  }
  out[blockIdx.x] = ev;
}

}  // namespace guac
