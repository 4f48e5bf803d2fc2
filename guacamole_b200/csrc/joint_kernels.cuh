// joint_kernels.cuh — somatic-joint per-locus kernels (SURVEY.md §8 rows A8–A12, A15 inputs).
//
//   sj_discover_kernel  dense, one warp per pileup locus x: AlleleAtLocus.variantAlleles at x + 1
//                       (jointcaller/AlleleAtLocus.scala:85-149) over ReadSubsequence.ofNextAltAllele
//                       (jointcaller/pileup_summarization/ReadSubsequence.scala:106-135).  Emits a site
//                       record only where an allele clears the support threshold or the site is force-called.
//   sj_evidence_kernel  sparse, one CTA per (site, allele): PileupStats of every sample and of the two pools
//                       (pileup_summarization/PileupStats.scala:23-91, MultiplePileupStats.scala:12-18):
//                       ReadSubsequence.ofFixedReferenceLength per element, allelic depths, strand counts and
//                       the log10 mixture likelihoods the evidence classes ask for
//                       (evidence/{NormalDNA,TumorDNA,TumorRNA}SingleSampleSingleAlleleEvidence.scala).
//   sj_strings_kernel   writes the allele strings of the emitted records into the result arena.
// These are the generic kernels: any allele shape, any number of samples.  Sites whose alleles are single-base
// substitutions take the fast site kernel of joint_stream_kernels.cuh instead, and posteriors, call logic, annotations
// and the emit policy follow on the device for both (sj_posterior_kernel, sj_final_kernel there).
// Reference file names below are relative to src/main/scala/org/hammerlab/guacamole/.
#pragma once
#include "kernels.cuh"

namespace guac {

constexpr int SJ_MAX_ALLELES = 7;
constexpr int SJ_MAX_SLOTS = GUAC_SJ_MAX_SAMPLES + 2;  // samples, pooled normal, pooled tumor
constexpr int SJ_TABLE = 512;                           // distinct (sample, sequence) pairs per (site, allele)

struct SJParamsDev {
  int n_samples;
  int kind[GUAC_SJ_MAX_SAMPLES];
  int n_discovery;
  int discovery[GUAC_SJ_MAX_SAMPLES];  // normal DNA samples, then tumor DNA samples (MultiSampleMultiAlleleEvidence.scala:97)
  int max_alleles, min_reads;
  double min_percent, vaf_floor;
};

struct SJAlleleDev {
  uint64_t key;      // seq_key of the alt sequence; 0 = the "ref > N" placeholder of a forced site
  int64_t rep_read;  // first supporting read
  int32_t span;      // reference length (endLocus - startLocus)
  int32_t alt_len;
  int32_t rep_start_rp;
  int32_t _pad;
};

struct SJSiteDev {
  int64_t slot, locus;  // pileup locus x; alleles start at x + 1
  int64_t read_lo, read_hi;
  int32_t contig_id, ref_index, n_alleles, forced;
  int32_t first_job, _pad;  // the site's alleles own the raw evidence records [first_job, first_job + n_alleles)
  SJAlleleDev alleles[SJ_MAX_ALLELES];
};

// 64-bit identity of a base sequence: exact 2-bit packing for <= 27 standard bases, FNV-1a of the ASCII otherwise
template <class GetBase>
__device__ inline uint64_t seq_key(int n, GetBase get, bool* all_standard = nullptr) {
  bool exact = n <= 27;
  bool standard = true;
  uint64_t packed = 0, h = 1469598103934665603ull;
  for (int k = 0; k < n; ++k) {
    const uint8_t b = get(k);
    const int code = b == 'A' ? 0 : b == 'C' ? 1 : b == 'G' ? 2 : b == 'T' ? 3 : -1;
    if (code < 0) {
      exact = false;
      standard = false;
    }
    if (k < 27) packed |= (uint64_t)(code & 3) << (2 * k);
    h = (h ^ b) * 1099511628211ull;
  }
  if (all_standard) *all_standard = standard;
  if (exact) return (1ull << 61) | ((uint64_t)n << 54) | packed;
  return (2ull << 61) | (h & ((1ull << 61) - 1));
}

__device__ __forceinline__ bool elem_is_variant(const Elem& e) { return e.kind != GUAC_ALN_MATCH && e.kind != GUAC_ALN_CLIPPED; }
// ReadSubsequence.isVariantOrFollowedByDeletion (:108-112)
__device__ __forceinline__ bool elem_vofd(const Elem& e) { return elem_is_variant(e) || e.followed_by_ref_skip; }

struct SubSeq {
  int64_t start_locus, end_locus;
  int32_t start_rp, end_rp;
};

// ReadSubsequence.ofNextAltAllele (:106-135); e0 is the element at x.  Returns 1 found, 0 none, < 0 -DERR.
__device__ inline int next_alt_allele(const ReadsView& R, int64_t r, int64_t x, const RefView& ref, const Elem& e0, SubSeq& s) {
  const int64_t end = __ldg(R.end + r);
  if (elem_vofd(e0) || x >= end - 1) return 0;
  Elem first, cur;
  int ec = element_at(R, r, x + 1, ref, first);
  if (ec != DERR_NONE) return -ec;
  cur = first;
  int64_t cx = x + 1;
  while (cx < end - 1 && elem_vofd(cur)) {
    ++cx;
    ec = element_at(R, r, cx, ref, cur);
    if (ec != DERR_NONE) return -ec;
  }
  if (cx == x + 1 || cx == end) return 0;
  s.start_locus = x + 1;
  s.end_locus = cx;
  s.start_rp = first.read_position;
  s.end_rp = cur.read_position;
  return 1;
}

// ReadSubsequence.ofFixedReferenceLength (:68-93)
__device__ inline int fixed_length_subsequence(const ReadsView& R, int64_t r, int64_t x, const RefView& ref, const Elem& e0, int length,
                                               SubSeq& s) {
  const int64_t end = __ldg(R.end + r);
  if (elem_is_variant(e0) || x >= end - 1) return 0;
  Elem first, cur;
  int ec = element_at(R, r, x + 1, ref, first);
  if (ec != DERR_NONE) return -ec;
  cur = first;
  int64_t cx = x + 1;
  int ref_offset = 1;
  while (cx < end - 1 && ref_offset < length) {
    ++cx;
    ec = element_at(R, r, cx, ref, cur);
    if (ec != DERR_NONE) return -ec;
    ++ref_offset;
  }
  if (cx >= end - 1) return 0;
  Elem after;
  ec = element_at(R, r, cx + 1, ref, after);
  if (ec != DERR_NONE) return -ec;
  s.start_locus = x + 1;
  s.end_locus = cx + 1;
  s.start_rp = first.read_position;
  s.end_rp = after.read_position;
  return 1;
}

// force-call loci as sorted ranges (contig, start, end); site x is forced when x + 1 lies in one
__device__ inline bool locus_forced(const guac_range* force, int n_force, int contig, int64_t locus1) {
  for (int i = 0; i < n_force; ++i)
    if (force[i].contig_id == contig && force[i].start <= locus1 && locus1 < force[i].end) return true;
  return false;
}

__global__ void __launch_bounds__(CTA_THREADS)
sj_discover_kernel(ReadsView R, const RefView* __restrict__ refs, const Tile* __restrict__ tiles,
                   const int* __restrict__ tile_list /* nullptr: identity */, int n_tiles,
                   int tile_per_warp /* 0: the CTA's warps share one tile; 1: one (single-locus) tile per warp */, SJParamsDev P,
                   const guac_range* __restrict__ force, int n_force, uint8_t* __restrict__ visited /* nullptr: already written */,
                   SJSiteDev* __restrict__ sites, int* __restrict__ n_sites, int max_sites, int* __restrict__ n_jobs, DeviceError* err) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tsel = tile_per_warp ? (int)blockIdx.x * WARPS_PER_CTA + warp : (int)blockIdx.x;
  if (tsel >= n_tiles) return;  // no block-wide barrier below
  Tile tile = tiles[tile_list ? tile_list[tsel] : tsel];
  const RefView ref = refs[tile.ref_index];
  const unsigned FULL = 0xffffffffu;
  if (tile_per_warp && tile.n_loci == 1) {
    // a single-locus tile arrives with the read span of the tile it came from: narrow it to the reads that can overlap
    // the locus (start <= x, running-max end > x); the passes below and the evidence kernel after them walk this span
    const int64_t x = tile.locus0, key = ((int64_t)tile.contig_id << 40) | x;
    tile.read_hi = warp_partition_point(tile.read_lo, tile.read_hi, lane, [&](int64_t i) { return __ldg(R.start + i) > x; });
    tile.read_lo = warp_partition_point(tile.read_lo, tile.read_hi, lane, [&](int64_t i) { return __ldg(R.runmax_key + i) > key; });
  }
  for (int li = tile_per_warp ? 0 : warp; li < tile.n_loci; li += tile_per_warp ? 1 : WARPS_PER_CTA) {
    const int64_t x = tile.locus0 + li, slot = tile.slot0 + li;
    // visited: any sample's window is non-empty (skipEmpty = true)
    bool any = false;
    for (int64_t base = tile.read_lo; base < tile.read_hi && !any; base += 32) {
      const int64_t r = base + lane;
      const bool ov = r < tile.read_hi && __ldg(R.start + r) <= x && x < __ldg(R.end + r) && __ldg(R.sample_id + r) < P.n_samples;
      any = __any_sync(FULL, ov);
    }
    if (lane == 0 && visited) visited[slot] = any ? 1 : 0;
    if (!any) continue;
    const uint8_t rb1 = ref.at(x + 1);
    if (rb1 == 0) {  // reference.getReferenceBase(contig, locus + 1) (MultiSampleMultiAlleleEvidence.scala:93)
      if (lane == 0) report_error(err, DERR_REF_MISSING, tile.contig_id, x + 1, -1);
      continue;
    }
    if (!(rb1 == 'A' || rb1 == 'C' || rb1 == 'G' || rb1 == 'T')) continue;
    const bool forced = locus_forced(force, n_force, tile.contig_id, x + 1);

    // result alleles (warp-uniform) and, for forced sites, support totals over all samples (one entry per lane)
    int n_res = 0;
    uint64_t res_key[SJ_MAX_ALLELES];
    int res_span[SJ_MAX_ALLELES], res_len[SJ_MAX_ALLELES], res_rp[SJ_MAX_ALLELES];
    int64_t res_first[SJ_MAX_ALLELES];
    uint64_t tot_key = 0;
    int tot_span = 0, tot_cnt = 0, tot_len = 0, tot_rp = 0, n_tot = 0;
    int64_t tot_first = 0;

    for (int di = 0; di < P.n_discovery; ++di) {
      const int smp = P.discovery[di];
      int depth = 0;
      uint64_t tkey = 0;
      int tspan = 0, tcnt = 0, tlen = 0, trp = 0, ntab = 0;
      int64_t tfirst = 0;
      for (int64_t base = tile.read_lo; base < tile.read_hi; base += 32) {
        const int64_t r = base + lane;
        bool mine = r < tile.read_hi && __ldg(R.start + r) <= x && x < __ldg(R.end + r) && (int)__ldg(R.sample_id + r) == smp;
        Elem e0;
        e0.kind = GUAC_ALN_CLIPPED;
        if (mine) {
          const int ec = element_at(R, r, x, ref, e0);
          if (ec != DERR_NONE) {
            report_error(err, ec, tile.contig_id, x, r);
            mine = false;
          }
        }
        mine = mine && e0.kind != GUAC_ALN_CLIPPED;  // Clipped elements are dropped first (:85-86)
        depth += __popc(__ballot_sync(FULL, mine));
        bool var = false;
        uint64_t key = 0;
        SubSeq s;
        s.start_locus = s.end_locus = 0;
        s.start_rp = s.end_rp = 0;
        if (mine) {
          const int st = next_alt_allele(R, r, x, ref, e0, s);
          if (st < 0) report_error(err, -st, tile.contig_id, x, r);
          if (st > 0) {
            const uint64_t s0 = __ldg(R.seq_off + r);
            bool standard;
            key = seq_key(s.end_rp - s.start_rp, [&](int k) { return read_base(R, s0 + s.start_rp + k); }, &standard);
            var = standard;  // onlyStandardBases (:105)
          }
        }
        const int span = (int)(s.end_locus - s.start_locus), alen = s.end_rp - s.start_rp;
        unsigned vm = __ballot_sync(FULL, var);
        while (vm) {  // groupBy((endLocus, sequence))
          const int L = __ffs(vm) - 1;
          const uint64_t k = __shfl_sync(FULL, key, L);
          const int sp = __shfl_sync(FULL, span, L), al = __shfl_sync(FULL, alen, L), rp = __shfl_sync(FULL, s.start_rp, L);
          const unsigned same = __ballot_sync(FULL, var && key == k && span == sp);
          const int c = __popc(same);
          const unsigned hit = __ballot_sync(FULL, lane < ntab && tkey == k && tspan == sp);
          if (hit) {
            if (lane == __ffs(hit) - 1) tcnt += c;
          } else if (ntab < 32) {
            if (lane == ntab) {
              tkey = k;
              tspan = sp;
              tcnt = c;
              tfirst = base + L;
              tlen = al;
              trp = rp;
            }
            ++ntab;
          } else if (lane == 0) {
            report_error(err, DERR_ALLELE_TABLE_FULL, tile.contig_id, x, base + L);
          }
          vm &= ~same;
        }
      }
      const double required = fmax((double)P.min_reads, depth * P.min_percent / 100.0);  // :99-101
      // visit the sample's alleles by (count desc, first read asc)
      int my_rank = 0;
      for (int j = 0; j < ntab; ++j) {
        const int cj = __shfl_sync(FULL, tcnt, j);
        const int64_t fj = __shfl_sync(FULL, tfirst, j);
        if (lane < ntab && (cj > tcnt || (cj == tcnt && fj < tfirst))) ++my_rank;
      }
      for (int i = 0; i < ntab; ++i) {
        const unsigned who = __ballot_sync(FULL, lane < ntab && my_rank == i);
        const int L = __ffs(who) - 1;
        const uint64_t k = __shfl_sync(FULL, tkey, L);
        const int sp = __shfl_sync(FULL, tspan, L), c = __shfl_sync(FULL, tcnt, L), al = __shfl_sync(FULL, tlen, L),
                  rp = __shfl_sync(FULL, trp, L);
        const int64_t f = __shfl_sync(FULL, tfirst, L);
        if ((double)c >= required) {  // .filter(_._3 >= _._2).map(_._1).distinct (:119-124)
          bool dup = false;
          for (int q = 0; q < n_res; ++q) dup = dup || (res_key[q] == k && res_span[q] == sp);
          if (!dup && n_res < P.max_alleles) {
            res_key[n_res] = k;
            res_span[n_res] = sp;
            res_len[n_res] = al;
            res_rp[n_res] = rp;
            res_first[n_res] = f;
            ++n_res;
          }
        }
        if (forced) {  // groupBy(allele) sums for atLeastOneAllele (:126-133)
          const unsigned hit = __ballot_sync(FULL, lane < n_tot && tot_key == k && tot_span == sp);
          if (hit) {
            if (lane == __ffs(hit) - 1) {
              tot_cnt += c;
              if (f < tot_first) {  // the representative read moves with the first supporting read
                tot_first = f;
                tot_rp = rp;
              }
            }
          } else if (n_tot < 32) {
            if (lane == n_tot) {
              tot_key = k;
              tot_span = sp;
              tot_cnt = c;
              tot_first = f;
              tot_len = al;
              tot_rp = rp;
            }
            ++n_tot;
          }
        }
      }
    }
    if (n_res == 0 && forced) {
      if (n_tot > 0) {  // the best supported allele overall
        int bc = lane < n_tot ? tot_cnt : -1;
        int64_t bf = lane < n_tot ? tot_first : INT64_MAX;
        int bl = lane;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
          const int oc = __shfl_xor_sync(FULL, bc, d);
          const int64_t of = __shfl_xor_sync(FULL, bf, d);
          const int ol = __shfl_xor_sync(FULL, bl, d);
          if (oc > bc || (oc == bc && of < bf)) {
            bc = oc;
            bf = of;
            bl = ol;
          }
        }
        res_key[0] = __shfl_sync(FULL, tot_key, bl);
        res_span[0] = __shfl_sync(FULL, tot_span, bl);
        res_len[0] = __shfl_sync(FULL, tot_len, bl);
        res_rp[0] = __shfl_sync(FULL, tot_rp, bl);
        res_first[0] = __shfl_sync(FULL, tot_first, bl);
      } else {  // AlleleAtLocus(contig, variantStart, refBase, "N") (:136-141)
        res_key[0] = 0;
        res_span[0] = 1;
        res_len[0] = 1;
        res_rp[0] = 0;
        res_first[0] = -1;
      }
      n_res = 1;
    }
    if (n_res > 0 && lane == 0) {
      const int idx = atomicAdd(n_sites, 1);
      const int job0 = atomicAdd(n_jobs, n_res);
      if (idx < max_sites) {
        SJSiteDev s;
        s.first_job = job0;
        s._pad = 0;
        s.slot = slot;
        s.locus = x;
        s.read_lo = tile.read_lo;
        s.read_hi = tile.read_hi;
        s.contig_id = tile.contig_id;
        s.ref_index = tile.ref_index;
        s.n_alleles = n_res;
        s.forced = forced ? 1 : 0;
        for (int q = 0; q < SJ_MAX_ALLELES; ++q) {
          SJAlleleDev a;
          a.key = q < n_res ? res_key[q] : 0;
          a.rep_read = q < n_res ? res_first[q] : -1;
          a.span = q < n_res ? res_span[q] : 0;
          a.alt_len = q < n_res ? res_len[q] : 0;
          a.rep_start_rp = q < n_res ? res_rp[q] : 0;
          a._pad = 0;
          s.alleles[q] = a;
        }
        sites[idx] = s;
      }
    }
  }
}

// ---- evidence -------------------------------------------------------------------------------------

struct SJDepthDev {
  uint64_t key;
  int64_t rep_read;
  int32_t rep_start_rp, len, count, forward;
};

struct SJSlotDev {  // one SingleSampleSingleAlleleEvidence's raw numbers
  int32_t total_depth;  // elements (Clipped removed), including those contributing no allele
  int32_t n_depths;
  int32_t n_subsequences, forward_subsequences;
  int32_t variant_count, variant_forward;  // subsequences equal to the alt allele
  int32_t ref_count, _pad;                 // allelicDepths(ref), untruncated
  double alt_vaf;                          // max(vaf floor, count(alt) / total_depth)
  double ll[3];                            // normal kind: ref/ref, ref/alt, alt/alt; tumor kind: {ref}, {alt: vaf, ref: 1 - vaf}
  SJDepthDev depths[GUAC_SJ_MAX_DEPTH_ENTRIES];
};

struct SJEvidenceDev {
  int32_t site, allele;
  SJSlotDev slots[SJ_MAX_SLOTS];
};

struct SJTableEntry {
  unsigned long long tag;  // 0 = empty
  uint64_t key;
  long long first;
  int sample, count, forward, rep_start_rp, len;
};

__device__ inline int sj_table_find(SJTableEntry* tab, uint64_t key, int sample, bool insert, int* overflow) {
  const unsigned long long tag = (key ^ ((unsigned long long)(sample + 1) * 0x9E3779B97F4A7C15ull)) | (1ull << 63);
  int slot = (int)((tag * 0xD6E8FEB86659FD93ull) >> 55) & (SJ_TABLE - 1);
  for (int tries = 0; tries < SJ_TABLE; ++tries) {
    unsigned long long old = tab[slot].tag;
    if (old == 0 && insert) old = atomicCAS(&tab[slot].tag, 0ull, tag);
    if (old == 0 && insert) {
      tab[slot].key = key;
      tab[slot].sample = sample;
      return slot;
    }
    if (old == tag) return slot;
    if (old == 0) return -1;
    slot = (slot + 1) & (SJ_TABLE - 1);
  }
  if (overflow) *overflow = 1;
  return -1;
}

// Grid: one CTA per (site, allele index) of the sites [site0, site0 + n) — blockIdx.x = site * max_alleles + allele;
// CTAs of allele indices the site does not have leave at once.  Record (site.first_job + allele) of `out` is written.
__global__ void __launch_bounds__(128)
sj_evidence_kernel(ReadsView R, const RefView* __restrict__ refs, SJParamsDev P, const SJSiteDev* __restrict__ sites, int site0,
                   unsigned char* __restrict__ out, size_t out_stride, DeviceError* err) {
  __shared__ SJTableEntry tab[SJ_TABLE];
  __shared__ int total_depth[SJ_MAX_SLOTS];
  __shared__ double vaf[SJ_MAX_SLOTS];
  __shared__ double ll[SJ_MAX_SLOTS][3];
  __shared__ int overflow;
  __shared__ double s_term[3][128];
  __shared__ uint8_t s_smp[128];
  const int2 job = make_int2(site0 + (int)(blockIdx.x / (unsigned)P.max_alleles), (int)(blockIdx.x % (unsigned)P.max_alleles));
  const SJSiteDev& site = sites[job.x];
  if (job.y >= site.n_alleles) return;
  const SJAlleleDev al = site.alleles[job.y];
  const RefView ref = refs[site.ref_index];
  const int64_t x = site.locus;
  const int S = P.n_samples, NP = S, TP = S + 1;
  const int span = al.span;
  for (int i = threadIdx.x; i < SJ_TABLE; i += blockDim.x) {
    tab[i].tag = 0;
    tab[i].count = tab[i].forward = 0;
    tab[i].first = LLONG_MAX;
    tab[i].rep_start_rp = tab[i].len = 0;
  }
  for (int i = threadIdx.x; i < SJ_MAX_SLOTS; i += blockDim.x) {
    total_depth[i] = 0;
    vaf[i] = 0;
    ll[i][0] = ll[i][1] = ll[i][2] = 0;
  }
  if (threadIdx.x == 0) overflow = 0;
  __syncthreads();
  // identity of the reference sequence over [x + 1, x + 1 + span) and of the alt allele
  const uint64_t ref_key = seq_key(span, [&](int k) { return ref.at(x + 1 + k); });
  const uint64_t alt_key = al.key != 0 ? al.key : seq_key(1, [&](int) { return (uint8_t)'N'; });

  // pass A: subsequences, allelic depths, strand counts
  for (int64_t r = site.read_lo + threadIdx.x; r < site.read_hi; r += blockDim.x) {
    if (!(__ldg(R.start + r) <= x && x < __ldg(R.end + r))) continue;
    const int smp = __ldg(R.sample_id + r);
    if (smp >= S) continue;
    Elem e0;
    int ec = element_at(R, r, x, ref, e0);
    if (ec != DERR_NONE) {
      report_error(err, ec, site.contig_id, x, r);
      continue;
    }
    if (e0.kind == GUAC_ALN_CLIPPED) continue;
    atomicAdd(&total_depth[smp], 1);
    SubSeq s;
    const int st = fixed_length_subsequence(R, r, x, ref, e0, span, s);
    if (st < 0) report_error(err, -st, site.contig_id, x, r);
    if (st <= 0) continue;
    const uint64_t s0 = __ldg(R.seq_off + r);
    const uint64_t key = seq_key(s.end_rp - s.start_rp, [&](int k) { return read_base(R, s0 + s.start_rp + k); });
    const int slot = sj_table_find(tab, key, smp, true, &overflow);
    if (slot >= 0) {
      atomicAdd(&tab[slot].count, 1);
      if (__ldg(R.flags + r) & GUAC_READ_POSITIVE_STRAND) atomicAdd(&tab[slot].forward, 1);
      atomicMin(&tab[slot].first, (long long)r);
    }
  }
  __syncthreads();
  if (overflow) {
    if (threadIdx.x == 0) report_error(err, DERR_ALLELE_TABLE_FULL, site.contig_id, x, -1);
    return;
  }
  if (threadIdx.x == 0) {  // alt fractions per evidence slot (TumorDNA...Evidence.scala:38)
    int pool_alt = 0, pool_total = 0;
    for (int s = 0; s < S; ++s) {
      const int slot = sj_table_find(tab, alt_key, s, false, nullptr);
      const int c = slot >= 0 ? tab[slot].count : 0;
      vaf[s] = fmax(P.vaf_floor, (double)c / (double)total_depth[s]);
      if (P.kind[s] == GUAC_SJ_TUMOR_DNA) {
        pool_alt += c;
        pool_total += total_depth[s];
      }
    }
    vaf[TP] = fmax(P.vaf_floor, (double)pool_alt / (double)pool_total);
  }
  __syncthreads();
  // pass B: log10 mixture likelihoods (PileupStats.logLikelihoodPileup :77-91).  Each chunk of 128 reads leaves its terms
  // in shared memory; one thread per (evidence slot, mixture) then adds them in read order, so the sums do not depend on
  // how the threads were scheduled.
  double my_sum = 0.0;
  const int my_slot = threadIdx.x / 3, my_k = threadIdx.x % 3;
  for (int64_t base = site.read_lo; base < site.read_hi; base += blockDim.x) {
    const int64_t r = base + threadIdx.x;
    s_smp[threadIdx.x] = 255;
    do {
      if (r >= site.read_hi) break;
      if (!(__ldg(R.start + r) <= x && x < __ldg(R.end + r))) break;
      const int smp = __ldg(R.sample_id + r);
      if (smp >= S) break;
      Elem e0;
      if (element_at(R, r, x, ref, e0) != DERR_NONE || e0.kind == GUAC_ALN_CLIPPED) break;
      SubSeq s;
      if (fixed_length_subsequence(R, r, x, ref, e0, span, s) <= 0) break;
      const uint64_t s0 = __ldg(R.seq_off + r);
      const int n = s.end_rp - s.start_rp;
      const uint64_t key = seq_key(n, [&](int k) { return read_base(R, s0 + s.start_rp + k); });
      const int slot = sj_table_find(tab, key, smp, false, nullptr);
      if (slot >= 0 && tab[slot].first == (long long)r) {  // representative read of this (sample, sequence)
        tab[slot].rep_start_rp = s.start_rp;
        tab[slot].len = n;
      }
      const int mapq = __ldg(R.mapq + r);
      if (mapq == 0) break;  // contributes exactly 0.0
      long qsum = 0;
      int qn = 0;
      if (n == 0) {
        qsum = __ldg(R.qual + s0 + s.start_rp);
        qn = 1;
      } else {
        for (int k = 0; k < n; ++k) qsum += __ldg(R.qual + s0 + s.start_rp + k);
        qn = n;
      }
      const int q = (int)floor((double)qsum / qn + 0.5);  // meanBaseQuality.round.toInt
      const double p = c_success_probability[q] * c_success_probability[mapq];
      const bool is_ref = key == ref_key, is_alt = key == alt_key;
      auto term = [&](double f) { return log10(f * p + (1 - f) * (1 - p)); };
      const int kind = P.kind[smp];
      s_smp[threadIdx.x] = (uint8_t)smp;
      s_term[0][threadIdx.x] = term(is_ref ? 1.0 : 0.0);
      if (kind == GUAC_SJ_NORMAL_DNA) {
        s_term[1][threadIdx.x] = term((is_ref || is_alt) ? 0.5 : 0.0);
        s_term[2][threadIdx.x] = term(is_alt ? 1.0 : 0.0);
      } else {
        const double v = vaf[smp], vp = vaf[TP];
        s_term[1][threadIdx.x] = term(is_alt ? v : (is_ref ? 1.0 - v : 0.0));
        s_term[2][threadIdx.x] = kind == GUAC_SJ_TUMOR_DNA ? term(is_alt ? vp : (is_ref ? 1.0 - vp : 0.0)) : 0.0;  // pooled tumor mixture
      }
    } while (false);
    __syncthreads();
    if (my_slot < S + 2) {
      for (int i = 0; i < (int)blockDim.x; ++i) {
        const int smp = s_smp[i];
        if (smp == 255) continue;
        const int kind = P.kind[smp];
        if (my_slot < S) {
          if (smp == my_slot && (my_k < 2 || kind == GUAC_SJ_NORMAL_DNA)) my_sum += s_term[my_k][i];
        } else if (my_slot == NP) {
          if (kind == GUAC_SJ_NORMAL_DNA) my_sum += s_term[my_k][i];
        } else if (kind == GUAC_SJ_TUMOR_DNA && my_k < 2) {
          my_sum += s_term[my_k == 0 ? 0 : 2][i];
        }
      }
    }
    __syncthreads();
  }
  if (my_slot < S + 2) ll[my_slot][my_k] = my_sum;
  __syncthreads();
  // The table is nearly empty (a handful of distinct sequences per site): its used entries are listed first, in table
  // order, so that the assembly below walks that handful instead of all SJ_TABLE entries once per slot.
  __shared__ uint16_t s_used[SJ_TABLE];
  __shared__ int s_warp_total[4], s_n_used;
  static_assert(SJ_TABLE == 4 * 128, "four table entries per thread of the 128-thread CTA");
  {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int cnt = 0;
    for (int k = 0; k < 4; ++k) cnt += tab[4 * threadIdx.x + k].tag != 0;
    int incl = cnt;
    for (int d = 1; d < 32; d <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += o;
    }
    if (lane == 31) s_warp_total[warp] = incl;
    __syncthreads();
    int pos = incl - cnt;
    for (int w = 0; w < warp; ++w) pos += s_warp_total[w];
    for (int k = 0; k < 4; ++k)
      if (tab[4 * threadIdx.x + k].tag != 0) s_used[pos++] = (uint16_t)(4 * threadIdx.x + k);
    if (threadIdx.x == 127) s_n_used = pos;
    __syncthreads();
  }
  if (threadIdx.x >= S + 2) return;
  // assemble: per-sample slots, then the pools (MultiplePileupStats.scala:12-18), one thread per slot
  // records are packed with only the S + 2 slots in use (out_stride bytes each)
  SJEvidenceDev& E = *reinterpret_cast<SJEvidenceDev*>(out + (size_t)(site.first_job + job.y) * out_stride);
  if (threadIdx.x == 0) {
    E.site = job.x;
    E.allele = job.y;
  }
  const int n_used = s_n_used;
  {
    const int slot_i = threadIdx.x;
    SJSlotDev& o = E.slots[slot_i];
    const bool pool = slot_i >= S;
    const int want_kind = slot_i == NP ? GUAC_SJ_NORMAL_DNA : GUAC_SJ_TUMOR_DNA;
    o.total_depth = 0;
    o.n_depths = 0;
    o.n_subsequences = o.forward_subsequences = o.variant_count = o.variant_forward = o.ref_count = 0;
    o._pad = 0;
    o.alt_vaf = pool ? vaf[slot_i] : vaf[slot_i];
    o.ll[0] = ll[slot_i][0];
    o.ll[1] = ll[slot_i][1];
    o.ll[2] = ll[slot_i][2];
    for (int s = 0; s < S; ++s)
      if (pool ? P.kind[s] == want_kind : s == slot_i) o.total_depth += total_depth[s];
    // merge the member samples' table entries by sequence (a handful of entries: quadratic is fine)
    SJDepthDev merged[32];
    int nm = 0;
    bool too_many = false;
    for (int u = 0; u < n_used; ++u) {
      const int i = s_used[u];
      const int s = tab[i].sample;
      if (!(pool ? P.kind[s] == want_kind : s == slot_i)) continue;
      o.n_subsequences += tab[i].count;
      o.forward_subsequences += tab[i].forward;
      int j = 0;
      for (; j < nm; ++j)
        if (merged[j].key == tab[i].key) break;
      if (j == nm) {
        if (nm == 32) {
          too_many = true;
          continue;
        }
        merged[nm].key = tab[i].key;
        merged[nm].rep_read = tab[i].first;
        merged[nm].rep_start_rp = tab[i].rep_start_rp;
        merged[nm].len = tab[i].len;
        merged[nm].count = 0;
        merged[nm].forward = 0;
        ++nm;
      } else if (tab[i].first < merged[j].rep_read) {
        merged[j].rep_read = tab[i].first;
        merged[j].rep_start_rp = tab[i].rep_start_rp;
        merged[j].len = tab[i].len;
      }
      merged[j].count += tab[i].count;
      merged[j].forward += tab[i].forward;
    }
    if (too_many) report_error(err, DERR_ALLELE_TABLE_FULL, site.contig_id, x, -1);
    for (int j = 0; j < nm; ++j) {
      if (merged[j].key == alt_key) {
        o.variant_count = merged[j].count;
        o.variant_forward = merged[j].forward;
      }
      if (merged[j].key == ref_key) o.ref_count = merged[j].count;
    }
    // truncatedAllelicDepths(maxAllelesPerSite + 1): by count desc, first supporting read asc
    const int keep = min(nm, P.max_alleles + 1);
    for (int k = 0; k < keep; ++k) {
      int best = -1;
      for (int j = 0; j < nm; ++j) {
        if (merged[j].count < 0) continue;
        if (best < 0 || merged[j].count > merged[best].count ||
            (merged[j].count == merged[best].count && merged[j].rep_read < merged[best].rep_read))
          best = j;
      }
      o.depths[k] = merged[best];
      merged[best].count = -1 - merged[best].count;  // consumed
    }
    o.n_depths = keep;
  }
}

// ---- strings ----------------------------------------------------------------------------------------
struct SJStringJob {
  int64_t a;         // read index, or first locus of a reference slice
  int32_t b;         // read position, or reference table index
  int32_t len;
  uint32_t arena_off;
  int32_t kind;      // 0 read bases, 1 reference slice, 2 the literal "N", 3 the single character `b`
};

__global__ void sj_strings_kernel(ReadsView R, const RefView* __restrict__ refs, const SJStringJob* __restrict__ jobs, int n,
                                  char* __restrict__ arena) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const SJStringJob j = jobs[i];
  char* dst = arena + j.arena_off;
  if (j.kind == 0) {
    const uint64_t s0 = __ldg(R.seq_off + j.a);
    for (int k = 0; k < j.len; ++k) dst[k] = (char)read_base(R, s0 + j.b + k);
  } else if (j.kind == 1) {
    const RefView ref = refs[j.b];
    for (int k = 0; k < j.len; ++k) dst[k] = (char)ref.at(j.a + k);
  } else if (j.kind == 3) {
    dst[0] = (char)j.b;
  } else {
    dst[0] = 'N';
  }
}

}  // namespace guac
