// joint_stream_kernels.cuh — somatic-joint on the streaming architecture, back half on the device.
//
//   sj_snv_site_kernel   one warp per candidate site the locus pass (ss_tile_kernel MODE 2) classified as "simple"
//                        (every alternate allele that can start at x + 1 is a single-base substitution).  One pass over
//                        the reads of the site does ReadSubsequence.ofNextAltAllele, AlleleAtLocus.variantAlleles,
//                        PileupStats of every sample and of both pools and the mixture likelihoods
//                        (jointcaller/pileup_summarization/*.scala, jointcaller/AlleleAtLocus.scala:85-149,
//                        evidence/{NormalDNA,TumorDNA,TumorRNA}SingleSampleSingleAlleleEvidence.scala): single-operator
//                        reads are classified from three packed bases, the others through element_at().  Sums are
//                        accumulated per lane and combined by fixed butterflies: the same input gives the same bits.
//                        A site that turns out not to be simple (multi-base run, indel, N) is handed to the generic
//                        kernels of joint_kernels.cuh.
//   sj_posterior_kernel  one thread per site: posteriors, germline / somatic call logic, StrandBias (Fisher),
//                        InsufficientNormal and the emit policy (evidence/MultiSampleSingleAlleleEvidence.scala:66-341,
//                        annotation/StrandBias.scala:47-58, annotation/InsufficientNormal.scala:21-53,
//                        filters/FishersExactTest.scala:12-19, evidence/MultiSampleMultiAlleleEvidence.scala:131-177);
//                        only emitted sites are listed.
//   sj_final_kernel      one thread per emitted site, in (range order, locus) order: the finished guac_sj_* records and
//                        their strings, so that only emitted records cross PCIe.
// Reference file names are relative to src/main/scala/org/hammerlab/guacamole/.
#pragma once
#include "joint_kernels.cuh"
#include "tile_kernels.cuh"

namespace guac {

// ---- posteriors / annotations / emit policy ---------------------------------------------------------

// ln C(n, k) as commons-math3 3.0 ArithmeticUtils.binomialCoefficientLog returns it to 1e-15 relative: its special cases
// literally, everything else from a table of ln n! (correctly rounded sums made on the host in extended precision).
__device__ __forceinline__ double sj_binomial_log(const double* __restrict__ lf, int lf_n, int n, int k) {
  if (n == k || k == 0) return 0.0;
  if (k == 1 || k == n - 1) return log((double)n);
  if (n < lf_n) return __ldg(lf + n) - __ldg(lf + k) - __ldg(lf + (n - k));
  return lgamma((double)n + 1.0) - lgamma((double)k + 1.0) - lgamma((double)(n - k) + 1.0);
}

struct SJAlleleCall {
  int is_germline, is_somatic, germline_genotype, rna_expressed;
  int has_in, in_ref, in_total, in_filtered;
  unsigned failing;
  __device__ bool is_call() const { return is_germline || is_somatic; }
};

// One MultiSampleSingleAlleleEvidence from the raw per-slot numbers.  WRITE: also fills the S + 2 sample records.
template <bool WRITE>
__device__ inline void sj_evaluate_allele(const guac_sj_params& P, const double* __restrict__ lf, int lf_n,
                                          const SJSlotDev* __restrict__ slots, SJAlleleCall& call,
                                          guac_sj_sample_evidence* __restrict__ out) {
  const int S = P.n_samples, NP = S, TP = S + 1;
  const double LN10 = 2.302585092994046;  // math.log(10)
  unsigned sb_filtered = 0, triggered = 0;
  auto kind_of = [&](int i) { return i == NP ? (int)GUAC_SJ_NORMAL_DNA : i == TP ? (int)GUAC_SJ_TUMOR_DNA : P.sample_kind[i]; };
  for (int i = 0; i < S + 2; ++i) {
    const SJSlotDev& r = slots[i];
    // StrandBias (annotation/StrandBias.scala:47-58) over FishersExactTest.asLog10
    const int vf = r.variant_forward, vr = r.variant_count - r.variant_forward;
    const int tf = r.forward_subsequences, tr = r.n_subsequences - r.forward_subsequences;
    const double sb = 10.0 * ((sj_binomial_log(lf, lf_n, tf, vf) + sj_binomial_log(lf, lf_n, tr, vr) -
                               sj_binomial_log(lf, lf_n, tf + tr, vf + vr)) / LN10);
    if (sb > P.filter_strand_bias_phred) sb_filtered |= 1u << i;
    if (WRITE) {
      guac_sj_sample_evidence& e = out[i];
      const int kind = kind_of(i);
      e.kind = kind;
      e.total_depth = r.total_depth;
      e.n_depths = r.n_depths;
      e.n_likelihoods = kind == GUAC_SJ_NORMAL_DNA ? 3 : 2;
      for (int k = 0; k < GUAC_SJ_MAX_DEPTH_ENTRIES; ++k) {
        e.depths[k].seq_off = e.depths[k].seq_len = 0;
        e.depths[k].count = k < r.n_depths ? r.depths[k].count : 0;
        e.depths[k]._pad = 0;
      }
      for (int k = 0; k < 3; ++k) {
        e.log_likelihoods[k] = k < e.n_likelihoods ? r.ll[k] : 0.0;
        e.posteriors[k] = __longlong_as_double(0x7ff8000000000000ll);  // NaN until a posterior is defined
      }
      e.alt_vaf = kind == GUAC_SJ_NORMAL_DNA ? 0.0 : r.alt_vaf;
      e.sb_variant_forward = vf;
      e.sb_variant_reverse = vr;
      e.sb_total_forward = tf;
      e.sb_total_reverse = tr;
      e.sb_phred = sb;
      e.sb_filtered = (sb_filtered >> i) & 1u;
      e.triggered = 0;
    }
  }
  // germline posteriors: likelihood - prior (MultiSampleSingleAlleleEvidence.scala:66-123)
  const double gprior[3] = {0.0, P.germline_neg_log10_het_prior, P.germline_neg_log10_hom_alt_prior};
  if (WRITE)
    for (int i = 0; i <= NP; ++i)
      if (kind_of(i) == GUAC_SJ_NORMAL_DNA)
        for (int g = 0; g < 3; ++g) out[i].posteriors[g] = slots[i].ll[g] - gprior[g];
  int germline = 0;  // maxBy keeps the first of equal maxima
  {
    double best = slots[NP].ll[0] - gprior[0];
    for (int g = 1; g < 3; ++g) {
      const double v = slots[NP].ll[g] - gprior[g];
      if (v > best) {
        best = v;
        germline = g;
      }
    }
  }
  call.germline_genotype = germline;
  call.is_germline = germline != 0;
  // A tumor mixture is {ref: 1} or {alt: vaf, ref: 1 - vaf}; its prior depends on which alleles have a positive fraction
  // (:172-183), whether it triggers on its key set against the germline alleles (:220-224).
  auto somatic_prior = [](int mixture, double vaf, double variant_prior) {
    if (mixture == 0) return 0.0;
    const bool has_alt = vaf > 0, has_ref = 1.0 - vaf > 0;
    if (has_ref && !has_alt) return 0.0;
    if (has_ref && has_alt) return variant_prior;
    return 1.7976931348623157e308;  // Double.MaxValue
  };
  auto best_mixture = [&](int i, double variant_prior) {
    const SJSlotDev& r = slots[i];
    const double p0 = r.ll[0] - somatic_prior(0, r.alt_vaf, variant_prior), p1 = r.ll[1] - somatic_prior(1, r.alt_vaf, variant_prior);
    if (WRITE) {
      out[i].posteriors[0] = p0;
      out[i].posteriors[1] = p1;
    }
    return p1 > p0 ? 1 : 0;
  };
  auto triggers = [&](int best) { return best == 0 ? germline != 0 : germline != 1; };
  bool rna_expressed = false;
  for (int i = 0; i < S; ++i) {
    if (kind_of(i) != GUAC_SJ_TUMOR_RNA) continue;
    if (triggers(best_mixture(i, P.somatic_neg_log10_variant_prior_rna))) {
      triggered |= 1u << i;
      rna_expressed = true;
    }
  }
  call.rna_expressed = rna_expressed;
  const double dna_prior = rna_expressed ? P.somatic_neg_log10_variant_prior_with_rna_evidence : P.somatic_neg_log10_variant_prior;
  bool any_triggered = false;
  for (int i = 0; i <= TP; ++i) {
    if (i == NP || kind_of(i) != GUAC_SJ_TUMOR_DNA) continue;
    if (triggers(best_mixture(i, dna_prior))) {
      triggered |= 1u << i;
      any_triggered = true;
    }
  }
  call.is_somatic = !call.is_germline && any_triggered;
  // InsufficientNormal (annotation/InsufficientNormal.scala:21-53): only for non-germline alleles
  call.has_in = call.in_ref = call.in_total = call.in_filtered = 0;
  if (!call.is_germline) {
    call.has_in = 1;
    call.in_ref = slots[NP].ref_count;
    call.in_total = slots[NP].total_depth;
    call.in_filtered = call.in_total < P.filter_somatic_normal_depth ||
                       (call.in_total - call.in_ref) * 100.0 / call.in_total > P.filter_somatic_normal_nonreference_percent;
  }
  // failingFilterNames (:250-275)
  bool have_trigger = false, all_fail = true;
  if (call.is_germline) {
    have_trigger = true;
    all_fail = (sb_filtered >> NP) & 1u;
  } else if (call.is_somatic) {
    for (int i = 0; i <= TP; ++i)
      if (i != NP && kind_of(i) == GUAC_SJ_TUMOR_DNA && ((triggered >> i) & 1u)) {
        have_trigger = true;
        all_fail = all_fail && ((sb_filtered >> i) & 1u);
      }
  }
  call.failing = 0;
  if (have_trigger && all_fail) call.failing |= 1u;
  if (call.has_in && call.in_filtered) call.failing |= 2u;
  if (WRITE)
    for (int i = 0; i < S + 2; ++i) out[i].triggered = (triggered >> i) & 1u;
}

// MultiSampleMultiAlleleEvidence.make's emit policy (:131-177): bit a of the result = allele a is emitted
__device__ inline unsigned sj_emit_mask(const guac_sj_params& P, bool forced, const SJAlleleCall* calls, int n_alleles) {
  unsigned evidences = 0;
  bool any_germline = false;
  for (int a = 0; a < n_alleles; ++a)
    if (calls[a].is_germline || calls[a].is_somatic) {
      evidences |= 1u << a;
      any_germline = any_germline || calls[a].is_germline;
    }
  const bool any_call = evidences != 0;
  if (!any_call) evidences = 1u;  // allEvidences.take(1)
  if (!forced) {
    if (!any_call) return 0;
    if (P.only_somatic && any_germline) return 0;
  }
  if (forced || P.include_filtered) return evidences;
  unsigned passing = 0;
  for (int a = 0; a < n_alleles; ++a)
    if (((evidences >> a) & 1u) && calls[a].failing == 0) passing |= 1u << a;
  return passing;
}

constexpr int SJ_SITE_GENERIC = 1 << 30;  // SJEmitDev::site: the site is in the generic kernels' list, not the fast kernel's

struct SJEmitDev {  // one emitted site
  int64_t slot;
  int32_t site;       // index into its site list (| SJ_SITE_GENERIC)
  int32_t n_emit;     // alleles emitted
  uint32_t mask;      // which ones
  uint32_t str_len;   // bytes of their strings
};

__device__ __forceinline__ const SJEvidenceDev* sj_raw_record(const unsigned char* raw, size_t stride, int job) {
  return reinterpret_cast<const SJEvidenceDev*>(raw + (size_t)job * stride);
}

__global__ void __launch_bounds__(128)
sj_posterior_kernel(const __grid_constant__ guac_sj_params P, const double* __restrict__ lf, int lf_n, const SJSiteDev* __restrict__ sites,
                    int n_sites, int site_tag, const unsigned char* __restrict__ raw, size_t raw_stride, SJEmitDev* __restrict__ emits,
                    int* __restrict__ n_emits) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_sites) return;
  const SJSiteDev& site = sites[i];
  const int S = P.n_samples;
  SJAlleleCall calls[SJ_MAX_ALLELES];
  for (int a = 0; a < site.n_alleles; ++a)
    sj_evaluate_allele<false>(P, lf, lf_n, sj_raw_record(raw, raw_stride, site.first_job + a)->slots, calls[a], nullptr);
  const unsigned mask = sj_emit_mask(P, site.forced != 0, calls, site.n_alleles);
  if (mask == 0) return;
  uint32_t len = 0;
  for (int a = 0; a < site.n_alleles; ++a) {
    if (!((mask >> a) & 1u)) continue;
    len += (uint32_t)site.alleles[a].span + (uint32_t)(site.alleles[a].key == 0 ? 1 : site.alleles[a].alt_len);
    const SJSlotDev* slots = sj_raw_record(raw, raw_stride, site.first_job + a)->slots;
    for (int s = 0; s < S + 2; ++s)
      for (int k = 0; k < slots[s].n_depths; ++k) len += (uint32_t)slots[s].depths[k].len;
  }
  SJEmitDev e;
  e.slot = site.slot;
  e.site = i | site_tag;
  e.n_emit = __popc(mask);
  e.mask = mask;
  e.str_len = len;
  emits[atomicAdd(n_emits, 1)] = e;
}

// rank -> per-rank arrays for the two offset scans
__global__ void sj_order_kernel(const SJEmitDev* __restrict__ emits, const int* __restrict__ rank, int n, int* __restrict__ order,
                                int32_t* __restrict__ n_by_rank, int32_t* __restrict__ len_by_rank) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int r = rank[i];
  order[r] = i;
  n_by_rank[r] = emits[i].n_emit;
  len_by_rank[r] = (int32_t)emits[i].str_len;
}

// A string of a record: bases of a read, or (rep_read < 0) the single character in rep_start_rp
__device__ inline void sj_write_read_string(const ReadsView& R, int64_t rep_read, int32_t rep_start_rp, int len, char* dst) {
  if (rep_read < 0) {
    dst[0] = (char)rep_start_rp;
    return;
  }
  const uint64_t s0 = __ldg(R.seq_off + rep_read);
  for (int k = 0; k < len; ++k) dst[k] = (char)read_base(R, s0 + rep_start_rp + k);
}

__global__ void __launch_bounds__(128)
sj_final_kernel(ReadsView R, const RefView* __restrict__ refs, const __grid_constant__ guac_sj_params P, const double* __restrict__ lf,
                int lf_n, const SJSiteDev* __restrict__ sites_fast, const SJSiteDev* __restrict__ sites_gen,
                const unsigned char* __restrict__ raw_fast, const unsigned char* __restrict__ raw_gen, size_t raw_stride,
                const SJEmitDev* __restrict__ emits, const int* __restrict__ order, int n_emits,
                const int64_t* __restrict__ allele_off, const int64_t* __restrict__ string_off,
                guac_sj_allele_evidence* __restrict__ out_alleles, guac_sj_sample_evidence* __restrict__ out_samples,
                char* __restrict__ arena) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_emits) return;
  const SJEmitDev em = emits[order[j]];
  const bool generic = (em.site & SJ_SITE_GENERIC) != 0;
  const SJSiteDev& site = (generic ? sites_gen : sites_fast)[em.site & ~SJ_SITE_GENERIC];
  const unsigned char* raw = generic ? raw_gen : raw_fast;
  const RefView ref = refs[site.ref_index];
  const int S = P.n_samples;
  int64_t ai = allele_off[j];
  uint32_t so = (uint32_t)string_off[j];
  for (int a = 0; a < site.n_alleles; ++a) {
    if (!((em.mask >> a) & 1u)) continue;
    const SJAlleleDev& al = site.alleles[a];
    const SJSlotDev* slots = sj_raw_record(raw, raw_stride, site.first_job + a)->slots;
    guac_sj_sample_evidence* samples = out_samples + ai * (S + 2);
    SJAlleleCall call;
    sj_evaluate_allele<true>(P, lf, lf_n, slots, call, samples);
    guac_sj_allele_evidence o;
    o.contig_id = site.contig_id;
    o.site_index = j;
    o.start = site.locus + 1;
    o.ref_off = so;
    o.ref_len = (uint32_t)al.span;
    for (int k = 0; k < al.span; ++k) arena[so + k] = (char)ref.at(site.locus + 1 + k);
    so += (uint32_t)al.span;
    o.alt_off = so;
    if (al.key == 0) {  // the "ref > N" placeholder of a forced site (AlleleAtLocus.scala:136-141)
      o.alt_len = 1;
      arena[so] = 'N';
    } else {
      o.alt_len = (uint32_t)al.alt_len;
      sj_write_read_string(R, al.rep_read, al.rep_start_rp, al.alt_len, arena + so);
    }
    so += o.alt_len;
    o.is_germline_call = call.is_germline;
    o.is_somatic_call = call.is_somatic;
    o.germline_genotype = call.germline_genotype;
    o.tumor_rna_expressed = call.rna_expressed;
    o.has_insufficient_normal = call.has_in;
    o.in_reference_reads = call.in_ref;
    o.in_total_reads = call.in_total;
    o.in_filtered = call.in_filtered;
    o.failing_filters = call.failing;
    o.forced = site.forced;
    o.first_sample_evidence = ai * (S + 2);
    out_alleles[ai] = o;
    for (int s = 0; s < S + 2; ++s)
      for (int k = 0; k < slots[s].n_depths; ++k) {
        const SJDepthDev& d = slots[s].depths[k];
        samples[s].depths[k].seq_off = so;
        samples[s].depths[k].seq_len = (uint32_t)d.len;
        sj_write_read_string(R, d.rep_read, d.rep_start_rp, d.len, arena + so);
        so += (uint32_t)d.len;
      }
    ++ai;
  }
}

// ---- the fast site kernel ---------------------------------------------------------------------------

constexpr int SJF_SMAX = 4;         // samples the fast site kernel handles (more: every candidate takes the generic kernels)
constexpr int SJF_MAX_READS = 256;  // reads in the span of a site
constexpr int SJF_LIST = 256;       // subsequences of a site that differ from the reference base (a heterozygous site has many)

struct SJFListEntry {  // (P(correct) and its two logarithms are looked up again from the quality and mapq)
  uint16_t rel;        // read index relative to the site's first read
  uint8_t s, c, flags; // sample, base code, 1: positive strand | 2: mapq > 0 | 4: starts an alternate allele
  uint8_t q, mapq, _pad;
};

struct SJFWarpSmem {
  double p[SJF_MAX_READS];      // P(correct), log10 p, log10 (1 - p) of the reads that have a subsequence and mapq > 0
  double A[SJF_MAX_READS];
  double B[SJF_MAX_READS];
  uint8_t meta[SJF_MAX_READS];  // sample | 4: such a read | base code << 3
  SJFListEntry list[SJF_LIST];
  uint16_t slow[SJF_LIST];      // reads the exact cursor walks (several operators, N skips, non-standard bases)
};

struct SJFArgs {
  const uint4* hdr;
  const double2* log_table; // sj_log_table_kernel
  const RefView* refs;
  const Tile* cands;        // simple candidates as single-locus tiles (the read span of the tile they came from)
  const int* n_cands;
  int max_cands;
  const guac_range* force;
  int n_force;
  SJSiteDev* sites;
  int* n_sites;
  int max_sites;
  unsigned char* raw;
  size_t raw_stride;
  int* n_jobs;
  int max_jobs;
  Tile* bail;               // candidates handed to the generic kernels
  int* n_bail;
  int max_bail;
  DeviceError* err;
};

__device__ __forceinline__ double warp_sum_fixed(double v) {  // the same tree for every call: deterministic
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
  return v;
}

template <class T>
__device__ __forceinline__ T pick4(T a0, T a1, T a2, T a3, int i) { return i == 0 ? a0 : i == 1 ? a1 : i == 2 ? a2 : a3; }

// log10 p and log10 (1 - p) of p = P(quality) * P(mapq) (PileupStats.scala:83-86) for every (mapq, quality) pair: the
// fast site kernel looks its per-read terms up instead of taking two logarithms per read (1 MB, L2-resident).
__global__ void sj_log_table_kernel(double2* __restrict__ table) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 65536) return;
  const double p = c_success_probability[i & 255] * c_success_probability[i >> 8];
  table[i] = make_double2(log10(p), log10(1.0 - p));
}

// One read through the exact cursor, out of line (the fast kernel's code stays small): bit 0 in depth, bit 1 has a
// one-base subsequence (code in bits 8..9, quality in bits 16..23), bit 2 that base starts a one-base alternate allele,
// bit 3 starts an allele of another shape, bit 4 the site cannot be decided here.
__device__ __noinline__ unsigned sjf_classify_slow(const ReadsView& R, int64_t r, int64_t x, const RefView& ref, int cr) {
  Elem e0;
  if (element_at(R, r, x, ref, e0) != DERR_NONE) return 16u;  // the generic kernels report it
  if (e0.kind == GUAC_ALN_CLIPPED) return 0u;                 // Clipped elements are dropped first (:85-86)
  unsigned out = 1u;
  SubSeq sa, sf;
  const int fa = next_alt_allele(R, r, x, ref, e0, sa);
  const int ff = fixed_length_subsequence(R, r, x, ref, e0, 1, sf);
  if (fa < 0 || ff < 0) return 16u;
  const uint64_t s0 = __ldg(R.seq_off + r);
  int c = 4;
  if (ff > 0) {
    const uint8_t b = sf.end_rp - sf.start_rp == 1 ? read_base(R, s0 + sf.start_rp) : (uint8_t)0;
    c = base_code(b);
    if (c == 4) return 16u;  // a subsequence that is not one standard base (an insertion, an N)
    out |= 2u | ((unsigned)c << 8) | ((unsigned)__ldg(R.qual + s0 + sf.start_rp) << 16);
  }
  if (fa > 0) {  // a one-base allele is this very base over one locus; anything else is another shape
    if (sa.end_locus - sa.start_locus == 1 && sa.end_rp - sa.start_rp == 1 && ff > 0 && sa.start_rp == sf.start_rp && c != cr) out |= 4u;
    else out |= 8u;
  }
  return out;
}

// The same answer for most reads with several operators (RF_MEDIUM: M = X I D S H over standard bases, CIGAR consistent
// with the sequence) without the cursor: when x .. x + 3 lie inside one M operator, the elements at x, x + 1, x + 2 are
// plain Match / Mismatch elements and the read goes on past them — exactly the single-operator case, with the read
// position taken from the operator.  Returns false when the loci are not inside one M operator (the cursor decides).
__device__ __forceinline__ bool sjf_classify_medium(const ReadsView& R, int64_t r, const uint4& h, int64_t x, uint8_t rb0, uint8_t rb2, int cr,
                                                    unsigned& cls) {
  const int off = (int)((uint32_t)x - h.x);  // x - start (the caller checked the overlap: 0 <= off < span)
  const uint32_t c0 = __ldg(R.cigar_off + r), c1 = __ldg(R.cigar_off + r + 1);
  int cl = 0, rp = 0, p0 = -1;
  for (uint32_t ci = c0; ci < c1 && cl <= off; ++ci) {
    const uint32_t w = __ldg(R.cigar + ci);
    const int op = (int)(w & 15), len = (int)(w >> 4);
    if (op == OPC_M) {
      if (off >= cl && off + 3 < cl + len) p0 = rp + (off - cl);
      cl += len;
      rp += len;
    } else if (op == OPC_EQ || op == OPC_X) {
      cl += len;
      rp += len;
    } else if (op == OPC_D) {
      cl += len;
    } else if (op == OPC_I || op == OPC_S) {
      rp += len;
    }
  }
  if (p0 < 0) return false;
  const uint64_t s0 = (uint64_t)h.z | ((uint64_t)h.w << 32);
  auto base_at = [&](int q) {
    const uint64_t gi = s0 + (uint64_t)q;
    return (int)((__ldg(R.bases2 + (gi >> 2)) >> ((gi & 3) * 2)) & 3);
  };
  cls = 1u;
  if ("ACGT"[base_at(p0)] == (char)rb0) {
    const int c = base_at(p0 + 1);
    cls |= 2u | ((unsigned)c << 8) | ((unsigned)__ldg(R.qual + s0 + p0 + 1) << 16);
    if (c != cr) cls |= "ACGT"[base_at(p0 + 2)] != (char)rb2 ? 8u : 4u;
  }
  return true;
}

#ifndef GUAC_SJF_MINB
#define GUAC_SJF_MINB 2
#endif
#ifndef GUAC_SJF_WARPS
#define GUAC_SJF_WARPS 8
#endif
constexpr int SJF_WARPS = GUAC_SJF_WARPS;  // sites in flight per CTA, a warp each
#ifdef GUAC_TK_PHASE_CLOCKS
#define SJF_PHASE(k)                                                                      \
  do {                                                                                    \
    if (threadIdx.x == 0) {                                                               \
      const long long now_ = clock64();                                                   \
      atomicAdd(&g_tk_phase_clocks[16 + (k)], (unsigned long long)(now_ - sjf_t_));       \
      sjf_t_ = now_;                                                                      \
    }                                                                                     \
  } while (0)
#else
#define SJF_PHASE(k) do { } while (0)
#endif
__global__ void __launch_bounds__(32 * SJF_WARPS, GUAC_SJF_MINB)
sj_snv_site_kernel(ReadsView R, const __grid_constant__ SJParamsDev P, const __grid_constant__ guac_sj_params EP,
                   const __grid_constant__ SJFArgs A) {
  extern __shared__ __align__(16) unsigned char sjf_smem_raw[];  // SJF_WARPS x SJFWarpSmem (dynamic: beyond the 48 KB static limit)
  const unsigned FULL = 0xffffffffu;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  SJFWarpSmem& W = reinterpret_cast<SJFWarpSmem*>(sjf_smem_raw)[warp];
  const int S = P.n_samples, NP = S, TP = S + 1;
  const int n_cands = min(*A.n_cands, A.max_cands);
  // sample kinds and the discovery order as packed words: no dynamically indexed parameter arrays in the loops
  unsigned kinds = 0, disc_order = 0, disc_mask = 0;
  bool has_rna = false;
  for (int s = 0; s < S; ++s) {
    kinds |= (unsigned)P.kind[s] << (2 * s);
    has_rna = has_rna || P.kind[s] == GUAC_SJ_TUMOR_RNA;
  }
  for (int d = 0; d < P.n_discovery; ++d) {
    disc_order |= (unsigned)P.discovery[d] << (4 * d);
    disc_mask |= 1u << P.discovery[d];
  }
  const double LOG10_HALF = -0.3010299956639812;  // math.log10(0.5)

  for (int ci = blockIdx.x * SJF_WARPS + warp; ci < n_cands; ci += gridDim.x * SJF_WARPS) {
    Tile tile = A.cands[ci];
#ifdef GUAC_TK_PHASE_CLOCKS
    long long sjf_t_ = clock64();
#endif
    const RefView ref = A.refs[tile.ref_index];
    const int64_t x = tile.locus0, key = ((int64_t)tile.contig_id << 40) | x;
    // reads that can overlap x: start <= x, running-max end > x
    tile.read_hi = warp_partition_point(tile.read_lo, tile.read_hi, lane, [&](int64_t i) { return __ldg(R.start + i) > x; });
    tile.read_lo = warp_partition_point(tile.read_lo, tile.read_hi, lane, [&](int64_t i) { return __ldg(R.runmax_key + i) > key; });
    const int n_span = (int)min((int64_t)(tile.read_hi - tile.read_lo), (int64_t)(SJF_MAX_READS + 1));
    const uint8_t rb0 = ref.at(x), rb1 = ref.at(x + 1), rb2 = ref.at(x + 2);
    if (rb1 == 0) {  // reference.getReferenceBase(contig, locus + 1) (MultiSampleMultiAlleleEvidence.scala:93)
      if (lane == 0) report_error(A.err, DERR_REF_MISSING, tile.contig_id, x + 1, -1);
      continue;
    }
    const int cr = base_code(rb1);
    if (cr == 4) continue;  // not a standard base: no evidence at this site
    // force-called sites (their fallback alleles can have any shape) and spans beyond the stash take the generic kernels
    bool bail = n_span > SJF_MAX_READS || rb0 == 0 || rb2 == 0 || locus_forced(A.force, A.n_force, tile.contig_id, x + 1);

    SJF_PHASE(0);
    int depth[SJF_SMAX], rcnt[SJF_SMAX], rfwd[SJF_SMAX], rfirst[SJF_SMAX], rll[SJF_SMAX], ns_alt[SJF_SMAX];
    double rA[SJF_SMAX], rB[SJF_SMAX];
#pragma unroll
    for (int s = 0; s < SJF_SMAX; ++s) {
      depth[s] = rcnt[s] = rfwd[s] = rll[s] = ns_alt[s] = 0;
      rfirst[s] = INT_MAX;
      rA[s] = rB[s] = 0.0;
    }
    int n_list = 0, n_slow = 0;
    const uint32_t x_lo = (uint32_t)x;
    // One read's contribution.  in_depth: it has a (non-Clipped) element at x; has_sub: its fixed-length-1 subsequence
    // exists and is the standard base c with quality q; disc: that base starts a one-base alternate allele; other_alt:
    // the read starts an alternate allele of another shape (counted: it only matters if a sample has enough of them).
    bool low_p = false;  // a tumor read on the reference base with P(correct) <= 1/2: the call bound below does not hold
    auto accumulate = [&](int idx, int smp, int mapq, bool fwd, bool in_depth, bool has_sub, int c, int q, bool disc, bool other_alt) {
      uint8_t meta = (uint8_t)(smp & 3);
      double pq = 0.0, la = 0.0, lb = 0.0;
#pragma unroll
      for (int s = 0; s < SJF_SMAX; ++s) {
        depth[s] += in_depth && smp == s;
        ns_alt[s] += other_alt && smp == s;
      }
      if (!has_sub) mapq = 0;
      if (mapq != 0) {  // mapq 0 contributes exactly 0.0 (PileupStats.scala:79-80)
        pq = c_success_probability[q] * c_success_probability[mapq];
        const double2 t = __ldg(A.log_table + ((mapq << 8) | q));
        la = t.x;
        lb = t.y;
      }
      const bool is_ref = has_sub && c == cr;
      if (is_ref && mapq != 0 && pq <= 0.5 && ((kinds >> (2 * smp)) & 3u) != GUAC_SJ_NORMAL_DNA) low_p = true;
      if (mapq != 0) {  // (has_sub)
        meta |= (uint8_t)(4 | (c << 3));
        W.p[idx] = pq;
        W.A[idx] = la;
        W.B[idx] = lb;
      }
      if (is_ref) {
#pragma unroll
        for (int s = 0; s < SJF_SMAX; ++s)
          if (smp == s) {
            rcnt[s] += 1;
            rfwd[s] += fwd;
            rfirst[s] = min(rfirst[s], idx);
            if (mapq != 0) {
              rll[s] += 1;
              rA[s] += la;
              rB[s] += lb;
            }
          }
      }
      if (idx >= 0 && idx < n_span) W.meta[idx] = meta;
      // the other subsequences go to the list (ballot compaction: the same order every run)
      const unsigned om = __ballot_sync(FULL, has_sub && !is_ref);
      if (om) {
        const int pos = n_list + __popc(om & ((1u << lane) - 1u));
        if (has_sub && !is_ref && pos < SJF_LIST) {
          SJFListEntry e;
          e.rel = (uint16_t)idx;
          e.s = (uint8_t)smp;
          e.c = (uint8_t)c;
          e.flags = (uint8_t)((fwd ? 1 : 0) | (mapq != 0 ? 2 : 0) | (disc ? 4 : 0));
          e.q = (uint8_t)q;
          e.mapq = (uint8_t)mapq;
          e._pad = 0;
          W.list[pos] = e;
        }
        n_list += __popc(om);
      }
    };
    // single-operator reads inline; the others are listed and walked by the exact cursor afterwards, all at once
    // the header, mapq and flags of a read depend on its index alone: those of the next chunk are requested while this
    // one is worked on
    uint4 next_h = make_uint4(0, 0, 0, 0);
    int next_mapq = 0, next_flags = 0;
    if (lane < n_span && !bail) {
      next_h = __ldg(A.hdr + tile.read_lo + lane);
      next_mapq = (int)__ldg(R.mapq + tile.read_lo + lane);
      next_flags = (int)__ldg(R.flags + tile.read_lo + lane);
    }
    for (int base = 0; base < n_span && !bail; base += 32) {
      const int idx = base + lane;
      const int64_t r = tile.read_lo + idx;
      bool in_depth = false, has_sub = false, disc = false, other_alt = false, slow = false;
      int smp = 0, c = 0, q = 0;
      const uint4 h = next_h;
      const int mapq = next_mapq;
      const bool fwd = (next_flags & GUAC_READ_POSITIVE_STRAND) != 0;
      if (idx + 32 < n_span) {
        next_h = __ldg(A.hdr + r + 32);
        next_mapq = (int)__ldg(R.mapq + r + 32);
        next_flags = (int)__ldg(R.flags + r + 32);
      }
      if (idx < n_span) {
        smp = (int)((h.y >> 16) & 0xFF);
        const uint8_t rf = (uint8_t)(h.y >> 24);
        const int st = (int)(h.x - x_lo);  // start - x
        int span = (int)(h.y & 0xFFFF);
        if (span == 0xFFFF) span = (int)min((long long)(__ldg(R.end + r) - __ldg(R.start + r)), (1ll << 30));
        const bool overlap = st <= 0 && st + span > 0 && smp < S;
        if (overlap && (rf & RF_FAST)) {
          // one aligned operator over standard bases: elements are Match / Mismatch, read position = locus - start
          const uint64_t s0 = (uint64_t)h.z | ((uint64_t)h.w << 32);
          const int p0 = -st;
          in_depth = true;
          auto base_at = [&](int rp) {
            const uint64_t gi = s0 + (uint64_t)rp;
            return (int)((__ldg(R.bases2 + (gi >> 2)) >> ((gi & 3) * 2)) & 3);
          };
          const bool e0_variant = "ACGT"[base_at(p0)] != (char)rb0;
          if (!e0_variant && p0 + 1 < span - 1) {  // ReadSubsequence.scala:71,114: the read goes on past x + 1
            c = base_at(p0 + 1);
            q = (int)__ldg(R.qual + s0 + p0 + 1);
            has_sub = true;
            if (c != cr) {  // ofNextAltAllele: a one-base allele when the run of variant elements ends at x + 2
              if (p0 + 2 < span - 1 && "ACGT"[base_at(p0 + 2)] != (char)rb2) other_alt = true;
              else disc = true;
            }
          }
        } else if (overlap) {
          slow = true;
        }
      }
      const unsigned sm = __ballot_sync(FULL, slow);
      if (sm) {
        const int pos = n_slow + __popc(sm & ((1u << lane) - 1u));
        if (slow && pos < SJF_LIST) W.slow[pos] = (uint16_t)idx;
        n_slow += __popc(sm);
      }
      accumulate(idx, smp, mapq, fwd, in_depth, has_sub, c, q, disc, other_alt);
    }
    SJF_PHASE(1);
    if (n_slow > SJF_LIST) bail = true;
    __syncwarp();
    for (int base = 0; base < n_slow && !bail; base += 32) {  // the exact cursor, one lane per listed read
      unsigned cls = 0;
      int smp = 0, idx = -1, mapq = 0;
      bool fwd = false;
      if (base + lane < n_slow) {
        idx = W.slow[base + lane];
        const int64_t r = tile.read_lo + idx;
        smp = (int)__ldg(R.sample_id + r);
        mapq = (int)__ldg(R.mapq + r);
        fwd = (__ldg(R.flags + r) & GUAC_READ_POSITIVE_STRAND) != 0;
        const uint4 h = __ldg(A.hdr + r);
        const uint8_t rf = (uint8_t)(h.y >> 24);
        if (!((rf & RF_MEDIUM) && sjf_classify_medium(R, r, h, x, rb0, rb2, cr, cls))) cls = sjf_classify_slow(R, r, x, ref, cr);
      }
      if (__any_sync(FULL, (cls & 16u) != 0)) bail = true;
      accumulate(idx, smp, mapq, fwd, (cls & 1u) != 0, (cls & 2u) != 0, (int)((cls >> 8) & 3u), (int)((cls >> 16) & 255u), (cls & 4u) != 0,
                 (cls & 8u) != 0);
    }
    if (n_list > SJF_LIST) bail = true;
    // alternate alleles of other shapes: harmless while no discovery sample has enough of them to keep one
    // (AlleleAtLocus.scala:99-124 needs at least minReads equal ones)
    if (!bail) {
#pragma unroll
      for (int s = 0; s < SJF_SMAX; ++s) {
        ns_alt[s] = __reduce_add_sync(FULL, ns_alt[s]);
        if (((disc_mask >> s) & 1u) && ns_alt[s] >= max(1, P.min_reads)) bail = true;
      }
    }
    low_p = __any_sync(FULL, low_p);
    if (bail) {  // not a simple site after all: the generic kernels take it
      if (lane == 0) {
        const int b = atomicAdd(A.n_bail, 1);
        if (b < A.max_bail) A.bail[b] = A.cands[ci];
      }
      continue;
    }
    __syncwarp();
    SJF_PHASE(2);
    // totals over the lanes (fixed trees)
#pragma unroll
    for (int s = 0; s < SJF_SMAX; ++s) {
      depth[s] = __reduce_add_sync(FULL, depth[s]);
      rcnt[s] = __reduce_add_sync(FULL, rcnt[s]);
      rfwd[s] = __reduce_add_sync(FULL, rfwd[s]);
      rll[s] = __reduce_add_sync(FULL, rll[s]);
      rfirst[s] = __reduce_min_sync(FULL, rfirst[s]);
      rA[s] = warp_sum_fixed(rA[s]);
      rB[s] = warp_sum_fixed(rB[s]);
    }
    // cells: lane 4 s + c holds the numbers of sample s's subsequences equal to base c.  The reference base's come from
    // the totals above; the others from the list, one cell at a time: every lane takes the listed reads k, k + 32, ... and
    // the per-lane numbers are combined by fixed trees (a heterozygous site lists half its reads, in four cells or so)
    int c_cnt = 0, c_fwd = 0, c_first = INT_MAX, c_ll = 0, c_disc = 0, c_dfirst = INT_MAX;
    double c_A = 0.0, c_B = 0.0;
    {
      const int s = (lane >> 2) & 3, c = lane & 3;
      if (lane < 16 && c == cr) {
        c_cnt = pick4(rcnt[0], rcnt[1], rcnt[2], rcnt[3], s);
        c_fwd = pick4(rfwd[0], rfwd[1], rfwd[2], rfwd[3], s);
        c_first = pick4(rfirst[0], rfirst[1], rfirst[2], rfirst[3], s);
        c_ll = pick4(rll[0], rll[1], rll[2], rll[3], s);
        c_A = pick4(rA[0], rA[1], rA[2], rA[3], s);
        c_B = pick4(rB[0], rB[1], rB[2], rB[3], s);
      }
      unsigned present = 0;
      for (int k = lane; k < n_list; k += 32) present |= 1u << (4 * (W.list[k].s & 3) + (W.list[k].c & 3));
      present = __reduce_or_sync(FULL, present);
      for (; present; present &= present - 1) {
        const int cell = __ffs(present) - 1, cs = cell >> 2, cc = cell & 3;
        int t_cnt = 0, t_fwd = 0, t_first = INT_MAX, t_ll = 0, t_disc = 0, t_dfirst = INT_MAX;
        double t_A = 0.0, t_B = 0.0;
        for (int k = lane; k < n_list; k += 32) {
          const SJFListEntry e = W.list[k];
          if (e.s != cs || e.c != cc) continue;
          t_cnt += 1;
          t_fwd += e.flags & 1;
          t_first = min(t_first, (int)e.rel);
          if (e.flags & 2) {
            t_ll += 1;
            t_A += W.A[e.rel];
            t_B += W.B[e.rel];
          }
          if (e.flags & 4) {
            t_disc += 1;
            t_dfirst = min(t_dfirst, (int)e.rel);
          }
        }
        t_cnt = __reduce_add_sync(FULL, t_cnt);
        t_fwd = __reduce_add_sync(FULL, t_fwd);
        t_first = __reduce_min_sync(FULL, t_first);
        t_ll = __reduce_add_sync(FULL, t_ll);
        t_disc = __reduce_add_sync(FULL, t_disc);
        t_dfirst = __reduce_min_sync(FULL, t_dfirst);
        t_A = warp_sum_fixed(t_A);
        t_B = warp_sum_fixed(t_B);
        if (lane == cell) {
          c_cnt = t_cnt;
          c_fwd = t_fwd;
          c_first = t_first;
          c_ll = t_ll;
          c_disc = t_disc;
          c_dfirst = t_dfirst;
          c_A = t_A;
          c_B = t_B;
        }
      }
    }
    SJF_PHASE(3);
    // AlleleAtLocus.variantAlleles (jointcaller/AlleleAtLocus.scala:85-149): per discovery sample, its alternate bases
    // by (count desc, first read asc), kept when count >= max(minReads, depth * percent / 100); union, distinct, capped
    int n_res = 0, res_c[4] = {-1, -1, -1, -1};
    for (int di = 0; di < P.n_discovery; ++di) {
      const int smp = (int)((disc_order >> (4 * di)) & 15u);
      const int dep = pick4(depth[0], depth[1], depth[2], depth[3], smp);
      const double required = fmax((double)P.min_reads, dep * P.min_percent / 100.0);
      int cnt[4], first[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        cnt[c] = __shfl_sync(FULL, c_disc, 4 * smp + c);
        first[c] = __shfl_sync(FULL, c_dfirst, 4 * smp + c);
      }
      for (int round = 0; round < 3; ++round) {
        int best = -1;
#pragma unroll
        for (int c = 0; c < 4; ++c)
          if (cnt[c] > 0 && (best < 0 || cnt[c] > cnt[best] || (cnt[c] == cnt[best] && first[c] < first[best]))) best = c;
        if (best < 0) break;
        if ((double)cnt[best] >= required) {
          bool dup = false;
          for (int k = 0; k < n_res; ++k) dup = dup || res_c[k] == best;
          if (!dup && n_res < P.max_alleles) res_c[n_res++] = best;
        }
        cnt[best] = 0;
      }
    }
    if (n_res == 0) continue;
    // per evidence slot (lane i < S + 2: sample i, pooled normal DNA, pooled tumor DNA): the member samples' cells
    const int my_kind = lane < S ? (int)((kinds >> (2 * lane)) & 3u) : lane == NP ? (int)GUAC_SJ_NORMAL_DNA : (int)GUAC_SJ_TUMOR_DNA;
    int g_depth = 0, g_cnt[4], g_fwd[4], g_first[4], g_ll[4];
    double g_A[4], g_B[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      g_cnt[c] = g_fwd[c] = g_ll[c] = 0;
      g_first[c] = INT_MAX;
      g_A[c] = g_B[c] = 0.0;
    }
    for (int s = 0; s < S; ++s) {
      const int sk = (int)((kinds >> (2 * s)) & 3u);
      const bool member = lane < S ? lane == s : (lane == NP ? sk == GUAC_SJ_NORMAL_DNA : sk == GUAC_SJ_TUMOR_DNA);
      if (member) g_depth += pick4(depth[0], depth[1], depth[2], depth[3], s);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int vc = __shfl_sync(FULL, c_cnt, 4 * s + c), vf = __shfl_sync(FULL, c_fwd, 4 * s + c);
        const int v1 = __shfl_sync(FULL, c_first, 4 * s + c), vl = __shfl_sync(FULL, c_ll, 4 * s + c);
        const double va = __shfl_sync(FULL, c_A, 4 * s + c), vb = __shfl_sync(FULL, c_B, 4 * s + c);
        if (member) {
          g_cnt[c] += vc;
          g_fwd[c] += vf;
          g_first[c] = min(g_first[c], v1);
          g_ll[c] += vl;
          g_A[c] += va;
          g_B[c] += vb;
        }
      }
    }
    const int n_sub = g_cnt[0] + g_cnt[1] + g_cnt[2] + g_cnt[3], f_sub = g_fwd[0] + g_fwd[1] + g_fwd[2] + g_fwd[3];
    // One allele: every slot's likelihoods, then the call logic across the slots
    // (MultiSampleSingleAlleleEvidence.scala:66-230), and the raw record for the posterior / final kernels.
    // Returns 1: called, 0: not called, 2: certainly not called and NO record written (only with allow_shortcut).
    auto eval_allele = [&](int a, bool allow_shortcut, int job0) -> int {
      const int ca = res_c[a];
      const int alt_cnt = pick4(g_cnt[0], g_cnt[1], g_cnt[2], g_cnt[3], ca);
      const int alt_fwd = pick4(g_fwd[0], g_fwd[1], g_fwd[2], g_fwd[3], ca);
      const int alt_ll = pick4(g_ll[0], g_ll[1], g_ll[2], g_ll[3], ca);
      // alt fraction of a tumor mixture (TumorDNASingleSampleSingleAlleleEvidence.scala:38)
      const double my_vaf = fmax(P.vaf_floor, (double)alt_cnt / (double)g_depth);
      // likelihoods of this lane's slot that need no per-read pass (PileupStats.logLikelihoodPileup :77-91 per mixture)
      double other_B = 0.0, nonref_B = 0.0, nonalt_B = 0.0;  // sums of log10(1 - p) over base classes
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        if (c != cr) nonref_B += g_B[c];
        if (c != ca) nonalt_B += g_B[c];
        if (c != cr && c != ca) other_B += g_B[c];
      }
      const double ref_A = pick4(g_A[0], g_A[1], g_A[2], g_A[3], cr);
      const double alt_A = pick4(g_A[0], g_A[1], g_A[2], g_A[3], ca);
      const int ref_ll = pick4(g_ll[0], g_ll[1], g_ll[2], g_ll[3], cr);
      const double ll0 = ref_A + nonref_B;                                   // {ref: 1}
      double ll1 = (double)(ref_ll + alt_ll) * LOG10_HALF + other_B;         // normal DNA: {ref: 0.5, alt: 0.5}
      const double ll2 = my_kind == GUAC_SJ_NORMAL_DNA ? alt_A + nonalt_B : 0.0;  // normal DNA: {alt: 1}
      // germline genotype from the pooled normal posteriors (maxBy keeps the first maximum)
      int germline = 0;
      {
        const double p0 = ll0, p1 = ll1 - EP.germline_neg_log10_het_prior, p2 = ll2 - EP.germline_neg_log10_hom_alt_prior;
        int g = 0;
        double best = p0;
        if (p1 > best) {
          best = p1;
          g = 1;
        }
        if (p2 > best) g = 2;
        germline = __shfl_sync(FULL, g, NP);
      }
      // alt fractions of the tumor mixtures: every sample's own, and the pooled one
      double v[SJF_SMAX];
#pragma unroll
      for (int s = 0; s < SJF_SMAX; ++s) v[s] = __shfl_sync(FULL, my_vaf, s);
      const double vp = __shfl_sync(FULL, my_vaf, TP);
      // Before the pass over all reads: a reference-base read with P(correct) > 1/2 can only lower
      // {alt: vaf, ref: 1 - vaf} against {ref: 1}, and reads on other bases add the same to both, so the alternate
      // reads' terms alone bound a tumor mixture's gain.  No germline call and no tumor mixture that can win against its
      // prior: the allele is not called (MultiSampleSingleAlleleEvidence.scala:120-123,220-230) and that pass is not
      // needed.  (With RNA samples the DNA prior depends on their posteriors: no shortcut.)  One lane per listed read.
      if (allow_shortcut && !low_p && !has_rna && germline == 0) {
        double gain[SJF_SMAX] = {0.0, 0.0, 0.0, 0.0}, gain_p = 0.0;
        for (int k = lane; k < n_list; k += 32) {
          const SJFListEntry& e = W.list[k];
          if (!(e.flags & 2) || (int)e.c != ca) continue;
          const int s = e.s, sk = (int)((kinds >> (2 * s)) & 3u);
          if (sk != GUAC_SJ_TUMOR_DNA) continue;
          const double ep = W.p[e.rel], eB = W.B[e.rel];
          const double vs = pick4(v[0], v[1], v[2], v[3], s);
          const double t = log10(vs * ep + (1.0 - vs) * (1.0 - ep)) - eB;
#pragma unroll
          for (int j = 0; j < SJF_SMAX; ++j)
            if (s == j) gain[j] += t;
          gain_p += log10(vp * ep + (1.0 - vp) * (1.0 - ep)) - eB;
        }
        double best_gain = warp_sum_fixed(gain_p);
#pragma unroll
        for (int j = 0; j < SJF_SMAX; ++j) best_gain = fmax(best_gain, warp_sum_fixed(gain[j]));
        if (best_gain - EP.somatic_neg_log10_variant_prior < -1e-6) return 2;
      }
      // {alt: vaf, ref: 1 - vaf} over every read of the tumor samples (PileupStats.logLikelihoodPileup :77-91), one lane per
      // read, per-lane sums combined by fixed butterflies
      double acc[SJF_SMAX] = {0.0, 0.0, 0.0, 0.0}, acc_p = 0.0;
      for (int i = lane; i < n_span; i += 32) {
        const uint8_t m = W.meta[i];
        if (!(m & 4)) continue;
        const int s = m & 3, c = (m >> 3) & 3, sk = (int)((kinds >> (2 * s)) & 3u);
        if (sk == GUAC_SJ_NORMAL_DNA) continue;
        const double pp = W.p[i];
        const double vs = pick4(v[0], v[1], v[2], v[3], s);
        double t, u = 0.0;
        if (c == cr) {  // mixtureFrequency = 1 - vaf
          t = log10((1.0 - vs) * pp + (1.0 - (1.0 - vs)) * (1.0 - pp));
          if (sk == GUAC_SJ_TUMOR_DNA) u = log10((1.0 - vp) * pp + (1.0 - (1.0 - vp)) * (1.0 - pp));
        } else if (c == ca) {  // = vaf
          t = log10(vs * pp + (1.0 - vs) * (1.0 - pp));
          if (sk == GUAC_SJ_TUMOR_DNA) u = log10(vp * pp + (1.0 - vp) * (1.0 - pp));
        } else {  // = 0
          t = u = W.B[i];
        }
#pragma unroll
        for (int k = 0; k < SJF_SMAX; ++k)
          if (s == k) acc[k] += t;
        if (sk == GUAC_SJ_TUMOR_DNA) acc_p += u;
      }
#pragma unroll
      for (int s = 0; s < SJF_SMAX; ++s) acc[s] = warp_sum_fixed(acc[s]);
      acc_p = warp_sum_fixed(acc_p);
      if (my_kind != GUAC_SJ_NORMAL_DNA) ll1 = lane == TP ? acc_p : pick4(acc[0], acc[1], acc[2], acc[3], lane);
      auto mixture_triggers = [&](double variant_prior) {
        const bool has_alt = my_vaf > 0, has_ref = 1.0 - my_vaf > 0;
        const double prior1 = has_ref && !has_alt ? 0.0 : has_ref && has_alt ? variant_prior : 1.7976931348623157e308;
        const int best = (ll1 - prior1) > ll0 ? 1 : 0;
        return best == 0 ? germline != 0 : germline != 1;
      };
      const bool rna = __any_sync(FULL, lane < S && my_kind == GUAC_SJ_TUMOR_RNA && mixture_triggers(EP.somatic_neg_log10_variant_prior_rna));
      const double dna_prior = rna ? EP.somatic_neg_log10_variant_prior_with_rna_evidence : EP.somatic_neg_log10_variant_prior;
      const bool trig = __any_sync(FULL, lane < S + 2 && lane != NP && my_kind == GUAC_SJ_TUMOR_DNA && mixture_triggers(dna_prior));
      const int is_call = (germline != 0 || trig) ? 1 : 0;
      if (lane >= S + 2) return is_call;
      SJEvidenceDev* E = reinterpret_cast<SJEvidenceDev*>(A.raw + (size_t)(job0 + a) * A.raw_stride);
      if (lane == 0) {
        E->site = -1;  // (the site record, if any, points here through first_job)
        E->allele = a;
      }
      SJSlotDev& o = E->slots[lane];
      o.total_depth = g_depth;
      o.n_subsequences = n_sub;
      o.forward_subsequences = f_sub;
      o.variant_count = alt_cnt;
      o.variant_forward = alt_fwd;
      o.ref_count = pick4(g_cnt[0], g_cnt[1], g_cnt[2], g_cnt[3], cr);
      o._pad = 0;
      o.alt_vaf = my_vaf;
      o.ll[0] = ll0;
      o.ll[1] = ll1;
      o.ll[2] = ll2;
      // truncatedAllelicDepths(maxAllelesPerSite + 1): by count desc, first supporting read asc
      int cnt[4] = {g_cnt[0], g_cnt[1], g_cnt[2], g_cnt[3]};
      int nd = 0;
      for (int round = 0; round < 4 && nd < P.max_alleles + 1; ++round) {
        int best = -1;
#pragma unroll
        for (int c = 0; c < 4; ++c)
          if (cnt[c] > 0 && (best < 0 || cnt[c] > cnt[best] || (cnt[c] == cnt[best] && g_first[c] < g_first[best]))) best = c;
        if (best < 0) break;
        SJDepthDev d;
        d.key = (1ull << 61) | (1ull << 54) | (uint64_t)best;
        d.rep_read = -1;
        d.rep_start_rp = (int)"ACGT"[best];
        d.len = 1;
        d.count = cnt[best];
        d.forward = pick4(g_fwd[0], g_fwd[1], g_fwd[2], g_fwd[3], best);
        o.depths[nd++] = d;
        cnt[best] = 0;
      }
      o.n_depths = nd;
      return is_call;
    };
    // The raw records of the site's alleles are written as they are evaluated; a site none of whose alleles is called
    // is never emitted (MultiSampleMultiAlleleEvidence.scala:149-155; force-called sites are not here) and ends without
    // a site record — its job slots stay unused.
    SJF_PHASE(4);
    int site_idx = 0, job0 = 0;
    if (lane == 0) job0 = atomicAdd(A.n_jobs, n_res);
    job0 = __shfl_sync(FULL, job0, 0);
    if (job0 + n_res > A.max_jobs) continue;  // the host sees the counter and reruns with room
    bool any_call = false;
    unsigned unwritten = 0;
    for (int a = 0; a < n_res; ++a) {
      const int r = eval_allele(a, true, job0);
      any_call = any_call || r == 1;
      if (r == 2) unwritten |= 1u << a;
    }
    if (!any_call) continue;
    for (int a = 0; a < n_res; ++a)  // the emitted site's other alleles need their records too
      if ((unwritten >> a) & 1u) eval_allele(a, false, job0);
    SJF_PHASE(5);
    if (lane == 0) site_idx = atomicAdd(A.n_sites, 1);
    site_idx = __shfl_sync(FULL, site_idx, 0);
    if (site_idx >= A.max_sites) continue;
    if (lane == 0) {
      SJSiteDev& st = A.sites[site_idx];
      st.slot = tile.slot0;
      st.locus = x;
      st.read_lo = tile.read_lo;
      st.read_hi = tile.read_hi;
      st.contig_id = tile.contig_id;
      st.ref_index = tile.ref_index;
      st.n_alleles = n_res;
      st.forced = 0;
      st.first_job = job0;
      st._pad = 0;
      for (int k = 0; k < SJ_MAX_ALLELES; ++k) {
        SJAlleleDev al;
        const int c = k < n_res ? res_c[k & 3] : 0;
        al.key = k < n_res ? ((1ull << 61) | (1ull << 54) | (uint64_t)c) : 0;  // seq_key of one standard base
        al.rep_read = -1;  // its string is the single character in rep_start_rp
        al.span = k < n_res ? 1 : 0;
        al.alt_len = k < n_res ? 1 : 0;
        al.rep_start_rp = (int)"ACGT"[c];
        al._pad = 0;
        st.alleles[k] = al;
      }
    }
  }
}

}  // namespace guac
