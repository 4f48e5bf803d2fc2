// joint_host.hpp — host half of the somatic-joint path: turns the raw per-(site, allele) numbers of
// sj_evidence_kernel into posteriors, calls, annotations and the emitted records.
//
// This is scalar logic on a handful of doubles per candidate site (the reference spends its time in
// the per-element work, which runs on the device); it follows
//   jointcaller/evidence/MultiSampleSingleAlleleEvidence.scala:24-342   (posteriors, calls, filters)
//   jointcaller/evidence/MultiSampleMultiAlleleEvidence.scala:131-177   (emit policy of `make`)
//   jointcaller/annotation/StrandBias.scala:22-58, InsufficientNormal.scala:21-53
//   filters/FishersExactTest.scala:12-19 over commons-math3 3.0 ArithmeticUtils.binomialCoefficientLog
// (paths relative to src/main/scala/org/hammerlab/guacamole/ of the reference).
#pragma once
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

#include "../../include/guac_b200.h"
#include "joint_kernels.cuh"

namespace guac {
namespace joint {

// commons-math3 3.0 binomialCoefficientLog: exact for n < 67, double product for n < 1030, log sums above
inline double binomial_coefficient_log(int n, int k) {
  if (n == k || k == 0) return 0;
  if (k == 1 || k == n - 1) return std::log((double)n);
  if (n < 67) {
    if (k > n / 2) k = n - k;
    unsigned long long c = 1;
    for (int j = 1, i = n - k + 1; j <= k; ++i, ++j) c = c * i / j;
    return std::log((double)c);
  }
  if (n < 1030) {
    const int kk = k > n / 2 ? n - k : k;
    double c = 1;
    for (int i = 1; i <= kk; ++i) c *= (double)(n - kk + i) / (double)i;
    return std::log(std::floor(c + 0.5));
  }
  if (k > n / 2) return binomial_coefficient_log(n, n - k);
  double s = 0;
  for (int i = n - k + 1; i <= n; ++i) s += std::log((double)i);
  for (int i = 2; i <= k; ++i) s -= std::log((double)i);
  return s;
}

inline double fishers_exact_log10(int total_a, int total_b, int cond_a, int cond_b) {
  return (binomial_coefficient_log(total_a, cond_a) + binomial_coefficient_log(total_b, cond_b) -
          binomial_coefficient_log(total_a + total_b, cond_a + cond_b)) / std::log(10.0);
}

struct Record {  // one MultiSampleSingleAlleleEvidence being assembled
  guac_sj_allele_evidence allele;
  std::vector<guac_sj_sample_evidence> samples;
  const SJEvidenceDev* raw;
  bool is_call() const { return allele.is_germline_call || allele.is_somatic_call; }
};

// MultiSampleSingleAlleleEvidence: posteriors, germline / somatic calls, annotations, failing filters
inline Record evaluate(const guac_sj_params& P, const SJSiteDev& site, int allele_index, const SJEvidenceDev& raw) {
  Record rec;
  rec.raw = &raw;
  const int S = P.n_samples, NP = S, TP = S + 1;
  std::memset(&rec.allele, 0, sizeof rec.allele);
  rec.allele.contig_id = site.contig_id;
  rec.allele.start = site.locus + 1;
  rec.allele.forced = site.forced;
  rec.samples.resize(S + 2);
  const bool ref_is_alt = false;  // an alt discovered at a site always differs from the reference sequence
  (void)ref_is_alt;
  (void)allele_index;
  for (int i = 0; i < S + 2; ++i) {
    guac_sj_sample_evidence& e = rec.samples[i];
    std::memset(&e, 0, sizeof e);
    const SJSlotDev& r = raw.slots[i];
    e.kind = i == NP ? GUAC_SJ_NORMAL_DNA : i == TP ? GUAC_SJ_TUMOR_DNA : P.sample_kind[i];
    e.total_depth = r.total_depth;
    e.n_depths = r.n_depths;
    for (int k = 0; k < r.n_depths; ++k) e.depths[k].count = r.depths[k].count;
    e.n_likelihoods = e.kind == GUAC_SJ_NORMAL_DNA ? 3 : 2;
    for (int k = 0; k < e.n_likelihoods; ++k) e.log_likelihoods[k] = r.ll[k];
    for (int k = 0; k < 3; ++k) e.posteriors[k] = std::numeric_limits<double>::quiet_NaN();
    e.alt_vaf = e.kind == GUAC_SJ_NORMAL_DNA ? 0.0 : r.alt_vaf;
    // StrandBias (annotation/StrandBias.scala:47-58)
    e.sb_variant_forward = r.variant_forward;
    e.sb_variant_reverse = r.variant_count - r.variant_forward;
    e.sb_total_forward = r.forward_subsequences;
    e.sb_total_reverse = r.n_subsequences - r.forward_subsequences;
    e.sb_phred = 10.0 * fishers_exact_log10(e.sb_total_forward, e.sb_total_reverse, e.sb_variant_forward, e.sb_variant_reverse);
    e.sb_filtered = e.sb_phred > P.filter_strand_bias_phred;
  }
  // germline posteriors: likelihood - prior (MultiSampleSingleAlleleEvidence.scala:66-123)
  const double germline_prior[3] = {0, P.germline_neg_log10_het_prior, P.germline_neg_log10_hom_alt_prior};
  for (int i = 0; i <= NP; ++i) {
    if (rec.samples[i].kind != GUAC_SJ_NORMAL_DNA) continue;
    for (int g = 0; g < 3; ++g) rec.samples[i].posteriors[g] = rec.samples[i].log_likelihoods[g] - germline_prior[g];
  }
  int germline = 0;  // maxBy keeps the first of equal maxima
  for (int g = 1; g < 3; ++g)
    if (rec.samples[NP].posteriors[g] > rec.samples[NP].posteriors[germline]) germline = g;
  rec.allele.germline_genotype = germline;
  rec.allele.is_germline_call = germline != 0;
  // A tumor mixture is {ref: 1} or {alt: vaf, ref: 1 - vaf}.  Its prior depends on which alleles have a
  // positive fraction (:172-183); whether it "triggers" depends on its key set vs the germline alleles (:220-224).
  auto somatic_prior = [&](int mixture, double vaf, double variant_prior) {
    if (mixture == 0) return 0.0;
    const bool has_alt = vaf > 0, has_ref = 1.0 - vaf > 0;
    if (has_ref && !has_alt) return 0.0;
    if (has_ref && has_alt) return variant_prior;
    return std::numeric_limits<double>::max();
  };
  auto triggers = [&](int best_mixture) {  // key set of the top mixture != {germline alleles}
    if (best_mixture == 0) return germline != 0;  // {ref} vs germline allele set
    return germline != 1;                         // {alt, ref} vs germline allele set
  };
  bool rna_expressed = false;
  for (int i = 0; i < S; ++i) {
    guac_sj_sample_evidence& e = rec.samples[i];
    if (e.kind != GUAC_SJ_TUMOR_RNA) continue;
    int best = 0;
    for (int m = 0; m < 2; ++m) {
      e.posteriors[m] = e.log_likelihoods[m] - somatic_prior(m, e.alt_vaf, P.somatic_neg_log10_variant_prior_rna);
      if (e.posteriors[m] > e.posteriors[best]) best = m;
    }
    e.triggered = triggers(best);
    rna_expressed = rna_expressed || e.triggered;
  }
  rec.allele.tumor_rna_expressed = rna_expressed;
  const double dna_prior = rna_expressed ? P.somatic_neg_log10_variant_prior_with_rna_evidence : P.somatic_neg_log10_variant_prior;
  bool any_triggered = false;
  for (int i = 0; i <= TP; ++i) {
    guac_sj_sample_evidence& e = rec.samples[i];
    if (i == NP || e.kind != GUAC_SJ_TUMOR_DNA) continue;
    int best = 0;
    for (int m = 0; m < 2; ++m) {
      e.posteriors[m] = e.log_likelihoods[m] - somatic_prior(m, e.alt_vaf, dna_prior);
      if (e.posteriors[m] > e.posteriors[best]) best = m;
    }
    e.triggered = triggers(best);
    any_triggered = any_triggered || e.triggered;
  }
  rec.allele.is_somatic_call = !rec.allele.is_germline_call && any_triggered;
  // InsufficientNormal (annotation/InsufficientNormal.scala:21-53): only for non-germline alleles
  if (!rec.allele.is_germline_call) {
    rec.allele.has_insufficient_normal = 1;
    rec.allele.in_reference_reads = raw.slots[NP].ref_count;
    rec.allele.in_total_reads = raw.slots[NP].total_depth;
    rec.allele.in_filtered = rec.allele.in_total_reads < P.filter_somatic_normal_depth ||
                             (rec.allele.in_total_reads - rec.allele.in_reference_reads) * 100.0 / rec.allele.in_total_reads >
                                 P.filter_somatic_normal_nonreference_percent;
  }
  // failingFilterNames (:250-275)
  bool have_trigger = false, all_fail = true;
  if (rec.allele.is_germline_call) {
    have_trigger = true;
    all_fail = rec.samples[NP].sb_filtered;
  } else if (rec.allele.is_somatic_call) {
    for (int i = 0; i <= TP; ++i)
      if (i != NP && rec.samples[i].kind == GUAC_SJ_TUMOR_DNA && rec.samples[i].triggered) {
        have_trigger = true;
        all_fail = all_fail && rec.samples[i].sb_filtered;
      }
  }
  if (have_trigger && all_fail) rec.allele.failing_filters |= 1u;
  if (rec.allele.has_insufficient_normal && rec.allele.in_filtered) rec.allele.failing_filters |= 2u;
  return rec;
}

// MultiSampleMultiAlleleEvidence.make's emit policy (:131-177) for one site's evaluated alleles
inline std::vector<Record> emit(const guac_sj_params& P, bool forced, std::vector<Record>& all) {
  std::vector<Record> evidences;
  for (auto& r : all)
    if (r.is_call()) evidences.push_back(r);
  if (evidences.empty()) evidences.push_back(all[0]);
  if (!forced) {
    bool any_call = false, any_germline = false;
    for (auto& r : evidences) {
      any_call = any_call || r.is_call();
      any_germline = any_germline || r.allele.is_germline_call;
    }
    if (!any_call) return {};
    if (P.only_somatic && any_germline) return {};
  }
  if (forced || P.include_filtered) return evidences;
  std::vector<Record> passing;
  for (auto& r : evidences)
    if (r.allele.failing_filters == 0) passing.push_back(r);
  return passing;
}

}  // namespace joint
}  // namespace guac
