// oracle/guac_oracle.cpp — TEST INFRASTRUCTURE ONLY (see guac_oracle.hpp).
// C entry points so that tests / bench.py's cpu_baseline can drive the CPU restatement
// through ctypes.  Inputs use the same SoA shapes as include/guac_b200.h except that
// bases are plain ASCII (the oracle does not trust the product's 2-bit packer).
#include "guac_oracle.hpp"

#include <cstdlib>
#include <cstring>
#include <thread>

#include "../include/guac_b200.h"  // POD result structs only

using namespace guac_oracle;

extern "C" {

typedef struct {
  int64_t n_reads;
  int32_t n_samples;
  const int32_t* contig_id;
  const int64_t* start;
  const uint32_t* cigar_off;
  const uint32_t* cigar;
  const uint64_t* seq_off;
  const uint8_t* bases_ascii;
  const uint8_t* qual;
  const uint8_t* mapq;
  const uint8_t* flags;
  const uint8_t* sample_id;
} oracle_reads;

typedef struct {
  int32_t contig_id;
  int32_t _pad;
  int64_t offset;
  int64_t len;
  const uint8_t* bases;
} oracle_contig;
}

namespace {

thread_local std::string g_error;

std::vector<MappedRead> buildReads(const oracle_reads* b) {
  std::vector<MappedRead> reads((size_t)b->n_reads);
  for (int64_t i = 0; i < b->n_reads; ++i) {
    MappedRead& r = reads[i];
    r.index = i;
    r.contig = b->contig_id[i];
    r.start = b->start[i];
    for (uint32_t c = b->cigar_off[i]; c < b->cigar_off[i + 1]; ++c)
      r.cigarElements.push_back(CigarElement{(int)(b->cigar[c] >> 4), (uint8_t)(b->cigar[c] & 15)});
    uint64_t s0 = b->seq_off[i], s1 = b->seq_off[i + 1];
    r.sequence.assign((const char*)b->bases_ascii + s0, (size_t)(s1 - s0));
    r.baseQualities.assign(b->qual + s0, b->qual + s1);
    r.alignmentQuality = b->mapq[i];
    r.isPositiveStrand = b->flags[i] & GUAC_READ_POSITIVE_STRAND;
    r.isPaired = b->flags[i] & GUAC_READ_PAIRED;
    r.isDuplicate = b->flags[i] & GUAC_READ_DUPLICATE;
    r.failedVendorQualityChecks = b->flags[i] & GUAC_READ_FAILED_QC;
    r.sampleId = b->sample_id[i];
    r.finish();
  }
  return reads;
}

std::vector<ContigSequence> buildContigs(const oracle_contig* c, int n) {
  std::vector<ContigSequence> v;
  for (int i = 0; i < n; ++i) {
    ContigSequence s;
    s.contig_id = c[i].contig_id;
    s.offset = c[i].offset;
    s.len = c[i].len;
    s.bases = c[i].bases;
    v.push_back(s);
  }
  return v;
}

std::vector<Range> buildRanges(const guac_range* r, int n) {
  std::vector<Range> v;
  for (int i = 0; i < n; ++i) v.push_back(Range{r[i].contig_id, r[i].start, r[i].end});
  return v;
}

template <class T>
T* toC(const std::vector<T>& v) {
  T* p = (T*)std::malloc(std::max<size_t>(1, v.size()) * sizeof(T));
  if (!v.empty()) std::memcpy(p, v.data(), v.size() * sizeof(T));
  return p;
}

guac_allele_evidence toC(const AlleleEvidence& e) {
  guac_allele_evidence c;
  c.probability = e.probability;
  c.read_depth = e.readDepth;
  c.allele_read_depth = e.alleleReadDepth;
  c.forward_depth = e.forwardDepth;
  c.allele_forward_depth = e.alleleForwardDepth;
  c.mean_mapping_quality = e.meanMappingQuality;
  c.median_mapping_quality = e.medianMappingQuality;
  c.mean_base_quality = e.meanBaseQuality;
  c.median_base_quality = e.medianBaseQuality;
  c.median_mismatches_per_read = e.medianMismatchesPerRead;
  return c;
}

guac_element toC(const Element& e) {
  guac_element g;
  std::memset(&g, 0, sizeof g);
  g.read_index = e.pe.read->index;
  g.read_position = e.pe.readPosition;
  g.cigar_element_index = e.pe.cigarElementIndex;
  g.cigar_element_locus = e.pe.cigarElementLocus;
  g.index_within_cigar_element = e.pe.indexWithinCigarElement;
  g.kind = e.aln.kind;
  g.base = (e.aln.kind == MATCH || e.aln.kind == MISMATCH) ? (uint8_t)e.aln.sequencedBases[0] : 0;
  g.quality = (uint8_t)e.qualityScore;
  g.sample_id = (uint8_t)e.pe.read->sampleId;
  g.indel_length = e.aln.kind == INSERTION  ? (int)e.aln.sequencedBases.size() - 1
                   : e.aln.kind == DELETION ? (int)e.aln.referenceBases.size() - 1
                                            : 0;
  return g;
}

}  // namespace

extern "C" {

const char* oracle_last_error() { return g_error.c_str(); }
void oracle_free(void* p) { std::free(p); }

// pileupFlatMapMultipleSamples with a counting closure.
int oracle_pileup_counts(const oracle_reads* b, const oracle_contig* contigs, int n_contigs, const guac_range* ranges,
                         int n_ranges, int skip_empty, int32_t** out_contig, int64_t** out_locus,
                         guac_locus_counts** out_counts, int64_t* n_out) {
  try {
    auto reads = buildReads(b);
    auto cs = buildContigs(contigs, n_contigs);
    std::vector<int32_t> oc;
    std::vector<int64_t> ol;
    std::vector<guac_locus_counts> cnt;
    pileupFlatMapMultipleSamples(reads, b->n_samples, buildRanges(ranges, n_ranges), cs, skip_empty != 0,
                                 [&](const std::vector<Pileup>& pileups) {
                                   oc.push_back(pileups[0].contig);
                                   ol.push_back(pileups[0].locus);
                                   for (auto& p : pileups) {
                                     guac_locus_counts c;
                                     std::memset(&c, 0, sizeof c);
                                     c.depth = p.depth();
                                     c.reference_depth = p.referenceDepth();
                                     c.positive_depth = p.positiveDepth();
                                     for (auto& e : p.elements) switch (e.aln.kind) {
                                         case MISMATCH: ++c.mismatches; break;
                                         case INSERTION: ++c.insertions; break;
                                         case DELETION: ++c.deletions; break;
                                         case MID_DELETION: ++c.mid_deletions; break;
                                         case CLIPPED: ++c.clipped; break;
                                         default: break;
                                       }
                                     cnt.push_back(c);
                                   }
                                 });
    *out_contig = toC(oc);
    *out_locus = toC(ol);
    *out_counts = toC(cnt);
    *n_out = (int64_t)ol.size();
    return 0;
  } catch (const std::exception& e) {
    g_error = e.what();
    return -1;
  }
}

// Same operator, materialising the elements.  Elements of a locus are reported sorted by
// read index (the CUDA path's canonical order); the reference's own order (priority-queue
// array order, then stable filter + append) is returned in out_ref_order as the read
// indices in reference order, same offsets.
int oracle_pileup_elements(const oracle_reads* b, const oracle_contig* contigs, int n_contigs, const guac_range* ranges,
                           int n_ranges, int skip_empty, int32_t** out_contig, int64_t** out_locus,
                           int64_t** out_offsets, guac_element** out_elements, int64_t** out_ref_order, int64_t* n_out) {
  try {
    auto reads = buildReads(b);
    auto cs = buildContigs(contigs, n_contigs);
    std::vector<int32_t> oc;
    std::vector<int64_t> ol, off{0}, order;
    std::vector<guac_element> els;
    pileupFlatMapMultipleSamples(reads, b->n_samples, buildRanges(ranges, n_ranges), cs, skip_empty != 0,
                                 [&](const std::vector<Pileup>& pileups) {
                                   oc.push_back(pileups[0].contig);
                                   ol.push_back(pileups[0].locus);
                                   std::vector<guac_element> here;
                                   for (auto& p : pileups)
                                     for (auto& e : p.elements) {
                                       here.push_back(toC(e));
                                       order.push_back(e.pe.read->index);
                                     }
                                   std::stable_sort(here.begin(), here.end(), [](const guac_element& x, const guac_element& y) {
                                     return x.read_index < y.read_index;
                                   });
                                   els.insert(els.end(), here.begin(), here.end());
                                   off.push_back((int64_t)els.size());
                                 });
    *out_contig = toC(oc);
    *out_locus = toC(ol);
    *out_offsets = toC(off);
    *out_elements = toC(els);
    *out_ref_order = toC(order);
    *n_out = (int64_t)ol.size();
    return 0;
  } catch (const std::exception& e) {
    g_error = e.what();
    return -1;
  }
}

// Allele strings (ref, alt) of every element at one locus of one sample, '\t' separated
// pairs joined by '\n', in reference element order.  For PileupSuite-style vectors.
int oracle_pileup_alleles(const oracle_reads* b, const oracle_contig* contigs, int n_contigs, int32_t contig, int64_t locus,
                          int sample, char** out_text) {
  try {
    auto reads = buildReads(b);
    auto cs = buildContigs(contigs, n_contigs);
    const ContigSequence* c = nullptr;
    for (auto& x : cs)
      if (x.contig_id == contig) c = &x;
    if (!c) throw InvalidInput("contig not found");
    std::string text;
    for (auto& r : reads)
      if (r.contig == contig && r.sampleId == sample && r.overlapsLocus(locus)) {
        Element e = makeElement(PileupElement::make(r, locus, *c));
        text += e.allele.refBases + "\t" + e.allele.altBases + "\t" + std::to_string((int)e.aln.kind) + "\t" +
                std::to_string(e.qualityScore) + "\n";
      }
    *out_text = (char*)std::malloc(text.size() + 1);
    std::memcpy(*out_text, text.c_str(), text.size() + 1);
    return 0;
  } catch (const std::exception& e) {
    g_error = e.what();
    return -1;
  }
}

// Likelihood.likelihoodsOfGenotypes for the pileup of one sample at one locus.  Genotype g
// has n_alleles[g] alleles; allele strings are given as NUL-separated "ref\0alt\0" pairs in
// `allele_text`, fractions in `fractions`, both flattened in genotype order.
int oracle_likelihoods_of_genotypes(const oracle_reads* b, const oracle_contig* contigs, int n_contigs, int32_t contig,
                                    int64_t locus, int sample, int n_genotypes, const int32_t* n_alleles,
                                    const char* allele_text, const double* fractions, int include_alignment, int log_space,
                                    int normalize, double* out) {
  try {
    auto reads = buildReads(b);
    auto cs = buildContigs(contigs, n_contigs);
    const ContigSequence* c = nullptr;
    for (auto& x : cs)
      if (x.contig_id == contig) c = &x;
    if (!c) throw InvalidInput("contig not found");
    std::vector<const MappedRead*> rs;
    for (auto& r : reads)
      if (r.contig == contig && r.sampleId == sample && r.overlapsLocus(locus)) rs.push_back(&r);
    Pileup p = Pileup::make(rs, sample, contig, locus, *c);
    std::vector<Genotype> gs;
    const char* t = allele_text;
    int k = 0;
    for (int g = 0; g < n_genotypes; ++g) {
      Genotype G;
      for (int a = 0; a < n_alleles[g]; ++a) {
        std::string ref(t);
        t += ref.size() + 1;
        std::string alt(t);
        t += alt.size() + 1;
        G.alleleMixture.push_back({Allele{ref, alt}, fractions[k++]});
      }
      gs.push_back(G);
    }
    auto l = likelihoodsOfGenotypes(p.elements, gs, include_alignment != 0, log_space != 0, normalize != 0);
    for (size_t i = 0; i < l.size(); ++i) out[i] = l[i];
    return 0;
  } catch (const std::exception& e) {
    g_error = e.what();
    return -1;
  }
}

// Likelihood.probabilitiesOfAllPossibleGenotypesFromPileup; result text lines
// "ref1\talt1\tref2\talt2\tvalue" (homozygous genotypes repeat the allele).
int oracle_all_genotypes(const oracle_reads* b, const oracle_contig* contigs, int n_contigs, int32_t contig, int64_t locus,
                         int sample, int include_alignment, int log_space, char** out_text) {
  try {
    auto reads = buildReads(b);
    auto cs = buildContigs(contigs, n_contigs);
    const ContigSequence* c = nullptr;
    for (auto& x : cs)
      if (x.contig_id == contig) c = &x;
    if (!c) throw InvalidInput("contig not found");
    std::vector<const MappedRead*> rs;
    for (auto& r : reads)
      if (r.contig == contig && r.sampleId == sample && r.overlapsLocus(locus)) rs.push_back(&r);
    Pileup p = Pileup::make(rs, sample, contig, locus, *c);
    auto res = probabilitiesOfAllPossibleGenotypesFromPileup(p, include_alignment != 0, log_space != 0);
    std::string text;
    char buf[64];
    for (auto& gv : res) {
      auto& m = gv.first.alleleMixture;
      const Allele& a0 = m[0].first;
      const Allele& a1 = m.size() > 1 ? m[1].first : m[0].first;
      snprintf(buf, sizeof buf, "%.17g", gv.second);
      text += a0.refBases + "\t" + a0.altBases + "\t" + a1.refBases + "\t" + a1.altBases + "\t" + buf + "\n";
    }
    *out_text = (char*)std::malloc(text.size() + 1);
    std::memcpy(*out_text, text.c_str(), text.size() + 1);
    return 0;
  } catch (const std::exception& e) {
    g_error = e.what();
    return -1;
  }
}

namespace {
struct SSOut {
  std::vector<guac_ss_call> calls;
  std::string arena;
  std::vector<guac_locus_summary> summary;
  int64_t visited = 0;
};

void runSomaticStandard(const std::vector<MappedRead>& reads, const std::vector<ContigSequence>& cs,
                        const std::vector<Range>& ranges, const guac_ss_params* params, bool wantSummary, SSOut& out) {
  SomaticStandardParams p;
  p.normalOddsThreshold = params->normal_odds_threshold;
  p.tumorOddsThreshold = params->tumor_odds_threshold;
  p.maxReadDepth = params->max_read_depth;
  p.maxNormalAlternateReadDepth = params->max_normal_alternate_read_depth;
  p.minTumorVariantAlleleFrequency = params->min_tumor_variant_allele_frequency;
  int ns = std::max(params->normal_sample_id, params->tumor_sample_id) + 1;
  pileupFlatMapMultipleSamples(reads, ns, ranges, cs, true, [&](const std::vector<Pileup>& pileups) {
    const Pileup& normal = pileups[params->normal_sample_id];
    const Pileup& tumor = pileups[params->tumor_sample_id];
    CalledSomaticAllele c;
    LocusSummary ts, nsum;
    ++out.visited;
    bool called = findPotentialVariantAtLocus(tumor, normal, p, &c, &ts, &nsum);
    if (wantSummary) {
      auto sat = [](int v) { return (uint16_t)std::min(v, 65535); };
      out.summary.push_back(guac_locus_summary{sat(nsum.depth), sat(nsum.referenceDepth), sat(nsum.topAltCount), (uint16_t)nsum.flags});
      out.summary.push_back(guac_locus_summary{sat(ts.depth), sat(ts.referenceDepth), sat(ts.topAltCount), (uint16_t)ts.flags});
    }
    if (called) {
      guac_ss_call g;
      std::memset(&g, 0, sizeof g);
      g.contig_id = c.contig;
      g.start = c.start;
      g.ref_off = (uint32_t)out.arena.size();
      g.ref_len = (uint32_t)c.allele.refBases.size();
      out.arena += c.allele.refBases;
      g.alt_off = (uint32_t)out.arena.size();
      g.alt_len = (uint32_t)c.allele.altBases.size();
      out.arena += c.allele.altBases;
      g.somatic_log_odds = c.somaticLogOdds;
      g.tumor_ref_log_prob = c.tumorRefLogProb;
      g.tumor_alt_log_prob = c.tumorAltLogProb;
      g.normal_ref_log_prob = c.normalRefLogProb;
      g.normal_alt_log_prob = c.normalAltLogProb;
      g.tumor_variant_evidence = toC(c.tumorVariantEvidence);
      g.normal_reference_evidence = toC(c.normalReferenceEvidence);
      out.calls.push_back(g);
    }
  });
}
}  // namespace

// somatic-standard over one partition; n_threads > 1 splits the ranges into contiguous
// chunks run in parallel (mirrors Spark local[N] over LociPartitioning ranges) — results
// are concatenated in range order.
int oracle_somatic_standard(const oracle_reads* b, const oracle_contig* contigs, int n_contigs, const guac_ss_params* params,
                            const guac_range* ranges, int n_ranges, int n_threads, guac_ss_call** out_calls, int64_t* n_calls,
                            char** out_arena, int64_t* arena_len, guac_locus_summary** out_summary, int64_t* out_n_visited) {
  try {
    auto reads = buildReads(b);
    auto cs = buildContigs(contigs, n_contigs);
    auto all = buildRanges(ranges, n_ranges);
    // split every range into n_threads pieces (UniformPartitioner-style contiguous loci)
    std::vector<std::vector<Range>> parts(std::max(1, n_threads));
    if (n_threads <= 1) {
      parts[0] = all;
    } else {
      int64_t total = 0;
      for (auto& r : all) total += r.end - r.start;
      int64_t per = (total + n_threads - 1) / n_threads, acc = 0;
      for (auto& r : all) {
        int64_t s = r.start;
        while (s < r.end) {
          int t = (int)std::min<int64_t>(n_threads - 1, acc / per);
          int64_t room = (t + 1) * per - acc;
          int64_t e = std::min(r.end, s + room);
          parts[t].push_back(Range{r.contig, s, e});
          acc += e - s;
          s = e;
        }
      }
    }
    std::vector<SSOut> outs(parts.size());
    std::vector<std::string> errs(parts.size());
    std::vector<std::thread> th;
    for (size_t t = 0; t < parts.size(); ++t)
      th.emplace_back([&, t] {
        try {
          // each thread works on its own copy-free view: MappedRead::cachedCountOfMismatches is
          // the only mutable field; races there are benign (same value written).
          runSomaticStandard(reads, cs, parts[t], params, out_summary != nullptr, outs[t]);
        } catch (const std::exception& e) {
          errs[t] = e.what();
        }
      });
    for (auto& x : th) x.join();
    for (auto& e : errs)
      if (!e.empty()) throw InvalidInput(e);
    SSOut merged;
    for (auto& o : outs) {
      for (auto c : o.calls) {
        c.ref_off += (uint32_t)merged.arena.size();
        c.alt_off += (uint32_t)merged.arena.size();
        merged.calls.push_back(c);
      }
      merged.arena += o.arena;
      merged.summary.insert(merged.summary.end(), o.summary.begin(), o.summary.end());
      merged.visited += o.visited;
    }
    *out_calls = toC(merged.calls);
    *n_calls = (int64_t)merged.calls.size();
    *out_arena = (char*)std::malloc(merged.arena.size() + 1);
    std::memcpy(*out_arena, merged.arena.c_str(), merged.arena.size() + 1);
    *arena_len = (int64_t)merged.arena.size();
    if (out_summary) *out_summary = toC(merged.summary);
    if (out_n_visited) *out_n_visited = merged.visited;
    return 0;
  } catch (const std::exception& e) {
    g_error = e.what();
    return -1;
  }
}

int oracle_success_probability_to_phred(double p) { return successProbabilityToPhred(p); }
double oracle_phred_to_success_probability(int q) { return phredToSuccessProbability(q); }

}  // extern "C"

// ---------------------------------------------------------------------------------
// C twin of the synthetic-read generator (guacamole_b200/synth.py is the specification;
// arithmetic shared with the CUDA twin through csrc/synth_model.h).  Produces reads
// [read_lo, read_hi) of the batch into caller-allocated arrays, ASCII bases.
// ---------------------------------------------------------------------------------
#include "../guacamole_b200/csrc/synth_model.h"

extern "C" int oracle_synth_generate(int64_t contig_len, int32_t n_samples, const int32_t* depths, const uint8_t* tumor,
                                     const uint8_t* pattern, int32_t pattern_len, uint64_t seed, int64_t read_lo,
                                     int64_t read_hi, int64_t* start, uint32_t* cigar_off, uint32_t* cigar, uint8_t* bases_ascii,
                                     uint8_t* qual, uint8_t* mapq, uint8_t* flags, uint8_t* sample, int n_threads, int profile) {
  using namespace guac::synth;
  Params P;
  std::memset(&P, 0, sizeof P);
  P.profile = profile;
  P.seed = seed;
  P.contig_len = contig_len;
  P.n_samples = n_samples;
  P.pattern_len = pattern_len;
  int64_t total = 0;
  for (int s = 0; s < n_samples; ++s) {
    total += depths[s];
    P.is_tumor[s] = tumor[s];
  }
  for (int i = 0; i < pattern_len; ++i) P.pattern[i] = pattern[i];
  P.n_reads = total * contig_len / READ_LEN;
  const int64_t n = read_hi - read_lo;
  uint32_t off = 0;
  for (int64_t i = 0; i < n; ++i) {
    ReadHeader h = read_header(P, read_lo + i);
    cigar_off[i] = off;
    write_cigar(h, cigar + off);
    off += h.n_ops;
    start[i] = h.start;
    mapq[i] = h.mapq;
    flags[i] = h.flags;
    sample[i] = (uint8_t)h.sample;
  }
  cigar_off[n] = off;
  std::vector<std::thread> th;
  int T = std::max(1, n_threads);
  for (int t = 0; t < T; ++t)
    th.emplace_back([&, t] {
      for (int64_t i = t; i < n; i += T) {
        ReadHeader h = read_header(P, read_lo + i);
        for (int k = 0; k < READ_LEN; ++k) {
          int code;
          uint8_t q;
          base_and_qual(P, h, read_lo + i, k, code, q);
          bases_ascii[i * READ_LEN + k] = (uint8_t)"ACGT"[code];
          qual[i * READ_LEN + k] = q;
        }
      }
    });
  for (auto& x : th) x.join();
  return 0;
}

extern "C" void oracle_synth_reference(uint64_t seed, int64_t lo, int64_t hi, uint8_t* out) {
  using namespace guac::synth;
  for (int64_t i = lo; i < hi; ++i) out[i - lo] = (uint8_t)"ACGT"[rnd_k(stream_key(seed, S_REF), (uint64_t)i) & 3];
}

// ---------------------------------------------------------------------------------
// somatic-joint: SomaticJoint.makeCalls (M/commands/SomaticJointCaller.scala:143-203) over one partition.
// `loci` are the caller's loci (the function applies lociSetMinusOne itself); a site is force-called
// when pileup locus + 1 lies in `force_loci`.
// ---------------------------------------------------------------------------------
#include "guac_oracle_joint.hpp"

extern "C" int oracle_somatic_joint(const oracle_reads* b, const oracle_contig* contigs, int n_contigs, const guac_sj_params* params,
                                    const guac_range* loci, int n_loci, const guac_range* force_loci, int n_force,
                                    guac_sj_allele_evidence** out_alleles, int64_t* n_alleles, guac_sj_sample_evidence** out_samples,
                                    int64_t* n_samples_out, char** out_arena, int64_t* arena_len, int64_t* out_n_visited) {
  try {
    auto reads = buildReads(b);
    auto cs = buildContigs(contigs, n_contigs);
    JointParameters P;
    for (int s = 0; s < params->n_samples; ++s) P.sampleKind.push_back(params->sample_kind[s]);
    P.maxAllelesPerSite = params->max_alleles_per_site;
    P.anyAlleleMinSupportingReads = params->any_allele_min_supporting_reads;
    P.anyAlleleMinSupportingPercent = params->any_allele_min_supporting_percent;
    P.germlineNegativeLog10HeterozygousPrior = params->germline_neg_log10_het_prior;
    P.germlineNegativeLog10HomozygousAlternatePrior = params->germline_neg_log10_hom_alt_prior;
    P.somaticNegativeLog10VariantPrior = params->somatic_neg_log10_variant_prior;
    P.somaticNegativeLog10VariantPriorRna = params->somatic_neg_log10_variant_prior_rna;
    P.somaticNegativeLog10VariantPriorWithRnaEvidence = params->somatic_neg_log10_variant_prior_with_rna_evidence;
    P.somaticVafFloor = params->somatic_vaf_floor;
    P.filterStrandBiasPhred = params->filter_strand_bias_phred;
    P.filterSomaticNormalNonreferencePercent = params->filter_somatic_normal_nonreference_percent;
    P.filterSomaticNormalDepth = params->filter_somatic_normal_depth;
    P.onlySomatic = params->only_somatic != 0;
    P.includeFiltered = params->include_filtered != 0;
    // lociSetMinusOne (SomaticJointCaller.scala:143-152)
    std::vector<Range> shifted;
    for (int i = 0; i < n_loci; ++i) {
      Range r{loci[i].contig_id, std::max<int64_t>(0, loci[i].start - 1), loci[i].end - 1};
      if (r.end > r.start) shifted.push_back(r);
    }
    std::vector<guac_sj_allele_evidence> alleles;
    std::vector<guac_sj_sample_evidence> samples;
    std::string arena;
    int64_t visited = 0;
    int site = 0;
    auto intern = [&](const std::string& s, uint32_t& off, uint32_t& len) {
      off = (uint32_t)arena.size();
      len = (uint32_t)s.size();
      arena += s;
    };
    pileupFlatMapMultipleSamples(reads, params->n_samples, shifted, cs, true, [&](const std::vector<Pileup>& pileups) {
      ++visited;
      const int64_t locus = pileups[0].locus;
      bool forceCall = false;
      for (int i = 0; i < n_force; ++i)
        forceCall = forceCall || (force_loci[i].contig_id == pileups[0].contig && force_loci[i].start <= locus + 1 && locus + 1 < force_loci[i].end);
      auto evs = makeMultiAlleleEvidence(pileups, P, forceCall);
      if (evs.empty()) return;
      for (auto& e : evs) {
        guac_sj_allele_evidence a;
        std::memset(&a, 0, sizeof a);
        a.contig_id = e.allele.contig;
        a.site_index = site;
        a.start = e.allele.start;
        intern(e.allele.ref, a.ref_off, a.ref_len);
        intern(e.allele.alt, a.alt_off, a.alt_len);
        a.is_germline_call = e.isGermlineCall;
        a.is_somatic_call = e.isSomaticCall;
        a.germline_genotype = e.germlineGenotype;
        a.tumor_rna_expressed = e.tumorRnaExpressed;
        a.has_insufficient_normal = e.hasInsufficientNormal;
        a.in_reference_reads = e.inReferenceReads;
        a.in_total_reads = e.inTotalReads;
        a.in_filtered = e.inFiltered;
        a.failing_filters = e.failingFilters;
        a.forced = forceCall;
        a.first_sample_evidence = (int64_t)samples.size();
        for (auto& se : e.samples) {
          guac_sj_sample_evidence g;
          std::memset(&g, 0, sizeof g);
          g.kind = se.kind;
          g.total_depth = se.totalDepth;
          g.n_depths = (int)se.allelicDepths.size();
          for (int k = 0; k < g.n_depths; ++k) {
            intern(se.allelicDepths[k].allele, g.depths[k].seq_off, g.depths[k].seq_len);
            g.depths[k].count = se.allelicDepths[k].count;
          }
          g.n_likelihoods = (int)se.logLikelihoods.size();
          for (int k = 0; k < g.n_likelihoods; ++k) {
            g.log_likelihoods[k] = se.logLikelihoods[k];
            g.posteriors[k] = k < (int)se.posteriors.size() ? se.posteriors[k] : NAN;
          }
          g.alt_vaf = se.altVaf;
          g.sb_variant_forward = se.sbVariantForward;
          g.sb_variant_reverse = se.sbVariantReverse;
          g.sb_total_forward = se.sbTotalForward;
          g.sb_total_reverse = se.sbTotalReverse;
          g.sb_phred = se.sbPhred;
          g.sb_filtered = se.sbFiltered;
          g.triggered = se.triggered;
          samples.push_back(g);
        }
        alleles.push_back(a);
      }
      ++site;
    });
    *out_alleles = toC(alleles);
    *n_alleles = (int64_t)alleles.size();
    *out_samples = toC(samples);
    *n_samples_out = (int64_t)samples.size();
    *out_arena = (char*)std::malloc(arena.size() + 1);
    std::memcpy(*out_arena, arena.c_str(), arena.size() + 1);
    *arena_len = (int64_t)arena.size();
    if (out_n_visited) *out_n_visited = visited;
    return 0;
  } catch (const std::exception& e) {
    g_error = e.what();
    return -1;
  }
}

// ReadSubsequence / PileupStats probes for the unit vectors of ReadSubsequenceSuite and PileupStatsSuite:
// text lines "kind\tsequence\tstartLocus\tendLocus" per element of `sample` at `locus`
// (kind F = ofFixedReferenceLength(length), A = ofNextAltAllele; "-" when None).
extern "C" int oracle_read_subsequences(const oracle_reads* b, const oracle_contig* contigs, int n_contigs, int32_t contig,
                                        int64_t locus, int sample, int length, char** out_text) {
  try {
    auto reads = buildReads(b);
    auto cs = buildContigs(contigs, n_contigs);
    const ContigSequence* c = nullptr;
    for (auto& x : cs)
      if (x.contig_id == contig) c = &x;
    if (!c) throw InvalidInput("contig not found");
    std::string text;
    for (auto& r : reads)
      if (r.contig == contig && r.sampleId == sample && r.overlapsLocus(locus)) {
        PileupElement pe = PileupElement::make(r, locus, *c);
        ReadSubsequence s;
        if (ofFixedReferenceLength(pe, length, &s))
          text += "F\t" + s.sequence() + "\t" + std::to_string(s.startLocus) + "\t" + std::to_string(s.endLocus) + "\n";
        else
          text += "F\t-\t0\t0\n";
        if (ofNextAltAllele(pe, &s))
          text += "A\t" + s.sequence() + "\t" + std::to_string(s.startLocus) + "\t" + std::to_string(s.endLocus) + "\n";
        else
          text += "A\t-\t0\t0\n";
      }
    *out_text = (char*)std::malloc(text.size() + 1);
    std::memcpy(*out_text, text.c_str(), text.size() + 1);
    return 0;
  } catch (const std::exception& e) {
    g_error = e.what();
    return -1;
  }
}

// PileupStats over the elements of `sample` at `locus` with reference sequence `ref_seq`:
// allelic depths as "allele=count" lines, then "LL\t<value>" per requested mixture
// (mixtures: NUL-separated "allele\0" keys + fractions, n_keys[m] keys each).
extern "C" int oracle_pileup_stats(const oracle_reads* b, const oracle_contig* contigs, int n_contigs, int32_t contig, int64_t locus,
                                   int sample, const char* ref_seq, int n_mixtures, const int32_t* n_keys, const char* keys_text,
                                   const double* fractions, char** out_text) {
  try {
    auto reads = buildReads(b);
    auto cs = buildContigs(contigs, n_contigs);
    const ContigSequence* c = nullptr;
    for (auto& x : cs)
      if (x.contig_id == contig) c = &x;
    if (!c) throw InvalidInput("contig not found");
    std::vector<const MappedRead*> rs;
    for (auto& r : reads)
      if (r.contig == contig && r.sampleId == sample && r.overlapsLocus(locus)) rs.push_back(&r);
    Pileup p = Pileup::make(rs, sample, contig, locus, *c);
    std::vector<const Element*> els;
    for (auto& e : p.elements) els.push_back(&e);
    PileupStats st(els, ref_seq);
    std::string text = "total=" + std::to_string(st.totalDepthIncludingReadsContributingNoAlleles()) + "\n";
    for (auto& d : st.allelicDepthsSorted) text += d.allele + "=" + std::to_string(d.count) + "\n";
    const char* t = keys_text;
    int k = 0;
    char buf[64];
    for (int m = 0; m < n_mixtures; ++m) {
      AlleleMixture mix;
      for (int j = 0; j < n_keys[m]; ++j) {
        std::string key(t);
        t += key.size() + 1;
        mix.push_back({key, fractions[k++]});
      }
      snprintf(buf, sizeof buf, "%.17g", st.logLikelihoodPileup(mix));
      text += std::string("LL\t") + buf + "\n";
    }
    *out_text = (char*)std::malloc(text.size() + 1);
    std::memcpy(*out_text, text.c_str(), text.size() + 1);
    return 0;
  } catch (const std::exception& e) {
    g_error = e.what();
    return -1;
  }
}

// ---------------------------------------------------------------------------------
// germline toy caller: the shortcut branch of GermlineAssemblyCaller.discoverGermlineVariants
// (M/commands/GermlineAssemblyCaller.scala:103-153) with callPileupVariant (:233-261) and
// AssemblyUtils.isActiveRegion (M/assembly/AssemblyUtils.scala:28-39).
// A locus whose region is "active" would be assembled by the reference; it is reported in
// out_active and not called here.  respect_active_regions = 0 calls every candidate locus.
// ---------------------------------------------------------------------------------
extern "C" int oracle_germline_standard(const oracle_reads* b, const oracle_contig* contigs, int n_contigs, const guac_gs_params* params,
                                        const guac_range* loci, int n_loci, guac_called_allele** out_calls, int64_t* n_calls,
                                        char** out_arena, int64_t* arena_len, int64_t** out_active, int64_t* n_active,
                                        int64_t* out_n_visited) {
  try {
    auto all_reads = buildReads(b);
    std::vector<MappedRead> reads;
    for (auto& r : all_reads)
      if (r.sampleId == params->sample_id) {
        reads.push_back(r);
        reads.back().sampleId = 0;
      }
    auto cs = buildContigs(contigs, n_contigs);
    std::vector<guac_called_allele> calls;
    std::vector<int64_t> active;
    std::string arena;
    int64_t visited = 0;
    pileupFlatMapMultipleSamples(reads, 1, buildRanges(loci, n_loci), cs, true, [&](const std::vector<Pileup>& pileups) {
      const Pileup& pileup = pileups[0];
      ++visited;
      const int pileupAltReads = pileup.depth() - pileup.referenceDepth();  // :135
      if (pileup.elements.empty() || pileupAltReads < params->min_alt_reads) return;
      if (params->respect_active_regions) {  // isActiveRegion (AssemblyUtils.scala:28-39)
        int n = 0;
        for (auto& e : pileup.elements)
          n += e.pe.read->countOfMismatches(*pileup.contigSequence) > 1 || e.pe.read->cigarElements.size() > 1;
        if ((float)n / (float)pileup.elements.size() > params->min_area_vaf) {
          active.push_back(pileup.contig);
          active.push_back(pileup.locus);
          return;
        }
      }
      // callPileupVariant (:233-261)
      auto gp = probabilitiesOfAllPossibleGenotypesFromPileup(pileup, false, true);
      if (gp.empty()) return;
      size_t best = 0;
      for (size_t i = 1; i < gp.size(); ++i)
        if (gp[i].second > gp[best].second) best = i;  // maxBy: first of equal maxima
      const double probability = std::exp(gp[best].second);
      std::vector<Allele> seen;
      for (auto& af : gp[best].first.alleleMixture) {
        const Allele& a = af.first;
        if (!a.isVariant() || a.altBases.empty()) continue;
        if (std::find(seen.begin(), seen.end(), a) != seen.end()) continue;
        seen.push_back(a);
        AlleleEvidence ev = makeAlleleEvidence(probability, a, pileup);
        if (!(ev.phredScaledLikelihood() > params->min_phred_scaled_likelihood)) continue;  // :147
        guac_called_allele c;
        std::memset(&c, 0, sizeof c);
        c.contig_id = pileup.contig;
        c.start = pileup.locus;
        c.ref_off = (uint32_t)arena.size();
        c.ref_len = (uint32_t)a.refBases.size();
        arena += a.refBases;
        c.alt_off = (uint32_t)arena.size();
        c.alt_len = (uint32_t)a.altBases.size();
        arena += a.altBases;
        c.genotype_log_probability = gp[best].second;
        c.n_genotypes = (int32_t)gp.size();
        c.evidence = toC(ev);
        calls.push_back(c);
      }
    });
    *out_calls = toC(calls);
    *n_calls = (int64_t)calls.size();
    *out_arena = (char*)std::malloc(arena.size() + 1);
    std::memcpy(*out_arena, arena.c_str(), arena.size() + 1);
    *arena_len = (int64_t)arena.size();
    *out_active = toC(active);
    *n_active = (int64_t)active.size() / 2;
    if (out_n_visited) *out_n_visited = visited;
    return 0;
  } catch (const std::exception& e) {
    g_error = e.what();
    return -1;
  }
}

// ---------------------------------------------------------------------------------
// germline-assembly: De Bruijn graph, alignment and the caller closure
// ---------------------------------------------------------------------------------
#include "guac_oracle_assembly.hpp"

namespace {
std::vector<MappedRead> readsFromText(const char* seqs, int n, const uint8_t* quals /* may be null: Phred 31 */) {
  std::vector<MappedRead> reads((size_t)n);
  const char* t = seqs;
  size_t qoff = 0;
  for (int i = 0; i < n; ++i) {
    reads[i].sequence = t;
    t += reads[i].sequence.size() + 1;
    const size_t L = reads[i].sequence.size();
    if (quals) reads[i].baseQualities.assign(quals + qoff, quals + qoff + L);
    else reads[i].baseQualities.assign(L, 31);
    qoff += L;
    reads[i].index = i;
  }
  return reads;
}
char* dupText(const std::string& s) {
  char* p = (char*)std::malloc(s.size() + 1);
  std::memcpy(p, s.c_str(), s.size() + 1);
  return p;
}
}  // namespace

// De Bruijn graph probe.  op: 0 kmerCounts ("kmer=count" lines), 1 depthFirstSearch(source, sink) merged path sequences,
// 2 mergeForward(source) k-mers, 3 mergeBackward(source) k-mers, 4 mergeIndex of every k-mer of `source` ("kmer=merged:idx").
extern "C" int oracle_debruijn(const char* seqs, int n_reads, const uint8_t* quals, int kmer_size, int min_occurrence,
                               int min_mean_kmer_quality, int merge_nodes, int op, const char* source, const char* sink, int max_paths,
                               char** out_text) {
  try {
    auto reads = readsFromText(seqs, n_reads, quals);
    std::vector<const MappedRead*> rp;
    for (auto& r : reads) rp.push_back(&r);
    DeBruijnGraph g = buildGraph(rp, kmer_size, min_occurrence, min_mean_kmer_quality, merge_nodes != 0);
    std::string text;
    if (op == 0) {
      for (auto& kv : g.kmerCounts) text += kv.first + "=" + std::to_string(kv.second) + "\n";
    } else if (op == 1) {
      for (auto& p : g.depthFirstSearch(source, sink, 1, std::numeric_limits<int>::max(), max_paths))
        text += DeBruijnGraph::mergeOverlappingSequences(p, kmer_size) + "\n";
    } else if (op == 2 || op == 3) {
      for (auto& k : (op == 2 ? g.mergeForward(source) : g.mergeBackward(source))) text += k + "\n";
    } else if (op == 4) {
      std::string s(source);
      for (size_t i = 0; i + kmer_size <= s.size(); ++i) {
        auto it = g.mergeIndex.find(s.substr(i, kmer_size));
        text += s.substr(i, kmer_size) + "=" + (it == g.mergeIndex.end() ? std::string("-") : it->second.first + ":" + std::to_string(it->second.second)) + "\n";
      }
    }
    *out_text = dupText(text);
    return 0;
  } catch (const std::exception& e) {
    g_error = e.what();
    return -1;
  }
}

// ReadAlignment(sequence, reference): "cigar\trefStartIdx\trefEndIdx\talignmentScore\trawScore"
extern "C" int oracle_read_alignment(const char* sequence, const char* reference, double mismatch_p, double open_gap_p, double close_gap_p,
                                     char** out_text) {
  try {
    ReadAlignmentResult r = mismatch_p > 0 ? readAlignment(sequence, reference, mismatch_p, open_gap_p, close_gap_p) : readAlignment(sequence, reference);
    char buf[64];
    snprintf(buf, sizeof buf, "%.17g", r.rawScore);
    *out_text = dupText(r.toCigarString() + "\t" + std::to_string(r.refStartIdx) + "\t" + std::to_string(r.refEndIdx) + "\t" +
                        std::to_string(r.alignmentScore) + "\t" + buf);
    return 0;
  } catch (const std::exception& e) {
    g_error = e.what();
    return -1;
  }
}

// GermlineAssemblyCaller.discoverGermlineVariants (M/commands/GermlineAssemblyCaller.scala:84-230) over one partition, one sample.
// Output text, one line per emitted CalledAllele in emission order (variants of a window sorted by (locus, ref, alt)):
//   "A\tcontig\tlocus\tref\talt\tdepth\tmeanMapq\tmedianMapq\tmeanBaseQ\tmedianBaseQ"   assembled
//   "P\tcontig\tlocus\tref\talt\tprobability\tdepth\talleleDepth"                         toy caller (shortcut)
//   "W\tcontig\tlocus\tnPaths"                                                             every assembled window
extern "C" int oracle_germline_assembly(const oracle_reads* b, const oracle_contig* contigs, int n_contigs, const guac_ga_params* P,
                                        const guac_range* loci, int n_loci, char** out_text) {
  try {
    auto all_reads = buildReads(b);
    std::vector<MappedRead> reads;
    for (auto& r : all_reads)
      if (r.sampleId == P->sample_id) reads.push_back(r);
    auto cs = buildContigs(contigs, n_contigs);
    auto ranges = buildRanges(loci, n_loci);
    std::string text;
    char buf[256];
    std::vector<int32_t> contigOrder;
    for (auto& r : ranges)
      if (std::find(contigOrder.begin(), contigOrder.end(), r.contig) == contigOrder.end()) contigOrder.push_back(r.contig);
    const int W = P->assembly_window_range, K = P->kmer_size;
    for (int32_t contig : contigOrder) {
      std::vector<std::pair<int64_t, int64_t>> rs;
      for (auto& r : ranges)
        if (r.contig == contig) rs.push_back({r.start, r.end});
      std::sort(rs.begin(), rs.end());
      const ContigSequence* c = nullptr;
      for (auto& x : cs)
        if (x.contig_id == contig) c = &x;
      std::vector<const MappedRead*> sampleReads;
      for (auto& r : reads)
        if (r.contig == contig) sampleReads.push_back(&r);
      if (!c) {
        if (sampleReads.empty()) continue;
        throw InvalidInput("contig not found in reference");
      }
      std::vector<SlidingWindow> windows(1);
      windows[0].halfWindowSize = W;
      windows[0].sortedRegions = &sampleReads;
      LociIterator it(rs);
      bool haveLast = false;
      int64_t lastCalledLocus = 0;
      while (advanceMultipleWindows(windows, it, true) >= 0) {
        SlidingWindow& window = windows[0];
        const int64_t locus = window.currentLocus;
        const int referenceStart = (int)(locus - W), referenceEnd = (int)(locus + W);
        auto regionReads = window.currentRegions();
        std::vector<const MappedRead*> currentLocusReads;
        for (auto* r : regionReads)
          if (r->overlapsLocus(locus)) currentLocusReads.push_back(r);
        if (currentLocusReads.empty()) continue;
        const std::string currentReference = c->slice(referenceStart, referenceEnd);  // :108-116 (evaluated before the early outs)
        Pileup pileup = Pileup::make(currentLocusReads, 0, contig, locus, *c);
        const int pileupAltReads = pileup.depth() - pileup.referenceDepth();
        if (pileupAltReads < P->min_alt_reads) continue;
        bool active = true;
        if (P->shortcut_assembly) {  // isActiveRegion (AssemblyUtils.scala:28-39)
          int n = 0;
          for (auto* r : currentLocusReads) n += r->countOfMismatches(*c) > 1 || r->cigarElements.size() > 1;
          active = (float)n / (float)currentLocusReads.size() > P->min_area_vaf;
        }
        if (!active) {  // callPileupVariant (:139-153)
          auto gp = probabilitiesOfAllPossibleGenotypesFromPileup(pileup, false, true);
          if (gp.empty()) continue;
          size_t best = 0;
          for (size_t i = 1; i < gp.size(); ++i)
            if (gp[i].second > gp[best].second) best = i;
          const double probability = std::exp(gp[best].second);
          std::vector<Allele> seen;
          bool any = false;
          for (auto& af : gp[best].first.alleleMixture) {
            const Allele& a = af.first;
            if (!a.isVariant() || a.altBases.empty() || std::find(seen.begin(), seen.end(), a) != seen.end()) continue;
            seen.push_back(a);
            AlleleEvidence ev = makeAlleleEvidence(probability, a, pileup);
            if (!(ev.phredScaledLikelihood() > P->min_phred_scaled_likelihood)) continue;
            snprintf(buf, sizeof buf, "P\t%d\t%lld\t%s\t%s\t%.17g\t%d\t%d\n", contig, (long long)locus, a.refBases.c_str(), a.altBases.c_str(),
                     probability, ev.readDepth, ev.alleleReadDepth);
            text += buf;
            any = true;
          }
          if (any) {
            haveLast = true;
            lastCalledLocus = locus;
          }
          continue;
        }
        // discoverHaplotypes over the whole window's reads (AssemblyUtils.scala:50-114)
        auto paths = discoverHaplotypes(regionReads, currentReference, K, P->min_occurrence, P->min_mean_kmer_quality);
        snprintf(buf, sizeof buf, "W\t%d\t%lld\t%d\n", contig, (long long)locus, (int)paths.size());
        text += buf;
        if (paths.empty()) continue;
        std::set<AssemblyVariant> variants;
        for (auto& path : paths)
          for (auto& v : buildVariantsFromPath(path, referenceStart, *c, currentReference))
            if (!haveLast || lastCalledLocus < v.locus) variants.insert(v);  // :205-209
        // window-wide evidence (:162-192): Float means / medians over all reads' mapq and all bases' qualities
        std::vector<double> mq, bq;
        for (auto* r : regionReads) {
          mq.push_back((float)r->alignmentQuality);
          for (uint8_t q : r->baseQualities) bq.push_back((float)q);
        }
        auto meanF = [](const std::vector<double>& v) {
          float mu = 0;
          long n = 0;
          for (double y : v) {
            n += 1;
            mu = mu + 1.0f / n * ((float)y - mu);
          }
          return (double)mu;
        };
        auto medianF = [](std::vector<double> v) {
          std::sort(v.begin(), v.end());
          const size_t n = v.size();
          return n % 2 ? v[(n - 1) / 2] : (double)(((float)v[n / 2 - 1] + (float)v[n / 2]) / 2);
        };
        for (auto& v : variants) {
          snprintf(buf, sizeof buf, "A\t%d\t%lld\t%s\t%s\t%d\t%.9g\t%.9g\t%.9g\t%.9g\n", contig, (long long)v.locus, v.ref.c_str(), v.alt.c_str(),
                   (int)regionReads.size(), meanF(mq), medianF(mq), meanF(bq), medianF(bq));
          text += buf;
          if (!haveLast || v.locus > lastCalledLocus) {
            haveLast = true;
            lastCalledLocus = std::max(lastCalledLocus, v.locus);
          }
        }
        window.setCurrentLocus(window.currentLocus + W - K);  // :213 (throws when W < K: "window can only move forward")
      }
    }
    *out_text = dupText(text);
    return 0;
  } catch (const std::exception& e) {
    g_error = e.what();
    return -1;
  }
}
