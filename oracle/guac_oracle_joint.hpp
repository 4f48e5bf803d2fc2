// oracle/guac_oracle_joint.hpp — TEST INFRASTRUCTURE ONLY (see guac_oracle.hpp).
// CPU restatement of the somatic-joint per-locus model (SURVEY.md §8 rows A8–A15).
// "M/" = src/main/scala/org/hammerlab/guacamole/ of the reference checkout.
// Pinned by the data row of tiny-sjc-output/all.vcf (SomaticJointCallerEndToEndSuite) and by the
// inline vectors of ReadSubsequenceSuite / PileupStatsSuite (tests/test_oracle_joint.py).
#pragma once
#include <set>

#include "guac_oracle.hpp"

namespace guac_oracle {

// ---------------------------------------------------------------------------------
// ReadSubsequence (M/jointcaller/pileup_summarization/ReadSubsequence.scala:18-147)
// ---------------------------------------------------------------------------------
struct ReadSubsequence {
  const MappedRead* read;
  int64_t startLocus, endLocus;
  int startReadPosition, endReadPosition;
  std::string sequence() const { return read->sequence.substr(startReadPosition, endReadPosition - startReadPosition); }  // :29
  bool sequenceIsAllStandardBases() const { return allStandardBases(sequence()); }                                         // :32
  double meanBaseQuality() const {                                                                                        // :36-50
    long sum = 0;
    int n = 0;
    if (startReadPosition == endReadPosition) {
      sum = read->baseQualities[startReadPosition];
      n = 1;
    } else {
      for (int i = startReadPosition; i < endReadPosition; ++i) {
        sum += read->baseQualities[i];
        ++n;
      }
    }
    return (double)sum / n;
  }
  std::string refSequence(const ContigSequence& c) const { return c.slice(startLocus, endLocus); }  // :53-55
};

inline bool elementIsVariant(const PileupElement& e) { return e.allele().isVariant(); }

// ReadSubsequence.ofFixedReferenceLength (:68-93)
inline bool ofFixedReferenceLength(const PileupElement& element, int length, ReadSubsequence* out) {
  if (elementIsVariant(element) || element.locus >= element.read->end - 1) return false;
  PileupElement firstElement = element.advanceToLocus(element.locus + 1);
  PileupElement currentElement = firstElement;
  int refOffset = 1;
  while (currentElement.locus < currentElement.read->end - 1 && refOffset < length) {
    currentElement = currentElement.advanceToLocus(currentElement.locus + 1);
    refOffset += 1;
  }
  if (currentElement.locus >= currentElement.read->end - 1) return false;
  *out = ReadSubsequence{element.read, firstElement.locus, currentElement.locus + 1, firstElement.readPosition,
                         currentElement.advanceToLocus(currentElement.locus + 1).readPosition};
  return true;
}

// ReadSubsequence.ofNextAltAllele (:106-135)
inline bool isVariantOrFollowedByDeletion(const PileupElement& e) {  // :108-112
  if (elementIsVariant(e)) return true;
  if (!e.isFinalCigarBase()) return false;
  const CigarElement* n = e.nextCigarElement();
  return n && !consumesReadBases(n->op) && consumesReferenceBases(n->op);
}
inline bool ofNextAltAllele(const PileupElement& element, ReadSubsequence* out) {
  if (isVariantOrFollowedByDeletion(element) || element.locus >= element.read->end - 1) return false;
  PileupElement firstElement = element.advanceToLocus(element.locus + 1);
  PileupElement currentElement = firstElement;
  while (currentElement.locus < currentElement.read->end - 1 && isVariantOrFollowedByDeletion(currentElement))
    currentElement = currentElement.advanceToLocus(currentElement.locus + 1);
  if (currentElement.locus == firstElement.locus || currentElement.locus == currentElement.read->end) return false;
  *out = ReadSubsequence{element.read, firstElement.locus, currentElement.locus, firstElement.readPosition, currentElement.readPosition};
  return true;
}

// ---------------------------------------------------------------------------------
// AlleleAtLocus (M/jointcaller/AlleleAtLocus.scala:22-149)
// ---------------------------------------------------------------------------------
struct AlleleAtLocus {
  int32_t contig;
  int64_t start;
  std::string ref, alt;
  int64_t end() const { return start + (int64_t)ref.size(); }
  bool operator==(const AlleleAtLocus& o) const { return contig == o.contig && start == o.start && ref == o.ref && alt == o.alt; }
};

// Ties between equally supported alleles: the reference's order comes out of Scala HashMap iteration
// (groupBy); oracle and CUDA path both break ties by the lowest first supporting read index.
inline std::vector<AlleleAtLocus> variantAlleles(const std::vector<const Pileup*>& pileups, int anyAlleleMinSupportingReads,
                                                 double anyAlleleMinSupportingPercent, int maxAlleles, bool atLeastOneAllele,
                                                 bool onlyStandardBases) {
  const ContigSequence& contigSequence = *pileups[0]->contigSequence;
  const int32_t contig = pileups[0]->contig;
  const int64_t variantStart = pileups[0]->locus + 1;
  struct Entry {
    AlleleAtLocus allele;
    double requiredReads;
    int actual;
    int64_t firstRead;
  };
  std::vector<Entry> all;
  for (const Pileup* pileup : pileups) {
    const double requiredReads = std::max<double>(anyAlleleMinSupportingReads, pileup->elements.size() * anyAlleleMinSupportingPercent / 100.0);
    struct G {
      int64_t endLocus;
      std::string seq;
      ReadSubsequence head;
      int n;
      int64_t firstRead;
    };
    std::vector<G> groups;
    for (auto& e : pileup->elements) {
      ReadSubsequence s;
      if (!ofNextAltAllele(e.pe, &s)) continue;
      if (onlyStandardBases && !s.sequenceIsAllStandardBases()) continue;
      std::string seq = s.sequence();
      bool found = false;
      for (auto& g : groups)
        if (g.endLocus == s.endLocus && g.seq == seq) {
          ++g.n;
          g.firstRead = std::min(g.firstRead, s.read->index);
          found = true;
        }
      if (!found) groups.push_back(G{s.endLocus, seq, s, 1, s.read->index});
    }
    std::stable_sort(groups.begin(), groups.end(), [](const G& a, const G& b) { return a.n != b.n ? a.n > b.n : a.firstRead < b.firstRead; });
    for (auto& g : groups)
      all.push_back(Entry{AlleleAtLocus{contig, variantStart, g.head.refSequence(contigSequence), g.seq}, requiredReads, g.n, g.firstRead});
  }
  std::vector<AlleleAtLocus> result;
  for (auto& t : all)
    if (t.actual >= t.requiredReads && std::find(result.begin(), result.end(), t.allele) == result.end()) result.push_back(t.allele);
  if (atLeastOneAllele && result.empty()) {
    struct T {
      AlleleAtLocus a;
      int total;
      int64_t firstRead;
    };
    std::vector<T> totals;
    for (auto& t : all) {
      bool found = false;
      for (auto& x : totals)
        if (x.a == t.allele) {
          x.total += t.actual;
          x.firstRead = std::min(x.firstRead, t.firstRead);
          found = true;
        }
      if (!found) totals.push_back(T{t.allele, t.actual, t.firstRead});
    }
    if (!totals.empty()) {
      const T* best = &totals[0];
      for (auto& x : totals)
        if (x.total > best->total || (x.total == best->total && x.firstRead < best->firstRead)) best = &x;
      return {best->a};
    }
    return {AlleleAtLocus{contig, variantStart, std::string(1, (char)contigSequence.apply(variantStart)), "N"}};
  }
  if (maxAlleles > 0 && (int)result.size() > maxAlleles) result.resize(maxAlleles);
  return result;
}

// ---------------------------------------------------------------------------------
// PileupStats (M/jointcaller/pileup_summarization/PileupStats.scala:23-91)
// ---------------------------------------------------------------------------------
using AlleleMixture = std::vector<std::pair<std::string, double>>;  // insertion ordered (scala Map1..Map4)

struct PileupStats {
  std::vector<const Element*> elements;
  std::string ref;
  std::vector<ReadSubsequence> subsequences;
  // allele -> depth in the canonical order: count desc, then first supporting read index
  struct Depth {
    std::string allele;
    int count;
    int64_t firstRead;
  };
  std::vector<Depth> allelicDepthsSorted;

  PileupStats(std::vector<const Element*> els, std::string refSequence) : elements(std::move(els)), ref(std::move(refSequence)) {
    if (ref.empty()) throw InvalidInput("assumption failed: referenceSequence.nonEmpty");
    for (auto* e : elements) {
      ReadSubsequence s;
      if (ofFixedReferenceLength(e->pe, (int)ref.size(), &s)) subsequences.push_back(s);  // :36
    }
    for (auto& s : subsequences) {  // :40-43
      std::string seq = s.sequence();
      bool found = false;
      for (auto& d : allelicDepthsSorted)
        if (d.allele == seq) {
          ++d.count;
          d.firstRead = std::min(d.firstRead, s.read->index);
          found = true;
        }
      if (!found) allelicDepthsSorted.push_back(Depth{seq, 1, s.read->index});
    }
    std::stable_sort(allelicDepthsSorted.begin(), allelicDepthsSorted.end(),
                     [](const Depth& a, const Depth& b) { return a.count != b.count ? a.count > b.count : a.firstRead < b.firstRead; });
  }
  int allelicDepth(const std::string& a) const {
    for (auto& d : allelicDepthsSorted)
      if (d.allele == a) return d.count;
    return 0;  // withDefaultValue(0)
  }
  std::vector<Depth> truncatedAllelicDepths(int max) const {  // :45-47
    std::vector<Depth> v = allelicDepthsSorted;
    if ((int)v.size() > max) v.resize(max);
    return v;
  }
  int totalDepthIncludingReadsContributingNoAlleles() const { return (int)elements.size(); }                        // :50
  double vaf(const std::string& allele) const { return (double)allelicDepth(allele) / totalDepthIncludingReadsContributingNoAlleles(); }  // :62
  double logLikelihoodPileup(const AlleleMixture& mixture) const {  // :77-91
    double total = 0;
    for (auto& s : subsequences) {
      double ll;
      if (s.read->alignmentQuality == 0) {
        ll = 0.0;
      } else {
        double mixtureFrequency = 0.0;
        std::string seq = s.sequence();
        for (auto& m : mixture)
          if (m.first == seq) mixtureFrequency = m.second;
        const int q = (int)std::floor(s.meanBaseQuality() + 0.5);  // .round.toInt = java Math.round
        const double probabilityCorrect = phredToSuccessProbability(q) * s.read->alignmentLikelihood();
        ll = std::log10(mixtureFrequency * probabilityCorrect + (1 - mixtureFrequency) * (1 - probabilityCorrect));
      }
      total += ll;
    }
    return total;
  }
};

// ---------------------------------------------------------------------------------
// commons-math3 3.0 ArithmeticUtils.binomialCoefficientLog (third party) and
// FishersExactTest.asLog10 (M/filters/FishersExactTest.scala:12-19)
// ---------------------------------------------------------------------------------
inline double binomialCoefficientLog(int n, int k) {
  if (n == k || k == 0) return 0;
  if (k == 1 || k == n - 1) return std::log((double)n);
  if (n < 67) {  // exact: commons-math divides by gcd(i, j) first for n > 61 so that no intermediate leaves 64 bits;
                 // a 128-bit intermediate gives the same integers
    if (k > n / 2) k = n - k;
    unsigned __int128 result = 1;
    for (int j = 1, i = n - k + 1; j <= k; ++i, ++j) result = result * (unsigned)i / (unsigned)j;
    return std::log((double)(unsigned long long)result);
  }
  if (n < 1030) {  // binomialCoefficientDouble
    int kk = k > n / 2 ? n - k : k;
    double result = 1;
    for (int i = 1; i <= kk; ++i) result *= (double)(n - kk + i) / (double)i;
    return std::log(std::floor(result + 0.5));
  }
  if (k > n / 2) return binomialCoefficientLog(n, n - k);
  double logSum = 0;
  for (int i = n - k + 1; i <= n; ++i) logSum += std::log((double)i);
  for (int i = 2; i <= k; ++i) logSum -= std::log((double)i);
  return logSum;
}
inline double fishersExactAsLog10(int totalA, int totalB, int conditionA, int conditionB) {
  return (binomialCoefficientLog(totalA, conditionA) + binomialCoefficientLog(totalB, conditionB) -
          binomialCoefficientLog(totalA + totalB, conditionA + conditionB)) / std::log(10.0);
}

// ---------------------------------------------------------------------------------
// Parameters (M/jointcaller/Parameters.scala:10-25), inputs (M/jointcaller/Input.scala)
// ---------------------------------------------------------------------------------
enum SampleKind { NORMAL_DNA = 0, TUMOR_DNA = 1, TUMOR_RNA = 2 };
struct JointParameters {
  std::vector<int> sampleKind;
  int maxAllelesPerSite = 4;
  int anyAlleleMinSupportingReads = 2;
  double anyAlleleMinSupportingPercent = 2.0;
  double germlineNegativeLog10HeterozygousPrior = 4, germlineNegativeLog10HomozygousAlternatePrior = 5;
  double somaticNegativeLog10VariantPrior = 6, somaticNegativeLog10VariantPriorRna = 1,
         somaticNegativeLog10VariantPriorWithRnaEvidence = 1;
  double somaticVafFloor = 0.05;
  double filterStrandBiasPhred = 60, filterSomaticNormalNonreferencePercent = 3.0;
  int filterSomaticNormalDepth = 30;
  bool onlySomatic = false, includeFiltered = false;
};

// One SingleSampleSingleAlleleEvidence + SingleSampleAnnotations
struct SampleEvidence {
  int kind;
  int totalDepth;
  std::vector<PileupStats::Depth> allelicDepths;  // truncated
  std::vector<double> logLikelihoods, posteriors;
  double altVaf = 0;
  int sbVariantForward = 0, sbVariantReverse = 0, sbTotalForward = 0, sbTotalReverse = 0;
  double sbPhred = 0;
  bool sbFiltered = false;
  bool triggered = false;
};

struct AlleleEvidenceJoint {
  AlleleAtLocus allele;
  std::vector<SampleEvidence> samples;  // per sample, then pooled normal, pooled tumor
  bool isGermlineCall = false, isSomaticCall = false;
  int germlineGenotype = 0;
  bool tumorRnaExpressed = false;
  bool hasInsufficientNormal = false;
  int inReferenceReads = 0, inTotalReads = 0;
  bool inFiltered = false;
  unsigned failingFilters = 0;
  bool isCall() const { return isGermlineCall || isSomaticCall; }
};

// MultiSampleSingleAlleleEvidence.apply + vals (M/jointcaller/evidence/MultiSampleSingleAlleleEvidence.scala:24-230,355-398),
// {NormalDNA,TumorDNA,TumorRNA}SingleSampleSingleAlleleEvidence.apply, annotate (:283-341), StrandBias, InsufficientNormal.
inline AlleleEvidenceJoint makeSingleAlleleEvidence(const JointParameters& P, const AlleleAtLocus& allele,
                                                    const std::vector<PileupStats>& sampleStats, const PileupStats& normalPooled,
                                                    const PileupStats& tumorPooled) {
  AlleleEvidenceJoint E;
  E.allele = allele;
  const int S = (int)sampleStats.size();
  auto normalEvidence = [&](const PileupStats& st) {  // NormalDNASingleSampleSingleAlleleEvidence.scala:42-52
    SampleEvidence ev;
    ev.kind = NORMAL_DNA;
    ev.totalDepth = st.totalDepthIncludingReadsContributingNoAlleles();
    ev.logLikelihoods = {st.logLikelihoodPileup({{allele.ref, 1.0}}),
                         st.logLikelihoodPileup(allele.ref == allele.alt ? AlleleMixture{{allele.ref, 1.0}}
                                                                         : AlleleMixture{{allele.ref, 0.5}, {allele.alt, 0.5}}),
                         st.logLikelihoodPileup({{allele.alt, 1.0}})};
    ev.allelicDepths = st.truncatedAllelicDepths(P.maxAllelesPerSite + 1);
    return ev;
  };
  auto tumorEvidence = [&](const PileupStats& st, int kind) {  // TumorDNA...scala:35-47 (RNA identical)
    SampleEvidence ev;
    ev.kind = kind;
    ev.totalDepth = st.totalDepthIncludingReadsContributingNoAlleles();
    if (allele.ref != st.ref) throw InvalidInput("assumption failed: " + allele.ref + " != " + st.ref);
    ev.altVaf = std::max(P.somaticVafFloor, st.vaf(allele.alt));
    ev.logLikelihoods = {st.logLikelihoodPileup({{allele.ref, 1.0}})};
    if (allele.ref != allele.alt) ev.logLikelihoods.push_back(st.logLikelihoodPileup({{allele.alt, ev.altVaf}, {allele.ref, 1.0 - ev.altVaf}}));
    ev.allelicDepths = st.truncatedAllelicDepths(P.maxAllelesPerSite + 1);
    return ev;
  };
  for (int s = 0; s < S; ++s) {
    if (P.sampleKind[s] == NORMAL_DNA) E.samples.push_back(normalEvidence(sampleStats[s]));
    else E.samples.push_back(tumorEvidence(sampleStats[s], P.sampleKind[s]));
  }
  E.samples.push_back(normalEvidence(normalPooled));
  E.samples.push_back(tumorEvidence(tumorPooled, TUMOR_DNA));
  const int normalDNAPooledIndex = S, tumorDNAPooledIndex = S + 1;

  // germline posteriors (:66-123)
  const double germlinePrior[3] = {0, P.germlineNegativeLog10HeterozygousPrior, P.germlineNegativeLog10HomozygousAlternatePrior};
  for (int i = 0; i <= normalDNAPooledIndex; ++i) {
    if (i < S && P.sampleKind[i] != NORMAL_DNA) continue;
    for (int g = 0; g < 3; ++g) E.samples[i].posteriors.push_back(E.samples[i].logLikelihoods[g] - germlinePrior[g]);
  }
  {
    auto& post = E.samples[normalDNAPooledIndex].posteriors;  // maxBy: first strictly greater
    int best = 0;
    for (int g = 1; g < 3; ++g)
      if (post[g] > post[best]) best = g;
    E.germlineGenotype = best;
    E.isGermlineCall = best != 0;
  }
  const std::set<std::string> germlineAlleleSet = E.germlineGenotype == 0   ? std::set<std::string>{allele.ref}
                                                  : E.germlineGenotype == 1 ? std::set<std::string>{allele.ref, allele.alt}
                                                                            : std::set<std::string>{allele.alt};
  // mixtures of a tumor evidence as key sets / priors
  auto mixtureContents = [&](int which, double altVaf) {  // keys with fraction > 0
    std::set<std::string> c;
    if (which == 0) c.insert(allele.ref);
    else {
      if (altVaf > 0) c.insert(allele.alt);
      if (1.0 - altVaf > 0) c.insert(allele.ref);
    }
    return c;
  };
  auto mixtureKeys = [&](int which) { return which == 0 ? std::set<std::string>{allele.ref} : std::set<std::string>{allele.alt, allele.ref}; };
  auto somaticPrior = [&](const std::set<std::string>& contents, double variantPrior) {
    if (contents == std::set<std::string>{allele.ref}) return 0.0;
    if (contents == std::set<std::string>{allele.ref, allele.alt}) return variantPrior;
    return std::numeric_limits<double>::max();
  };
  // RNA (:126-167)
  for (int i = 0; i < S; ++i)
    if (P.sampleKind[i] == TUMOR_RNA) {
      auto& ev = E.samples[i];
      int best = 0;
      for (size_t m = 0; m < ev.logLikelihoods.size(); ++m) {
        ev.posteriors.push_back(ev.logLikelihoods[m] - somaticPrior(mixtureContents((int)m, ev.altVaf), P.somaticNegativeLog10VariantPriorRna));
        if (ev.posteriors[m] > ev.posteriors[best]) best = (int)m;
      }
      ev.triggered = mixtureKeys(best) != germlineAlleleSet;
      E.tumorRnaExpressed = E.tumorRnaExpressed || ev.triggered;
    }
  // tumor DNA (:172-230)
  const double dnaPrior = E.tumorRnaExpressed ? P.somaticNegativeLog10VariantPriorWithRnaEvidence : P.somaticNegativeLog10VariantPrior;
  bool anyTriggered = false;
  for (int i = 0; i <= tumorDNAPooledIndex; ++i) {
    if (i == normalDNAPooledIndex) continue;
    if (i < S && P.sampleKind[i] != TUMOR_DNA) continue;
    auto& ev = E.samples[i];
    int best = 0;
    for (size_t m = 0; m < ev.logLikelihoods.size(); ++m) {
      ev.posteriors.push_back(ev.logLikelihoods[m] - somaticPrior(mixtureContents((int)m, ev.altVaf), dnaPrior));
      if (ev.posteriors[m] > ev.posteriors[best]) best = (int)m;
    }
    ev.triggered = mixtureKeys(best) != germlineAlleleSet;
    anyTriggered = anyTriggered || ev.triggered;
  }
  E.isSomaticCall = !E.isGermlineCall && anyTriggered;

  // annotations (:283-341): StrandBias per evidence, InsufficientNormal per allele
  auto strandBias = [&](SampleEvidence& ev, const PileupStats& st) {  // StrandBias.scala:22-58
    int varF = 0, varN = 0, totF = 0;
    for (auto& s : st.subsequences) {
      const bool fwd = s.read->isPositiveStrand;
      totF += fwd;
      if (s.sequence() == allele.alt) {
        ++varN;
        varF += fwd;
      }
    }
    ev.sbVariantForward = varF;
    ev.sbVariantReverse = varN - varF;
    ev.sbTotalForward = totF;
    ev.sbTotalReverse = (int)st.subsequences.size() - totF;
    ev.sbPhred = 10.0 * fishersExactAsLog10(ev.sbTotalForward, ev.sbTotalReverse, ev.sbVariantForward, ev.sbVariantReverse);
    ev.sbFiltered = ev.sbPhred > P.filterStrandBiasPhred;
  };
  for (int s = 0; s < S; ++s) strandBias(E.samples[s], sampleStats[s]);
  strandBias(E.samples[normalDNAPooledIndex], normalPooled);
  strandBias(E.samples[tumorDNAPooledIndex], tumorPooled);
  if (!E.isGermlineCall) {  // InsufficientNormal.scala:38-53,25-28
    E.hasInsufficientNormal = true;
    E.inReferenceReads = normalPooled.allelicDepth(allele.ref);
    E.inTotalReads = normalPooled.totalDepthIncludingReadsContributingNoAlleles();
    E.inFiltered = E.inTotalReads < P.filterSomaticNormalDepth ||
                   (E.inTotalReads - E.inReferenceReads) * 100.0 / E.inTotalReads > P.filterSomaticNormalNonreferencePercent;
  }
  // failingFilterNames (:250-275)
  std::vector<int> triggering;
  if (E.isGermlineCall) triggering = {normalDNAPooledIndex};
  else if (E.isSomaticCall)
    for (int i = 0; i <= tumorDNAPooledIndex; ++i)
      if (i != normalDNAPooledIndex && (i >= S || P.sampleKind[i] == TUMOR_DNA) && E.samples[i].triggered) triggering.push_back(i);
  bool allFail = true;
  for (int i : triggering) allFail = allFail && E.samples[i].sbFiltered;
  if (allFail && !triggering.empty()) E.failingFilters |= 1u;  // STRAND_BIAS
  if (E.hasInsufficientNormal && E.inFiltered) E.failingFilters |= 2u;  // INSUFFICIENT_NORMAL
  return E;
}

// MultiSampleMultiAlleleEvidence.make (M/jointcaller/evidence/MultiSampleMultiAlleleEvidence.scala:75-178)
inline std::vector<AlleleEvidenceJoint> makeMultiAlleleEvidence(const std::vector<Pileup>& pileups, const JointParameters& P, bool forceCall) {
  const int S = (int)pileups.size();
  // drop Clipped elements (:85-86)
  std::vector<Pileup> filtered(S);
  for (int s = 0; s < S; ++s) {
    filtered[s].sampleId = pileups[s].sampleId;
    filtered[s].contig = pileups[s].contig;
    filtered[s].locus = pileups[s].locus;
    filtered[s].contigSequence = pileups[s].contigSequence;
    for (auto& e : pileups[s].elements)
      if (e.aln.kind != CLIPPED) filtered[s].elements.push_back(e);
  }
  const ContigSequence& cs = *filtered[0].contigSequence;
  const int64_t locus = filtered[0].locus;
  if (!isStandardBase(cs.apply(locus + 1))) return {};  // :93
  std::vector<const Pileup*> ordered;  // normalDNA ++ tumorDNA (:97)
  for (int s = 0; s < S; ++s)
    if (P.sampleKind[s] == NORMAL_DNA) ordered.push_back(&filtered[s]);
  for (int s = 0; s < S; ++s)
    if (P.sampleKind[s] == TUMOR_DNA) ordered.push_back(&filtered[s]);
  if (ordered.empty() || P.sampleKind.empty()) throw InvalidInput("no DNA inputs");
  std::vector<AlleleAtLocus> possibleAlleles = variantAlleles(ordered, P.anyAlleleMinSupportingReads, P.anyAlleleMinSupportingPercent,
                                                               P.maxAllelesPerSite, forceCall, true);
  if (possibleAlleles.empty()) return {};
  std::vector<AlleleEvidenceJoint> all;
  for (auto& allele : possibleAlleles) {
    const std::string referenceSequence = cs.slice(allele.start, allele.end());
    std::vector<PileupStats> stats;
    std::vector<const Element*> normalPool, tumorPool;
    for (int s = 0; s < S; ++s) {
      std::vector<const Element*> els;
      for (auto& e : filtered[s].elements) els.push_back(&e);
      if (P.sampleKind[s] == NORMAL_DNA) normalPool.insert(normalPool.end(), els.begin(), els.end());
      if (P.sampleKind[s] == TUMOR_DNA) tumorPool.insert(tumorPool.end(), els.begin(), els.end());
      stats.emplace_back(els, referenceSequence);
    }
    PileupStats normalPooled(normalPool, referenceSequence), tumorPooled(tumorPool, referenceSequence);  // MultiplePileupStats.scala:12-18
    all.push_back(makeSingleAlleleEvidence(P, allele, stats, normalPooled, tumorPooled));
  }
  std::vector<AlleleEvidenceJoint> evidences;
  for (auto& e : all)
    if (e.isCall()) evidences.push_back(e);
  if (evidences.empty()) evidences.push_back(all[0]);  // :134
  if (!forceCall) {
    bool anyCall = false, anyGermline = false;
    for (auto& e : evidences) {
      anyCall = anyCall || e.isCall();
      anyGermline = anyGermline || e.isGermlineCall;
    }
    if (!anyCall) return {};
    if (P.onlySomatic && anyGermline) return {};
  }
  if (forceCall || P.includeFiltered) return evidences;
  std::vector<AlleleEvidenceJoint> passing;  // onlyPassingFilters (:53-55)
  for (auto& e : evidences)
    if (e.failingFilters == 0) passing.push_back(e);
  return passing;
}

}  // namespace guac_oracle
