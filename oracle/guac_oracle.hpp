// oracle/guac_oracle.hpp — TEST INFRASTRUCTURE ONLY.
//
// CPU restatement of hammerlab/guacamole's per-locus hot path, used as the parity
// checker for the CUDA library.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may use anything under oracle/.  The product
// (guacamole_b200/, libguac_b200.so) never includes, links or calls this.
//
// Each function cites the reference file:line it follows; "M/" abbreviates
// src/main/scala/org/hammerlab/guacamole/ of the reference checkout.
//
// Parity pins (see tests/test_oracle_*.py): tough.golden.vcf (somatic-standard end to
// end), PileupSuite / LikelihoodSuite / PileupFlatMapUtilsSuite / SlidingWindowSuite
// vectors transcribed into tests/golden/.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <limits>
#include <map>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace guac_oracle {

// ---------------------------------------------------------------------------------
// htsjdk 2.6.1 CigarOperator (third party, not under /root/reference): which operators
// consume read / reference bases; Cigar.getPaddedReferenceLength counts M D N = X P.
// ---------------------------------------------------------------------------------
enum CigarOp : uint8_t { OP_M = 0, OP_I = 1, OP_D = 2, OP_N = 3, OP_S = 4, OP_H = 5, OP_P = 6, OP_EQ = 7, OP_X = 8 };
inline bool consumesReadBases(uint8_t op) { return op == OP_M || op == OP_I || op == OP_S || op == OP_EQ || op == OP_X; }
inline bool consumesReferenceBases(uint8_t op) { return op == OP_M || op == OP_D || op == OP_N || op == OP_EQ || op == OP_X; }
inline char cigarOpChar(uint8_t op) { return "MIDNSHP=X"[op]; }

struct CigarElement {
  int length;
  uint8_t op;
};

struct InvalidInput : std::runtime_error {
  using std::runtime_error::runtime_error;
};

// ---------------------------------------------------------------------------------
// ADAM 0.20.3 org.bdgenomics.adam.util.PhredUtils (third party): success probability
// table 1 - 10^(-q/10); successProbabilityToPhred(p) = round(-10 log10(1 - p)).
// M/util/PhredUtils.scala:7-14 clamps p to 1 - 1e-16 first.
// ---------------------------------------------------------------------------------
inline double phredToSuccessProbability(int phred) { return 1.0 - std::pow(10.0, -phred / 10.0); }
inline int successProbabilityToPhred(double probability) {
  const double one = 1 - 1e-16;  // M/util/PhredUtils.scala:8
  double p = std::min(one, probability);
  return (int)std::llround(std::floor(-10.0 * std::log10(1.0 - p) + 0.5));  // java Math.round = floor(x + 0.5)
}

// ---------------------------------------------------------------------------------
// Reference contig (M/reference/ReferenceBroadcast.scala:42-83).  A window
// [offset, offset+len) of upper-cased ASCII; byte 0 = position missing (partial FASTA).
// ---------------------------------------------------------------------------------
struct ContigSequence {
  int32_t contig_id = 0;
  int64_t offset = 0;
  int64_t len = 0;
  const uint8_t* bases = nullptr;
  uint8_t apply(int64_t locus) const {
    if (locus < offset || locus >= offset + len || bases[locus - offset] == 0)
      throw InvalidInput("Position " + std::to_string(contig_id) + ":" + std::to_string(locus) + " missing from reference");
    return bases[locus - offset];
  }
  std::string slice(int64_t start, int64_t end) const {
    std::string s;
    for (int64_t i = start; i < end; ++i) s.push_back((char)apply(i));
    return s;
  }
};

// ---------------------------------------------------------------------------------
// MappedRead (M/reads/MappedRead.scala:21-81)
// ---------------------------------------------------------------------------------
struct MappedRead {
  int64_t index = 0;  // position in the uploaded batch (not in the reference; used for canonical tie-breaks)
  std::string sequence;
  std::vector<uint8_t> baseQualities;
  bool isDuplicate = false;
  int sampleId = 0;
  int32_t contig = 0;
  int alignmentQuality = 0;
  int64_t start = 0;
  std::vector<CigarElement> cigarElements;
  bool failedVendorQualityChecks = false;
  bool isPositiveStrand = true;
  bool isPaired = true;
  int64_t end = 0;  // start + cigar.getPaddedReferenceLength  (:75)
  mutable int cachedCountOfMismatches = -1;

  void finish() {
    int64_t len = 0;
    for (auto& e : cigarElements)
      if (consumesReferenceBases(e.op) || e.op == OP_P) len += e.length;
    end = start + len;
  }
  double alignmentLikelihood() const { return phredToSuccessProbability(alignmentQuality); }  // :66
  bool overlapsLocus(int64_t locus, int64_t halfWindowSize = 0) const {
    // genomic-loci Region.overlapsLocus: start - hw <= locus < end + hw
    return start - halfWindowSize <= locus && end + halfWindowSize > locus;
  }
  int countOfMismatches(const ContigSequence& contigSequence) const;  // :53-64
};

// ---------------------------------------------------------------------------------
// Allele (M/variants/Allele.scala:7-17), ordering via Bases.BasesOrdering
// (M/util/Bases.scala:27-31: String.compareTo of the base strings).
// ---------------------------------------------------------------------------------
struct Allele {
  std::string refBases, altBases;
  bool isVariant() const { return refBases != altBases; }
  bool operator==(const Allele& o) const { return refBases == o.refBases && altBases == o.altBases; }
  bool operator!=(const Allele& o) const { return !(*this == o); }
  bool operator<(const Allele& o) const {
    int c = refBases.compare(o.refBases);
    if (c != 0) return c < 0;
    return altBases.compare(o.altBases) < 0;
  }
};

inline bool isStandardBase(uint8_t b) { return b == 'A' || b == 'C' || b == 'T' || b == 'G'; }  // M/util/Bases.scala:49-51
inline bool allStandardBases(const std::string& s) {
  for (char c : s)
    if (!isStandardBase((uint8_t)c)) return false;
  return true;
}

// ---------------------------------------------------------------------------------
// Alignment ADT (M/pileup/Alignment.scala:14-75)
// ---------------------------------------------------------------------------------
enum AlignmentKind : uint8_t { MATCH = 0, MISMATCH = 1, INSERTION = 2, DELETION = 3, MID_DELETION = 4, CLIPPED = 5 };

struct Alignment {
  AlignmentKind kind = CLIPPED;
  std::string sequencedBases, referenceBases;
  std::vector<uint8_t> baseQualities;  // Insertion: all; Match/Mismatch/Deletion: one
};

// ---------------------------------------------------------------------------------
// PileupElement (M/pileup/PileupElement.scala:24-259)
// ---------------------------------------------------------------------------------
struct PileupElement {
  const MappedRead* read = nullptr;
  int64_t locus = 0;
  int readPosition = 0;
  int cigarElementIndex = 0;
  int64_t cigarElementLocus = 0;
  int indexWithinCigarElement = 0;
  const ContigSequence* contigSequence = nullptr;

  const CigarElement& cigarElement() const { return read->cigarElements[cigarElementIndex]; }
  const CigarElement* nextCigarElement() const {  // :39-44
    return cigarElementIndex + 1 < (int)read->cigarElements.size() ? &read->cigarElements[cigarElementIndex + 1] : nullptr;
  }
  uint8_t referenceBase() const { return contigSequence->apply(locus); }  // :36
  int cigarElementReferenceLength() const {                                 // :51, M/util/CigarUtils.scala:22
    return consumesReferenceBases(cigarElement().op) ? cigarElement().length : 0;
  }
  int64_t cigarElementEndLocus() const { return cigarElementLocus + cigarElementReferenceLength(); }  // :52
  bool isFinalCigarBase() const { return indexWithinCigarElement == cigarElement().length - 1; }       // :57

  // :59-117
  Alignment alignment() const {
    uint8_t op = cigarElement().op;
    const CigarElement* nextBase = isFinalCigarBase() ? nextCigarElement() : &cigarElement();
    int nextOp = nextBase ? (int)nextBase->op : -1;
    auto makeInsertion = [&](const CigarElement& e) {  // :64-74
      Alignment a;
      a.kind = INSERTION;
      int n = (consumesReadBases(e.op) ? e.length : 0) + 1;
      int until = std::min<int>(readPosition + n, (int)read->sequence.size());  // view(from, until) clamps
      a.sequencedBases = read->sequence.substr(readPosition, until - readPosition);
      a.baseQualities.assign(read->baseQualities.begin() + readPosition, read->baseQualities.begin() + until);
      a.referenceBases = a.sequencedBases.substr(0, 1);  // Alignment.scala:28
      return a;
    };
    Alignment a;
    if ((op == OP_M || op == OP_EQ) && nextOp == OP_I) return makeInsertion(*nextCigarElement());  // :82-83
    if (op == OP_I && nextOp >= 0 && cigarElementLocus == 0) return makeInsertion(cigarElement());  // :88-89
    if (op == OP_I)                                                                                  // :92
      throw InvalidInput("Should not have a PileupElement at non-reference-consuming cigar-operator I. Locus: " +
                         std::to_string(locus) + ", readPosition: " + std::to_string(readPosition));
    if ((op == OP_M || op == OP_EQ || op == OP_X) && nextOp == OP_D) {  // :94-97
      a.kind = DELETION;
      a.referenceBases = contigSequence->slice(locus, locus + nextCigarElement()->length + 1);
      a.sequencedBases = a.referenceBases.substr(0, 1);  // Alignment.scala:62
      a.baseQualities = {read->baseQualities[readPosition]};
      return a;
    }
    if (op == OP_D) {  // :98-99
      a.kind = MID_DELETION;
      a.referenceBases = std::string(1, (char)referenceBase());
      return a;
    }
    if (nextOp == OP_D)  // :100-104
      throw InvalidInput(std::string("Found deletion preceded by cigar operator ") + cigarOpChar(op) +
                         " at PileupElement at locus " + std::to_string(locus));
    if (op == OP_M || op == OP_EQ || op == OP_X) {  // :105-112
      uint8_t base = (uint8_t)read->sequence[readPosition];
      uint8_t ref = referenceBase();
      a.kind = base == ref ? MATCH : MISMATCH;
      a.sequencedBases = std::string(1, (char)base);
      a.referenceBases = std::string(1, (char)ref);
      a.baseQualities = {read->baseQualities[readPosition]};
      return a;
    }
    if (op == OP_S || op == OP_N || op == OP_H) {  // :113
      a.kind = CLIPPED;
      return a;
    }
    throw InvalidInput("`P` CIGAR-ops should have been ignored earlier in `findNextCigarElement`");  // :114-115
  }

  Allele allele() const {  // :140
    Alignment a = alignment();
    return Allele{a.referenceBases, a.sequencedBases};
  }

  static int qualityScoreOf(const Alignment& a, const MappedRead& r) {  // :149-154
    switch (a.kind) {
      case CLIPPED:
      case MID_DELETION: return r.alignmentQuality;
      case DELETION:
      case MATCH:
      case MISMATCH: return a.baseQualities[0];
      case INSERTION: return *std::min_element(a.baseQualities.begin(), a.baseQualities.end());
    }
    return 0;
  }
  int qualityScore() const { return qualityScoreOf(alignment(), *read); }
  double probabilityCorrectIgnoringAlignment() const { return phredToSuccessProbability(qualityScore()); }  // :156-157
  double probabilityCorrectIncludingAlignment() const {                                                      // :159-160
    return phredToSuccessProbability(qualityScore()) * read->alignmentLikelihood();
  }

  PileupElement advanceToNextCigarElement() const {  // :165-188
    int readPositionOffset = consumesReadBases(cigarElement().op) ? cigarElement().length - indexWithinCigarElement : 0;
    PileupElement e = *this;
    e.locus = locus + (cigarElementReferenceLength() - indexWithinCigarElement);
    e.readPosition = readPosition + readPositionOffset;
    e.cigarElementIndex = cigarElementIndex + 1;
    e.cigarElementLocus = cigarElementLocus + cigarElementReferenceLength();
    e.indexWithinCigarElement = 0;
    if (e.cigarElementIndex >= (int)read->cigarElements.size())
      throw InvalidInput("advanced past the last cigar element");
    return e;
  }
  bool currentCigarElementContainsLocus(int64_t referenceLocus) const {  // :195-197
    return cigarElementLocus <= referenceLocus && referenceLocus < cigarElementEndLocus();
  }
  PileupElement advanceToLocus(int64_t newLocus) const {  // :209-235 (tail recursion -> loop)
    PileupElement cur = *this;
    for (;;) {
      if (!(newLocus >= cur.locus))
        throw InvalidInput("Can't rewind to locus " + std::to_string(newLocus) + " from " + std::to_string(cur.locus) +
                           ". Pileups only advance.");
      if (!(newLocus < cur.read->end))
        throw InvalidInput("This read stops at position " + std::to_string(cur.read->end) + ". Can't advance to " +
                           std::to_string(newLocus));
      if (cur.currentCigarElementContainsLocus(newLocus)) {
        int readPositionOffset =
            consumesReadBases(cur.cigarElement().op) ? (int)(newLocus - cur.cigarElementLocus - cur.indexWithinCigarElement) : 0;
        cur.locus = newLocus;
        cur.readPosition += readPositionOffset;
        cur.indexWithinCigarElement = (int)(newLocus - cur.cigarElementLocus);
        return cur;
      } else if (newLocus == 0 && cur.cigarElement().op == OP_I) {  // :227-232
        return cur;
      } else {
        cur = cur.advanceToNextCigarElement();
      }
    }
  }
  // object PileupElement.apply (:250-259)
  static PileupElement make(const MappedRead& read, int64_t locus, const ContigSequence& contigSequence) {
    if (!(locus >= read.start) || !(locus < read.end)) throw InvalidInput("read does not overlap locus");  // :33-34
    PileupElement e;
    e.read = &read;
    e.locus = read.start;
    e.cigarElementLocus = read.start;
    e.contigSequence = &contigSequence;
    return e.advanceToLocus(locus);
  }
};

inline int MappedRead::countOfMismatches(const ContigSequence& contigSequence) const {  // M/reads/MappedRead.scala:53-64
  if (cachedCountOfMismatches == -1) {
    PileupElement element = PileupElement::make(*this, start, contigSequence);
    int count = element.alignment().kind == MISMATCH ? 1 : 0;
    while (element.locus < end - 1) {
      element = element.advanceToLocus(element.locus + 1);
      count += element.alignment().kind == MISMATCH ? 1 : 0;
    }
    cachedCountOfMismatches = count;
  }
  return cachedCountOfMismatches;
}

// ---------------------------------------------------------------------------------
// Pileup (M/pileup/Pileup.scala:19-126).  Elements carry their alignment, computed once
// (the Scala uses lazy vals).
// ---------------------------------------------------------------------------------
struct Element {
  PileupElement pe;
  Alignment aln;
  Allele allele;
  int qualityScore;
};
inline Element makeElement(const PileupElement& pe) {
  Element e;
  e.pe = pe;
  e.aln = pe.alignment();
  e.allele = Allele{e.aln.referenceBases, e.aln.sequencedBases};
  e.qualityScore = PileupElement::qualityScoreOf(e.aln, *pe.read);
  return e;
}

struct Pileup {
  int sampleId = 0;
  int32_t contig = 0;
  int64_t locus = 0;
  const ContigSequence* contigSequence = nullptr;
  std::vector<Element> elements;

  uint8_t referenceBase() const { return contigSequence->apply(locus); }  // :25
  int depth() const { return (int)elements.size(); }                      // :37
  int positiveDepth() const {                                             // :42
    int n = 0;
    for (auto& e : elements) n += e.pe.read->isPositiveStrand;
    return n;
  }
  int referenceDepth() const {  // :47-52
    int n = 0;
    for (auto& e : elements) n += e.aln.kind == MATCH;
    return n;
  }
  std::vector<Allele> distinctAlleles() const {  // :32
    std::vector<Allele> v;
    for (auto& e : elements)
      if (std::find(v.begin(), v.end(), e.allele) == v.end()) v.push_back(e.allele);
    std::sort(v.begin(), v.end());
    return v;
  }
  std::pair<int, int> alleleReadDepthAndPositiveDepth(const Allele& allele) const {  // :100-106
    int n = 0, p = 0;
    for (auto& e : elements)
      if (e.allele == allele) {
        ++n;
        p += e.pe.read->isPositiveStrand;
      }
    return {n, p};
  }
  // :63-92
  Pileup atGreaterLocus(int64_t newLocus, const std::vector<const MappedRead*>& newReads) const {
    if (!(elements.empty() || newLocus > locus)) throw InvalidInput("New locus not greater than current locus");
    Pileup p;
    p.sampleId = sampleId;
    p.contig = contig;
    p.locus = newLocus;
    p.contigSequence = contigSequence;
    for (auto& e : elements)
      if (e.pe.read->overlapsLocus(newLocus)) p.elements.push_back(makeElement(e.pe.advanceToLocus(newLocus)));
    for (auto* r : newReads) p.elements.push_back(makeElement(PileupElement::make(*r, newLocus, *contigSequence)));
    return p;
  }
  // object Pileup.apply (:119-126)
  static Pileup make(const std::vector<const MappedRead*>& reads, int sampleId, int32_t contig, int64_t locus,
                     const ContigSequence& cs) {
    Pileup p;
    p.sampleId = sampleId;
    p.contig = contig;
    p.locus = locus;
    p.contigSequence = &cs;
    for (auto* r : reads) p.elements.push_back(makeElement(PileupElement::make(*r, locus, cs)));
    return p;
  }
};

// ---------------------------------------------------------------------------------
// scala.collection.mutable.PriorityQueue (Scala 2.11.8, third party): binary heap in a
// 1-based array; `toVector` yields array order.  Restated so that the initial element
// order of a Pileup (M/windowing/SlidingWindow.scala:43-48) follows the reference.
// Ordering = Interval.orderByEndDesc (genomic-loci): the head is the region with the
// smallest end, i.e. "x < y" iff x.end > y.end.
// ---------------------------------------------------------------------------------
struct RegionQueue {
  std::vector<const MappedRead*> arr{nullptr};  // index 0 unused
  static bool lt(const MappedRead* x, const MappedRead* y) { return x->end > y->end; }
  size_t size() const { return arr.size() - 1; }
  bool nonEmpty() const { return size() > 0; }
  const MappedRead* head() const { return arr[1]; }
  void enqueue(const MappedRead* r) {
    arr.push_back(r);
    size_t k = arr.size() - 1;
    while (k > 1 && lt(arr[k / 2], arr[k])) {
      std::swap(arr[k], arr[k / 2]);
      k /= 2;
    }
  }
  const MappedRead* dequeue() {
    size_t last = arr.size() - 1;
    std::swap(arr[1], arr[last]);
    size_t n = last - 1;
    size_t k = 1;
    while (n >= 2 * k) {
      size_t j = 2 * k;
      if (j < n && lt(arr[j], arr[j + 1])) ++j;
      if (!lt(arr[k], arr[j])) break;
      std::swap(arr[k], arr[j]);
      k = j;
    }
    const MappedRead* r = arr[last];
    arr.pop_back();
    return r;
  }
  std::vector<const MappedRead*> toVector() const { return std::vector<const MappedRead*>(arr.begin() + 1, arr.end()); }
};

// ---------------------------------------------------------------------------------
// LociIterator over one contig's sorted disjoint ranges (genomic-loci, third party):
// head / next / skipTo.
// ---------------------------------------------------------------------------------
struct LociIterator {
  std::vector<std::pair<int64_t, int64_t>> ranges;
  size_t idx = 0;
  int64_t cur = 0;
  explicit LociIterator(std::vector<std::pair<int64_t, int64_t>> r) : ranges(std::move(r)) {
    if (!ranges.empty()) cur = ranges[0].first;
    normalise();
  }
  void normalise() {
    while (idx < ranges.size() && cur >= ranges[idx].second) {
      ++idx;
      if (idx < ranges.size()) cur = std::max(cur, ranges[idx].first);
    }
    if (idx < ranges.size() && cur < ranges[idx].first) cur = ranges[idx].first;
  }
  bool hasNext() const { return idx < ranges.size(); }
  int64_t head() const { return cur; }
  int64_t next() {
    int64_t r = cur;
    ++cur;
    normalise();
    return r;
  }
  void skipTo(int64_t locus) {
    if (locus > cur) {
      cur = locus;
      normalise();
    }
  }
};

// ---------------------------------------------------------------------------------
// SlidingWindow (M/windowing/SlidingWindow.scala:26-162)
// ---------------------------------------------------------------------------------
struct SlidingWindow {
  int halfWindowSize = 0;
  const std::vector<const MappedRead*>* sortedRegions = nullptr;  // this sample, this contig, sorted by start
  size_t nextRegion = 0;
  int64_t currentLocus = -1;
  std::vector<const MappedRead*> newRegions;
  RegionQueue queue;

  std::vector<const MappedRead*> currentRegions() const { return queue.toVector(); }  // :45-47
  // :58-85
  void setCurrentLocus(int64_t locus) {
    if (!(locus >= currentLocus)) throw InvalidInput("Pileup window can only move forward in locus");
    currentLocus = locus;
    while (queue.nonEmpty() && queue.head()->end <= locus - halfWindowSize) queue.dequeue();
    newRegions.clear();
    while (nextRegion < sortedRegions->size() && (*sortedRegions)[nextRegion]->start <= locus + halfWindowSize) {
      const MappedRead* region = (*sortedRegions)[nextRegion++];
      if (region->overlapsLocus(locus, halfWindowSize)) newRegions.push_back(region);
    }
    for (auto* r : newRegions) queue.enqueue(r);
  }
  // :93-103; returns -1 for None
  int64_t nextLocusWithRegions() const {
    for (size_t i = 1; i < queue.arr.size(); ++i)
      if (queue.arr[i]->overlapsLocus(currentLocus + 1, halfWindowSize)) return currentLocus + 1;
    if (nextRegion < sortedRegions->size())
      return std::max<int64_t>(0, (*sortedRegions)[nextRegion]->start - halfWindowSize);
    return -1;
  }
};

// object SlidingWindow.advanceMultipleWindows (:124-162); returns -1 for None
inline int64_t advanceMultipleWindows(std::vector<SlidingWindow>& windows, LociIterator& loci, bool skipEmpty) {
  if (skipEmpty) {
    while (loci.hasNext()) {
      int64_t nextNonEmptyLocus = -1;
      for (auto& w : windows) {
        int64_t n = w.nextLocusWithRegions();
        if (n >= 0 && (nextNonEmptyLocus < 0 || n < nextNonEmptyLocus)) nextNonEmptyLocus = n;
      }
      if (nextNonEmptyLocus < 0) return -1;
      if (nextNonEmptyLocus <= loci.head()) {
        int64_t nextLocus = loci.next();
        bool any = false;
        for (auto& w : windows) {
          w.setCurrentLocus(nextLocus);
          any = any || w.queue.nonEmpty();
        }
        if (any) return nextLocus;
      } else {
        loci.skipTo(nextNonEmptyLocus);
      }
    }
    return -1;
  } else if (loci.hasNext()) {
    int64_t nextLocus = loci.next();
    for (auto& w : windows) w.setCurrentLocus(nextLocus);
    return nextLocus;
  }
  return -1;
}

struct Range {
  int32_t contig;
  int64_t start, end;
};

// ---------------------------------------------------------------------------------
// windowFlatMapWithState + splitPartitionByContigAndMap
// (M/distributed/WindowFlatMapUtils.scala:40-76,119-145) specialised to the pileup
// operators of M/distributed/PileupFlatMapUtils.scala:35-141 (halfWindowSize 0; first
// locus Pileup(window.currentRegions()), later loci atGreaterLocus(newRegions)).
// `fn(pileups)` is the user closure, called once per visited locus.
// Contigs are processed in the order in which they first appear in `loci`.
// ---------------------------------------------------------------------------------
template <class F>
void pileupFlatMapMultipleSamples(const std::vector<MappedRead>& reads, int numSamples, const std::vector<Range>& loci,
                                  const std::vector<ContigSequence>& reference, bool skipEmpty, F fn) {
  std::vector<int32_t> contigOrder;
  for (auto& r : loci)
    if (std::find(contigOrder.begin(), contigOrder.end(), r.contig) == contigOrder.end()) contigOrder.push_back(r.contig);
  for (int32_t contig : contigOrder) {
    std::vector<std::pair<int64_t, int64_t>> rs;
    for (auto& r : loci)
      if (r.contig == contig) rs.push_back({r.start, r.end});
    std::sort(rs.begin(), rs.end());
    const ContigSequence* cs = nullptr;
    for (auto& c : reference)
      if (c.contig_id == contig) cs = &c;
    if (!cs) {
      // reference.getContig is only reached from the per-locus closure (PileupFlatMapUtils.scala:60,93):
      // a contig without reads is never visited when skipEmpty, so its absence goes unnoticed.
      bool anyRead = false;
      for (auto& r : reads) anyRead = anyRead || r.contig == contig;
      if (skipEmpty && !anyRead) continue;
      throw InvalidInput("contig " + std::to_string(contig) + " not found in reference");
    }
    // SplitIterator.split by sampleId + RegionsByContig (WindowFlatMapUtils.scala:93-106,128-135)
    std::vector<std::vector<const MappedRead*>> perSample(numSamples);
    int64_t lastStart = 0;
    for (auto& r : reads)
      if (r.contig == contig) {
        if (r.start < lastStart) throw InvalidInput("Regions must be sorted by start locus");  // SlidingWindow.scala:37
        lastStart = r.start;
        perSample.at(r.sampleId).push_back(&r);
      }
    std::vector<SlidingWindow> windows(numSamples);
    for (int s = 0; s < numSamples; ++s) windows[s].sortedRegions = &perSample[s];
    LociIterator it(rs);
    std::vector<Pileup> pileups;
    bool have = false;
    while (advanceMultipleWindows(windows, it, skipEmpty) >= 0) {
      if (!have) {
        pileups.clear();
        for (int s = 0; s < numSamples; ++s)
          pileups.push_back(Pileup::make(windows[s].currentRegions(), s, contig, windows[s].currentLocus, *cs));
        have = true;
      } else {
        for (int s = 0; s < numSamples; ++s)
          pileups[s] = pileups[s].atGreaterLocus(windows[s].currentLocus, windows[s].newRegions);
      }
      fn(pileups);
    }
  }
}

// ---------------------------------------------------------------------------------
// Genotype (M/variants/Genotype.scala:10-15): allele -> fraction, insertion ordered
// (scala Map1..Map4 keep insertion order).
// ---------------------------------------------------------------------------------
struct Genotype {
  std::vector<std::pair<Allele, double>> alleleMixture;
};

// breeze 0.12 (third party) softmax / logNormalize on a DenseVector[Double]:
// softmax(v) = max + log(sum exp(v - max)) (max itself when infinite); logNormalize = v - softmax(v).
inline std::vector<double> logNormalize(const std::vector<double>& v) {
  double mx = -std::numeric_limits<double>::infinity();
  for (double x : v) mx = std::max(mx, x);
  double sm;
  if (std::isinf(mx)) {
    sm = mx;
  } else {
    double acc = 0;
    for (double x : v) acc += std::exp(x - mx);
    sm = mx + std::log(acc);
  }
  std::vector<double> r;
  for (double x : v) r.push_back(x - sm);
  return r;
}

// Likelihood.likelihoodsOfGenotypes + computeAlleleElementProbabilities
// (M/likelihood/Likelihood.scala:99-151,177-210); prior is uniform (log 1 = 0) unless given.
inline std::vector<double> likelihoodsOfGenotypes(const std::vector<Element>& elements, const std::vector<Genotype>& genotypes,
                                                  bool includeAlignment, bool logSpace, bool normalize,
                                                  const std::vector<double>* priors = nullptr) {
  std::vector<double> successProbability;
  for (auto& e : elements) {
    double p = phredToSuccessProbability(e.qualityScore);
    if (includeAlignment) p = p * e.pe.read->alignmentLikelihood();
    successProbability.push_back(p);
  }
  std::vector<double> logLikelihoods;
  for (size_t g = 0; g < genotypes.size(); ++g) {
    double total = 0;
    for (size_t i = 0; i < elements.size(); ++i) {
      double s = 0;
      bool first = true;
      for (auto& af : genotypes[g].alleleMixture) {
        double probability = (af.first == elements[i].allele) ? successProbability[i] : 1 - successProbability[i];
        double term = probability * af.second;
        s = first ? term : s + term;
        first = false;
      }
      total += std::log(s);
    }
    logLikelihoods.push_back(total + std::log(priors ? (*priors)[g] : 1.0));
  }
  std::vector<double> r = normalize ? logNormalize(logLikelihoods) : logLikelihoods;
  if (!logSpace)
    for (auto& x : r) x = std::exp(x);
  return r;
}

// Likelihood.probabilitiesOfAllPossibleGenotypesFromPileup (M/likelihood/Likelihood.scala:31-63)
inline std::vector<std::pair<Genotype, double>> probabilitiesOfAllPossibleGenotypesFromPileup(const Pileup& pileup,
                                                                                            bool includeAlignment,
                                                                                            bool logSpace) {
  std::vector<Allele> alleles;
  for (auto& a : pileup.distinctAlleles())
    if (allStandardBases(a.altBases)) alleles.push_back(a);
  std::vector<Genotype> genotypes;
  for (size_t i = 0; i < alleles.size(); ++i)
    for (size_t j = i; j < alleles.size(); ++j) {
      Genotype g;
      if (i == j)
        g.alleleMixture = {{alleles[i], 1.0}};
      else
        g.alleleMixture = {{alleles[i], 0.5}, {alleles[j], 0.5}};
      genotypes.push_back(g);
    }
  std::vector<double> l = likelihoodsOfGenotypes(pileup.elements, genotypes, includeAlignment, logSpace, true);
  std::vector<std::pair<Genotype, double>> r;
  for (size_t i = 0; i < genotypes.size(); ++i) r.push_back({genotypes[i], l[i]});
  return r;
}

// ---------------------------------------------------------------------------------
// breeze 0.12 stats.mean / stats.median (third party).  mean: running mean
// mu += (x - mu)/n.  median: odd n -> middle; even n -> (lower + upper)/2 in the element
// type (integer division for DenseVector[Int]).  Even-length behaviour is not pinned by
// any reference test.
// ---------------------------------------------------------------------------------
inline double breezeMean(const std::vector<double>& v) {
  double mu = 0;
  long n = 0;
  for (double y : v) {
    n += 1;
    double d = y - mu;
    mu = mu + 1.0 / n * d;
  }
  return mu;
}
inline double breezeMedianDouble(std::vector<double> v) {
  std::sort(v.begin(), v.end());
  size_t n = v.size();
  if (n % 2 == 1) return v[(n - 1) / 2];
  return (v[n / 2 - 1] + v[n / 2]) / 2;
}
inline int breezeMedianInt(std::vector<int> v) {
  std::sort(v.begin(), v.end());
  size_t n = v.size();
  if (n % 2 == 1) return v[(n - 1) / 2];
  return (v[n / 2 - 1] + v[n / 2]) / 2;
}

// AlleleEvidence (M/variants/AlleleEvidence.scala:25-85)
struct AlleleEvidence {
  double probability = 0;
  int readDepth = 0, alleleReadDepth = 0, forwardDepth = 0, alleleForwardDepth = 0;
  double meanMappingQuality = NAN, medianMappingQuality = NAN, meanBaseQuality = NAN, medianBaseQuality = NAN,
         medianMismatchesPerRead = NAN;
  int phredScaledLikelihood() const { return successProbabilityToPhred(probability); }     // :36
  float variantAlleleFrequency() const { return (float)alleleReadDepth / (float)readDepth; }  // :37
};

inline AlleleEvidence makeAlleleEvidence(double likelihood, const Allele& allele, const Pileup& pileup) {  // :42-85
  auto dp = pileup.alleleReadDepthAndPositiveDepth(allele);
  AlleleEvidence ev;
  ev.probability = likelihood;
  ev.readDepth = pileup.depth();
  ev.alleleReadDepth = dp.first;
  ev.forwardDepth = pileup.positiveDepth();
  ev.alleleForwardDepth = dp.second;
  std::vector<double> alignmentScores, baseQualityScores;
  std::vector<int> mismatches;
  for (auto& e : pileup.elements)
    if (e.allele == allele) {
      alignmentScores.push_back(e.pe.read->alignmentQuality);
      baseQualityScores.push_back(e.qualityScore);
      mismatches.push_back(e.pe.read->countOfMismatches(*pileup.contigSequence));
    }
  if (!alignmentScores.empty()) {
    ev.meanMappingQuality = breezeMean(alignmentScores);
    ev.medianMappingQuality = breezeMedianDouble(alignmentScores);
    ev.meanBaseQuality = breezeMean(baseQualityScores);
    ev.medianBaseQuality = breezeMedianDouble(baseQualityScores);
    ev.medianMismatchesPerRead = breezeMedianInt(mismatches);
  }
  return ev;
}

// CalledSomaticAllele (M/variants/CalledSomaticAllele.scala:25-45)
struct CalledSomaticAllele {
  int32_t contig;
  int64_t start;
  Allele allele;
  double somaticLogOdds;
  double tumorRefLogProb, tumorAltLogProb, normalRefLogProb, normalAltLogProb;
  AlleleEvidence tumorVariantEvidence, normalReferenceEvidence;
  int phredScaledSomaticLikelihood() const {
    return successProbabilityToPhred(tumorVariantEvidence.probability * normalReferenceEvidence.probability);
  }
};

struct SomaticStandardParams {
  int normalOddsThreshold = 5, tumorOddsThreshold = 10;
  int maxReadDepth = std::numeric_limits<int>::max();
  int maxNormalAlternateReadDepth = 4;
  float minTumorVariantAlleleFrequency = 3 / 100.0f;
};

// Per-locus summary used to compare with the CUDA path's dense output.
struct LocusSummary {
  int depth, referenceDepth, topAltCount, flags;
};

// SomaticStandard.Caller.findPotentialVariantAtLocus (M/commands/SomaticStandardCaller.scala:155-253).
// Tie-break of `variantAlleleFractions.maxBy(_._2)` (:188): the reference's choice depends on
// Scala HashMap iteration order; oracle and CUDA path both take, among the most frequent
// variant alleles, the one whose first supporting read has the lowest batch index.
inline bool findPotentialVariantAtLocus(const Pileup& tumorPileup, const Pileup& normalPileup, const SomaticStandardParams& p,
                                        CalledSomaticAllele* out, LocusSummary* tumorSummary = nullptr,
                                        LocusSummary* normalSummary = nullptr) {
  int tumorDepth = tumorPileup.depth(), normalDepth = normalPileup.depth();
  int tumorRefDepth = tumorPileup.referenceDepth(), normalRefDepth = normalPileup.referenceDepth();
  if (tumorSummary) *tumorSummary = {tumorDepth, tumorRefDepth, 0, 0};
  if (normalSummary) *normalSummary = {normalDepth, normalRefDepth, 0, 0};
  if (tumorPileup.elements.empty() || normalPileup.elements.empty() || tumorDepth > p.maxReadDepth ||
      normalDepth > p.maxReadDepth || tumorRefDepth == tumorDepth ||
      normalDepth - normalRefDepth > p.maxNormalAlternateReadDepth)  // :164-172
    return false;

  uint8_t refBase = tumorPileup.referenceBase();
  Allele referenceAllele{std::string(1, (char)refBase), std::string(1, (char)refBase)};  // :174
  Genotype referenceGenotype{{{referenceAllele, 1.0}}};

  // :177-184 groupBy(identity) over variant alleles
  struct Cnt {
    Allele a;
    int n;
    int64_t firstRead;
  };
  std::vector<Cnt> counts;
  for (auto& e : tumorPileup.elements)
    if (e.allele.isVariant()) {
      bool found = false;
      for (auto& c : counts)
        if (c.a == e.allele) {
          ++c.n;
          c.firstRead = std::min(c.firstRead, e.pe.read->index);
          found = true;
        }
      if (!found) counts.push_back({e.allele, 1, e.pe.read->index});
    }
  if (counts.empty()) throw InvalidInput("empty.maxBy: every non-match tumor element is Clipped");  // :188 throws
  const Cnt* best = &counts[0];
  for (auto& c : counts)
    if (c.n > best->n || (c.n == best->n && c.firstRead < best->firstRead)) best = &c;
  Allele mostFrequentVariantAllele = best->a;
  double highestFrequency = best->n / (double)tumorDepth;
  double empiricalVariantAlleleFrequency = std::max((double)p.minTumorVariantAlleleFrequency, highestFrequency);  // :189
  if (tumorSummary) {
    tumorSummary->topAltCount = best->n;
    tumorSummary->flags |= 1;
  }
  if (normalSummary) {
    int n = 0;
    for (auto& e : normalPileup.elements) n += e.allele == mostFrequentVariantAllele;
    normalSummary->topAltCount = n;
    normalSummary->flags |= 1;
  }

  Genotype somaticVariantGenotype{
      {{referenceAllele, 1.0 - empiricalVariantAlleleFrequency}, {mostFrequentVariantAllele, empiricalVariantAlleleFrequency}}};
  auto t = likelihoodsOfGenotypes(tumorPileup.elements, {referenceGenotype, somaticVariantGenotype}, false, true, true);  // :201-208
  double tumorRefLogProb = t[0], tumorAltLogProb = t[1];
  double tumorAltLOD = tumorAltLogProb - tumorRefLogProb;

  Genotype germlineVariantGenotype{{{referenceAllele, 0.5}, {mostFrequentVariantAllele, 0.5}}};
  auto n = likelihoodsOfGenotypes(normalPileup.elements, {referenceGenotype, germlineVariantGenotype}, false, true, true);  // :220-227
  double normalRefLogProb = n[0], normalAltLogProb = n[1];
  double normalRefLOD = normalRefLogProb - normalAltLogProb;

  if (tumorAltLOD > p.tumorOddsThreshold && normalRefLOD > p.normalOddsThreshold &&
      !mostFrequentVariantAllele.altBases.empty()) {  // :231-233
    out->contig = tumorPileup.contig;
    out->start = tumorPileup.locus;
    out->allele = mostFrequentVariantAllele;
    out->somaticLogOdds = tumorAltLOD;
    out->tumorRefLogProb = tumorRefLogProb;
    out->tumorAltLogProb = tumorAltLogProb;
    out->normalRefLogProb = normalRefLogProb;
    out->normalAltLogProb = normalAltLogProb;
    out->tumorVariantEvidence = makeAlleleEvidence(std::exp(-tumorAltLogProb), mostFrequentVariantAllele, tumorPileup);  // :237
    out->normalReferenceEvidence = makeAlleleEvidence(std::exp(-normalRefLogProb), referenceAllele, normalPileup);     // :238
    if (tumorSummary) tumorSummary->flags |= 2;
    if (normalSummary) normalSummary->flags |= 2;
    return true;
  }
  return false;
}

}  // namespace guac_oracle
