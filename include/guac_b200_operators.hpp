// guac_b200_operators.hpp — Pileup / PileupElement views and the pileupFlatMap* operators (included by guac_b200.hpp).
#pragma once
#include <cctype>

namespace guac_b200 {

// pileup/PileupElement.scala:24-160 over one device-made record
struct PileupElement {
  const guac_element* e;
  int64_t readIndex() const { return e->read_index; }
  int readPosition() const { return e->read_position; }
  int cigarElementIndex() const { return e->cigar_element_index; }
  int64_t cigarElementLocus() const { return e->cigar_element_locus; }
  int indexWithinCigarElement() const { return e->index_within_cigar_element; }
  guac_alignment_kind alignment() const { return (guac_alignment_kind)e->kind; }
  bool isMatch() const { return e->kind == GUAC_ALN_MATCH; }
  bool isMismatch() const { return e->kind == GUAC_ALN_MISMATCH; }
  bool isInsertion() const { return e->kind == GUAC_ALN_INSERTION; }
  bool isDeletion() const { return e->kind == GUAC_ALN_DELETION; }
  bool isMidDeletion() const { return e->kind == GUAC_ALN_MID_DELETION; }
  bool isClipped() const { return e->kind == GUAC_ALN_CLIPPED; }
  char sequencedBase() const { return (char)e->base; }  // Match / Mismatch only
  int qualityScore() const { return e->quality; }
  int indelLength() const { return e->indel_length; }
};

// pileup/Pileup.scala:19-126 of one sample at one locus
struct Pileup {
  int32_t contigId;
  int64_t locus;
  int sampleId;
  std::vector<PileupElement> elements;
  int depth() const { return (int)elements.size(); }  // :37
  int referenceDepth() const {                        // :47-52
    int n = 0;
    for (auto& el : elements) n += el.isMatch();
    return n;
  }
};

using Ranges = std::vector<guac_range>;

namespace detail {

// Runs guac_pileup_elements and hands every visited locus's per-sample pileups to `visit`.
template <class Visit>
void for_each_locus(Context& ctx, const Ranges& loci, bool skipEmpty, Visit visit) {
  int32_t* contig = nullptr;
  int64_t* locus = nullptr;
  int64_t* offsets = nullptr;
  guac_element* elements = nullptr;
  int64_t n = 0;
  ctx.check(guac_pileup_elements(ctx.get(), loci.data(), (int32_t)loci.size(), skipEmpty ? 1 : 0, &contig, &locus, &offsets, &elements, &n));
  struct Release {
    guac_ctx* c;
    void* p[4];
    ~Release() {
      for (void* q : p) guac_free(c, q);
    }
  } release{ctx.get(), {contig, locus, offsets, elements}};
  std::vector<Pileup> per_sample((size_t)ctx.nSamples());
  for (int64_t i = 0; i < n; ++i) {
    for (int s = 0; s < ctx.nSamples(); ++s) {
      per_sample[(size_t)s].contigId = contig[i];
      per_sample[(size_t)s].locus = locus[i];
      per_sample[(size_t)s].sampleId = s;
      per_sample[(size_t)s].elements.clear();
    }
    for (int64_t k = offsets[i]; k < offsets[i + 1]; ++k) per_sample[elements[k].sample_id].elements.push_back(PileupElement{elements + k});
    visit(per_sample);
  }
}

}  // namespace detail

// M/distributed/PileupFlatMapUtils.scala:35-57.  `function` appends its results to `out`.
template <class T>
std::vector<T> pileupFlatMapOneSample(Context& ctx, const Ranges& loci, bool skipEmpty, int sample,
                                      const std::function<void(const Pileup&, std::vector<T>&)>& function) {
  std::vector<T> out;
  detail::for_each_locus(ctx, loci, skipEmpty, [&](const std::vector<Pileup>& p) {
    // with skipEmpty the reference only visits loci where this sample's window holds a read
    if (!skipEmpty || p[(size_t)sample].depth() > 0) function(p[(size_t)sample], out);
  });
  return out;
}

// :60-85
template <class T>
std::vector<T> pileupFlatMapTwoSamples(Context& ctx, const Ranges& loci, bool skipEmpty, int sample1, int sample2,
                                       const std::function<void(const Pileup&, const Pileup&, std::vector<T>&)>& function) {
  std::vector<T> out;
  detail::for_each_locus(ctx, loci, skipEmpty, [&](const std::vector<Pileup>& p) {
    if (!skipEmpty || p[(size_t)sample1].depth() + p[(size_t)sample2].depth() > 0) function(p[(size_t)sample1], p[(size_t)sample2], out);
  });
  return out;
}

// :88-141
template <class T>
std::vector<T> pileupFlatMapMultipleSamples(Context& ctx, const Ranges& loci, bool skipEmpty,
                                            const std::function<void(const std::vector<Pileup>&, std::vector<T>&)>& function) {
  std::vector<T> out;
  detail::for_each_locus(ctx, loci, skipEmpty, [&](const std::vector<Pileup>& p) { function(p, out); });
  return out;
}

}  // namespace guac_b200
