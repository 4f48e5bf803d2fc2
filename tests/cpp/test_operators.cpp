// PileupFlatMapUtilsSuite (src/test/scala/org/hammerlab/guacamole/distributed/PileupFlatMapUtilsSuite.scala:80-345) written
// against the C++ host mirror (include/guac_b200.hpp): the operators keep the reference's names and argument meaning, the
// closures are ordinary host lambdas, every pileup element comes from the device.
// Build: g++ -std=c++17 -I include tests/cpp/test_operators.cpp -L guacamole_b200/csrc -lguac_b200 -Wl,-rpath,... -o test_operators
#include <cstdio>
#include <string>
#include <vector>

#include "guac_b200.hpp"

using namespace guac_b200;

static int failures = 0;
#define CHECK(cond)                                                   \
  do {                                                                \
    if (!(cond)) {                                                    \
      std::printf("FAIL %s:%d  %s\n", __FILE__, __LINE__, #cond);     \
      ++failures;                                                     \
    }                                                                 \
  } while (0)

static const std::string N90(90, 'N'), N500(500, 'N');

// makeReference(sc, ("chr1", 0, "ATCGATCGA" + "N"*90), ("chr1", 99, "ACGTACGTACGT" + "N"*500), ("chr2", 10, "N"*10)) (:47-53)
static void suiteReference(Context& ctx) {
  ctx.reserveReference(1, 0, 99 + 12 + 500);
  ctx.uploadReference(1, "ATCGATCGA" + N90, 0);
  ctx.uploadReference(1, "ACGTACGTACGT" + N500, 99);
  ctx.uploadReference(2, std::string(10, 'N'), 10);
}

static void dummyReads(ReadBatchBuilder& b) {  // dummyReadsRDD (:55-60)
  for (int i = 0; i < 3; ++i) b.add("TCGATCGA", "8M", 1, 1);
}

int main() {
  try {
    {  // "test pileup flatmap parallelism 0; create pileups" (:80-101)
      Context ctx(0);
      suiteReference(ctx);
      ReadBatchBuilder b(1);
      dummyReads(b);
      ctx.uploadReads(b.finish());
      auto pileups = pileupFlatMapOneSample<Pileup>(ctx, {{1, 0, 1, 9}}, false, 0,
                                                     [](const Pileup& p, std::vector<Pileup>& out) { out.push_back(p); });
      CHECK(pileups.size() == 8);
      CHECK(pileups.front().locus == 1);
      for (auto& e : pileups.front().elements) CHECK(e.readPosition() == 0 && e.isMatch());
      CHECK(pileups.back().locus == 8);
      for (auto& e : pileups.back().elements) CHECK(e.readPosition() == 7);
      for (auto& p : pileups) CHECK(p.depth() == 3 && p.elements.front().isMatch());
      // "skip empty pileups" (:124-138): loci over several contigs, only the covered ones are visited
      auto loci = pileupFlatMapOneSample<int64_t>(ctx, {{0, 0, 5, 10}, {1, 0, 0, 100}, {2, 0, 0, 1000}, {2, 0, 5000, 6000}}, true, 0,
                                                  [](const Pileup& p, std::vector<int64_t>& out) { out.push_back(p.locus); });
      CHECK((loci == std::vector<int64_t>{1, 2, 3, 4, 5, 6, 7, 8}));
    }
    {  // "test pileup flatmap two rdds; skip empty pileups" (:140-178)
      Context ctx(0);
      suiteReference(ctx);
      ReadBatchBuilder b(2);
      // sorted by start, samples interleaved, as PartitionedRegions hands a partition over
      for (int i = 0; i < 3; ++i) b.add("TCGATCGA", "8M", 1, 1, 0);
      b.add("AAAAAAAA", "8M", 1, 1, 1);
      b.add("CCCCCCCC", "8M", 1, 1, 1);
      b.add("TTTTTTTT", "8M", 1, 1, 1);
      b.add("XXX", "3M", 99, 1, 1);
      for (int i = 0; i < 3; ++i) b.add("GGGGGGGG", "8M", 100, 1, 0);
      ctx.uploadReads(b.finish());
      auto loci = pileupFlatMapTwoSamples<int64_t>(ctx, {{0, 0, 0, 1000}, {1, 0, 1, 500}, {2, 0, 10, 20}}, true, 0, 1,
                                                   [](const Pileup& p1, const Pileup&, std::vector<int64_t>& out) { out.push_back(p1.locus); });
      CHECK((loci == std::vector<int64_t>{1, 2, 3, 4, 5, 6, 7, 8, 99, 100, 101, 102, 103, 104, 105, 106, 107}));
      // "multiple rdds" shape (:180-257): per-sample depths at a locus covered by both samples and at one covered by one
      auto depths = pileupFlatMapMultipleSamples<std::vector<int>>(
          ctx, {{1, 0, 1, 500}}, true, [](const std::vector<Pileup>& ps, std::vector<std::vector<int>>& out) {
            if (ps[0].locus == 1 || ps[0].locus == 99) out.push_back({ps[0].depth(), ps[1].depth()});
          });
      CHECK((depths == std::vector<std::vector<int>>{{3, 3}, {0, 1}}));
    }
    {  // "test two-rdd pileup flatmap; create pileup elements" (:276-318)
      Context ctx(0);
      ctx.uploadReference(1, "ATCGATCGA" + N90 + "AGGGGGGGGGG" + N500, 0);
      ReadBatchBuilder b(2);
      for (int s = 0; s < 2; ++s)
        for (int i = 0; i < 3; ++i) b.add("TCGATCGA", "8M", 1, 1, s);
      b.add("AGG", "3M", 99, 1, 1);
      for (int i = 0; i < 3; ++i) b.add("GGGGGGGG", "8M", 100, 1, 0);
      ctx.uploadReads(b.finish());
      std::string bases;
      bool all_match = true;
      pileupFlatMapTwoSamples<int>(ctx, {{1, 0, 1, 500}}, false, 0, 1, [&](const Pileup& p1, const Pileup& p2, std::vector<int>&) {
        for (auto& e : p1.elements) { all_match = all_match && e.isMatch(); bases.push_back(e.sequencedBase()); }
        for (auto& e : p2.elements) { all_match = all_match && e.isMatch(); bases.push_back(e.sequencedBase()); }
      });
      CHECK(all_match);
      CHECK(bases == "TTTTTTCCCCCCGGGGGGAAAAAATTTTTTCCCCCCGGGGGGAAAAAAAGGGGGGGGGGGGGGGGGGGGGGGGGG");
    }
    {  // "create pileup elements; with indel" (:320-344)
      Context ctx(0);
      ctx.uploadReference(1, "ATCGATCGATC", 1);
      ReadBatchBuilder b(1);
      b.add("TCGATCGA", "8M", 1, 1);
      b.add("TCGATCGA", "8M", 1, 1);
      b.add("TCGACCCTCGA", "4M3I4M", 1, 1);
      ctx.uploadReads(b.finish());
      int n = 0, insertions = 0;
      pileupFlatMapOneSample<int>(ctx, {{1, 0, 1, 12}}, false, 0, [&](const Pileup& p, std::vector<int>&) {
        for (auto& e : p.elements) { ++n; insertions += e.isInsertion(); }
      });
      CHECK(n == 24);
      CHECK(insertions == 1);
    }
    {  // error wording crosses the boundary: reads out of order (windowing/SlidingWindow.scala:36-37)
      Context ctx(0);
      ReadBatchBuilder b(1);
      b.add("ACGT", "4M", 20, 1);
      b.add("ACGT", "4M", 10, 1);
      bool thrown = false;
      try {
        ctx.uploadReads(b.finish());
      } catch (const Error& e) {
        thrown = std::string(e.what()).find("Regions must be sorted by start locus") != std::string::npos;
      }
      CHECK(thrown);
    }
  } catch (const std::exception& e) {
    std::printf("FAIL exception: %s\n", e.what());
    return 2;
  }
  std::printf(failures ? "FAILED (%d)\n" : "ok\n", failures);
  return failures ? 1 : 0;
}
