// guac_b200.hpp — C++ host mirror of the reference's operator interface over the C ABI of guac_b200.h.
//
// The reference's callers reach the hot path through three operators of
// org.hammerlab.guacamole.distributed.PileupFlatMapUtils (M/distributed/PileupFlatMapUtils.scala):
//   pileupFlatMapOneSample       (:35-57)   function: Pileup => Iterator[T]
//   pileupFlatMapTwoSamples      (:60-85)   function: (Pileup, Pileup) => Iterator[T]
//   pileupFlatMapMultipleSamples (:88-141)  function: PerSample[Pileup] => Iterator[T]
// each over one partition's reads and loci, with skipEmpty.  The same names, argument meaning and error behaviour are
// kept here (errors surface as guac_b200::Error carrying the reference's wording); the per-locus Pileup objects are
// views over the guac_element records the device materialised (guac_pileup_elements), so an arbitrary host closure
// runs on real pileups while every CIGAR walk happened on the GPU.  Header-only; link with libguac_b200.so.
#pragma once
#include <cctype>
#include <cstdint>
#include <cstring>
#include <functional>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "guac_b200.h"

namespace guac_b200 {

struct Error : std::runtime_error {
  int status;
  Error(int s, const std::string& m) : std::runtime_error(m), status(s) {}
};

// One partition's reads in the columnar layout of guac_read_batch, built read by read
// (the Iterator[MappedRead] of M/readsets/rdd/PartitionedRegions.scala:56-72).
class ReadBatchBuilder {
 public:
  explicit ReadBatchBuilder(int n_samples) : n_samples_(n_samples) { cigar_off_.push_back(0); seq_off_.push_back(0); }
  // cigar as text ("8M", "3M1D4M"); qualities default to 31 and the mapping quality to 30 like ReadsUtil.makeRead
  // (src/test/scala/org/hammerlab/guacamole/reads/ReadsUtil.scala:62-86).  Reads must arrive sorted by (contig, start).
  void add(const std::string& sequence, const std::string& cigar, int64_t start, int32_t contig_id = 0, int sample_id = 0,
           int mapq = 30, bool positive_strand = true, const std::vector<uint8_t>* qualities = nullptr) {
    static const char* ops = "MIDNSHP=X";
    size_t i = 0;
    while (i < cigar.size()) {
      uint32_t len = 0;
      while (i < cigar.size() && cigar[i] >= '0' && cigar[i] <= '9') len = len * 10 + (uint32_t)(cigar[i++] - '0');
      const char* p = i < cigar.size() ? std::strchr(ops, cigar[i]) : nullptr;
      if (!p) throw Error(GUAC_E_INVALID_INPUT, "bad CIGAR: " + cigar);
      cigar_.push_back((len << 4) | (uint32_t)(p - ops));
      ++i;
    }
    cigar_off_.push_back((uint32_t)cigar_.size());
    for (char c : sequence) ascii_.push_back((uint8_t)std::toupper((unsigned char)c));
    for (size_t k = 0; k < sequence.size(); ++k) qual_.push_back(qualities ? (*qualities)[k] : (uint8_t)31);
    seq_off_.push_back((uint64_t)ascii_.size());
    contig_.push_back(contig_id);
    start_.push_back(start);
    mapq_.push_back((uint8_t)mapq);
    flags_.push_back((uint8_t)((positive_strand ? GUAC_READ_POSITIVE_STRAND : 0) | GUAC_READ_PAIRED));
    sample_.push_back((uint8_t)sample_id);
  }
  // Packs the bases (guac_pack_bases) and returns a view valid while the builder lives.
  const guac_read_batch& finish() {
    const int64_t n = (int64_t)ascii_.size();
    bases2_.assign((size_t)(n + 3) / 4 + 8, 0);
    nmask_.assign((size_t)(n + 7) / 8 + 8, 0);
    exc_pos_.assign((size_t)n + 1, 0);
    exc_base_.assign((size_t)n + 1, 0);
    int64_t n_exc = 0;
    const int st = guac_pack_bases(ascii_.data(), n, 0, bases2_.data(), nmask_.data(), exc_pos_.data(), exc_base_.data(), n + 1, &n_exc);
    if (st != GUAC_OK) throw Error(st, "guac_pack_bases failed");
    batch_ = guac_read_batch{};
    batch_.n_reads = (int64_t)start_.size();
    batch_.n_samples = n_samples_;
    batch_.contig_id = contig_.data();
    batch_.start = start_.data();
    batch_.cigar_off = cigar_off_.data();
    batch_.cigar = cigar_.data();
    batch_.seq_off = seq_off_.data();
    batch_.bases2 = bases2_.data();
    batch_.nmask = nmask_.data();
    batch_.n_exc = n_exc;
    batch_.exc_pos = exc_pos_.data();
    batch_.exc_base = exc_base_.data();
    batch_.qual = qual_.data();
    batch_.mapq = mapq_.data();
    batch_.flags = flags_.data();
    batch_.sample_id = sample_.data();
    return batch_;
  }

 private:
  int n_samples_;
  std::vector<int32_t> contig_;
  std::vector<int64_t> start_;
  std::vector<uint32_t> cigar_off_, cigar_;
  std::vector<uint64_t> seq_off_, exc_pos_;
  std::vector<uint8_t> ascii_, qual_, mapq_, flags_, sample_, bases2_, nmask_, exc_base_;
  guac_read_batch batch_{};
};

// guac_ctx: one CUDA device, one stream; not thread-safe (a Spark task thread takes one from a pool).
class Context {
 public:
  explicit Context(int device = 0) {
    const int st = guac_ctx_create(device, &ctx_);
    if (st != GUAC_OK) throw Error(st, st == GUAC_E_NO_DEVICE ? "no CUDA device (there is no CPU fallback)" : "guac_ctx_create failed");
  }
  ~Context() { guac_ctx_destroy(ctx_); }
  Context(const Context&) = delete;
  Context& operator=(const Context&) = delete;
  guac_ctx* get() const { return ctx_; }
  void check(int status) const {
    if (status != GUAC_OK) throw Error(status, guac_last_error(ctx_));
  }
  // ReferenceBroadcast window of one contig: bases [offset, offset + bases.size()) (M/reference/ReferenceBroadcast.scala:42-83)
  void uploadReference(int32_t contig_id, const std::string& bases, int64_t offset = 0) {
    check(guac_ref_upload(ctx_, contig_id, reinterpret_cast<const uint8_t*>(bases.data()), (int64_t)bases.size(), offset));
  }
  void reserveReference(int32_t contig_id, int64_t offset, int64_t len) { check(guac_ref_reserve(ctx_, contig_id, offset, len)); }
  void uploadReads(const guac_read_batch& batch) {
    check(guac_reads_upload(ctx_, &batch));
    n_samples_ = batch.n_samples;
  }
  int nSamples() const { return n_samples_; }

 private:
  guac_ctx* ctx_ = nullptr;
  int n_samples_ = 0;
};

}  // namespace guac_b200

#include "guac_b200_operators.hpp"
