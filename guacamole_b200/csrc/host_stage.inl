// host_stage.inl — the host halves either side of the device path, in C++ behind the C ABI (part of the guac_b200.cu
// translation unit; no device work in this file):
//   guac_pack_reads          decoded SAM records -> MappedRead fields (reads/Read.scala:63-116), the ingest filters of
//                            ReadSets.filterRDD (readsets/ReadSets.scala:272-304), (contig, start) order as
//                            PartitionedRegions hands a partition over, and the compact wire form of guac_read_batch_compact
//   guac_ss_genotype_filter  SomaticGenotypeFilter.apply (filters/somatic/SomaticGenotypeFilter.scala:82-130)
//   guac_ss_vcf_rows         the rows the reference's VCF writer renders for CalledSomaticAllele.toBDGGenotype
//                            (variants/CalledSomaticAllele.scala:45-64, variants/GenotypeOutputArgs.scala:86-113)
//   guac_sj_vcf_rows         VCFOutput.writeVcf's rows (jointcaller/VCFOutput.scala:97-100,127-413 through htsjdk 2.6.1)
// Reference file names are relative to src/main/scala/org/hammerlab/guacamole/.

namespace {

template <class F>
void host_parallel_for(int n_threads, int64_t n, F fn) {  // fn(lo, hi, thread)
  n_threads = (int)std::max<int64_t>(1, std::min<int64_t>(n_threads, (n + 4095) / 4096));
  if (n_threads == 1) {
    fn((int64_t)0, n, 0);
    return;
  }
  std::vector<std::thread> pool;
  const int64_t per = (n + n_threads - 1) / n_threads;
  for (int t = 0; t < n_threads; ++t) pool.emplace_back([=] { fn(std::min(n, t * per), std::min(n, (t + 1) * per), t); });
  for (auto& th : pool) th.join();
}

// An array whose every element the passes below write before anything reads it: no serial zero-fill of tens of megabytes,
// and the pages are first touched by the threads that use them.
template <class T>
struct Uninit {
  std::unique_ptr<T[]> p;
  size_t n = 0;
  void alloc(size_t k) {
    p.reset(new T[k]);
    n = k;
  }
  T& operator[](size_t i) { return p[i]; }
  const T& operator[](size_t i) const { return p[i]; }
  T* begin() { return p.get(); }
  T* end() { return p.get() + n; }
};

// memory of guac_pack_reads' outputs: the context's pinned pool (fast H2D) or, without a context, the C heap
struct PackAlloc {
  guac_ctx* ctx;
  std::vector<void*> got;
  bool ok = true;
  void* get(size_t bytes) {
    void* p = nullptr;
    bytes = std::max<size_t>(bytes, 16);
    if (ctx) {
      if (pinned_alloc(ctx, &p, bytes) != GUAC_OK) p = nullptr;
    } else {
      p = std::malloc(bytes);
    }
    if (!p) ok = false; else got.push_back(p);
    return p;
  }
  void release_all() {
    for (void* p : got) {
      if (ctx) guac_free(ctx, p); else std::free(p);
    }
    got.clear();
  }
};

enum : uint16_t { SAM_PAIRED = 0x1, SAM_UNMAPPED = 0x4, SAM_REVERSE = 0x10, SAM_FAILED_QC = 0x200, SAM_DUPLICATE = 0x400 };

// ---- number formatting as the JVM does it ----------------------------------------------------------------------------
// Java's Formatter rounds HALF_UP on the exact decimal expansion of the double; printf rounds half-even.  glibc prints the
// exact expansion when asked for enough digits, and HALF_UP only looks at the first dropped digit: round by hand.
struct ExactDecimal {
  bool negative;
  std::string digits;  // significant digits, first one non-zero (or "0")
  int exponent;        // value = 0.d1d2d3... * 10^(exponent + 1), i.e. d1 is the 10^exponent digit
};

ExactDecimal exact_decimal(double x) {
  ExactDecimal d;
  d.negative = std::signbit(x);
  x = std::fabs(x);
  if (x == 0) {
    d.digits = "0";
    d.exponent = 0;
    return d;
  }
  char buf[128];
  std::snprintf(buf, sizeof buf, "%.40e", x);  // d.ddddd...e+XX, exact digits (glibc), far beyond any tie we round at
  d.digits.push_back(buf[0]);
  const char* p = buf + 2;
  while (*p && *p != 'e') d.digits.push_back(*p++);
  d.exponent = std::atoi(p + 1);
  return d;
}

// digits rounded HALF_UP to `keep` significant digits (keep >= 1); may carry into a new leading digit
void round_half_up(ExactDecimal& d, int keep) {
  if ((int)d.digits.size() <= keep) {
    d.digits.resize((size_t)keep, '0');
    return;
  }
  const bool up = d.digits[(size_t)keep] >= '5';
  d.digits.resize((size_t)keep);
  if (!up) return;
  int i = keep - 1;
  while (i >= 0 && d.digits[(size_t)i] == '9') d.digits[(size_t)i--] = '0';
  if (i >= 0) {
    ++d.digits[(size_t)i];
  } else {
    d.digits.insert(d.digits.begin(), '1');
    d.digits.resize((size_t)keep);
    ++d.exponent;
  }
}

// fixed notation with `frac` fractional digits, HALF_UP (String.format("%.<frac>f"))
std::string java_fixed(double x, int frac) {
  if (std::isnan(x)) return "NaN";
  if (std::isinf(x)) return x > 0 ? "Infinity" : "-Infinity";
  ExactDecimal d = exact_decimal(x);
  // significant digits to keep = digits before the point + frac
  int keep = d.exponent + 1 + frac;
  std::string out;
  if (keep <= 0) {  // rounds to zero unless the first dropped digit (keep == 0) is >= 5
    const bool up = keep == 0 && d.digits[0] >= '5';
    std::string s = "0";
    if (frac > 0) {
      s += "." + std::string((size_t)frac, '0');
      if (up) s[s.size() - 1] = '1';
    } else if (up) {
      s = "1";
    }
    return (d.negative && up ? "-" : "") + s;
  }
  round_half_up(d, keep);
  keep = d.exponent + 1 + frac;  // a carry lengthens the integer part
  d.digits.resize((size_t)keep, '0');
  const int int_digits = d.exponent + 1;
  if (int_digits <= 0) out = "0." + std::string((size_t)(-int_digits), '0') + d.digits;
  else out = d.digits.substr(0, (size_t)int_digits) + (frac > 0 ? "." + d.digits.substr((size_t)int_digits) : "");
  bool all_zero = true;
  for (char c : out)
    if (c != '0' && c != '.') all_zero = false;
  return (d.negative && !all_zero ? "-" : "") + out;
}

// String.format("%.<sig-1>e"): d.ddd e±XX with at least two exponent digits
std::string java_sci(double x, int sig) {
  ExactDecimal d = exact_decimal(x);
  round_half_up(d, sig);
  char e[16];
  std::snprintf(e, sizeof e, "e%c%02d", d.exponent < 0 ? '-' : '+', std::abs(d.exponent));
  return std::string(d.negative ? "-" : "") + d.digits.substr(0, 1) + (sig > 1 ? "." + d.digits.substr(1) : "") + e;
}

// String.format("%.8g", x): 8 significant digits; fixed notation for 1e-4 <= |x| < 1e8 (after rounding), trailing zeros kept
std::string java_g8(double x) {
  if (std::isnan(x)) return "NaN";
  if (std::isinf(x)) return x > 0 ? "Infinity" : "-Infinity";
  if (x == 0) return "0.0000000";
  ExactDecimal d = exact_decimal(x);
  round_half_up(d, 8);
  if (d.exponent >= -4 && d.exponent < 8) {
    const int frac = std::max(0, 7 - d.exponent);
    std::string out;
    const int int_digits = d.exponent + 1;
    if (int_digits <= 0) out = "0." + std::string((size_t)(-int_digits), '0') + d.digits;
    else out = d.digits.substr(0, (size_t)int_digits) + (frac > 0 ? "." + d.digits.substr((size_t)int_digits) : "");
    return (d.negative ? "-" : "") + out;
  }
  char e[16];
  std::snprintf(e, sizeof e, "e%c%02d", d.exponent < 0 ? '-' : '+', std::abs(d.exponent));
  return std::string(d.negative ? "-" : "") + d.digits.substr(0, 1) + "." + d.digits.substr(1) + e;
}

// htsjdk VCFEncoder.formatVCFDouble
std::string format_vcf_double(double d) {
  if (d < 1) {
    if (d < 0.01) {
      if (std::fabs(d) >= 1e-20) return java_sci(d, 4);
      return "0.00";
    }
    return java_fixed(d, 3);
  }
  return java_fixed(d, 2);
}

// util/PhredUtils.scala:7-14 over ADAM's successProbabilityToPhred
int success_probability_to_phred(double probability) {
  const double p = std::min(1 - 1e-16, probability);
  return (int)std::floor(-10.0 * std::log10(1.0 - p) + 0.5);
}
int phred_scaled_somatic_likelihood(const guac_ss_call& c) {  // variants/CalledSomaticAllele.scala:38-43
  return success_probability_to_phred(c.tumor_variant_evidence.probability * c.normal_reference_evidence.probability);
}

int copy_text_out(const std::string& text, char* out, int64_t cap, int64_t* needed) {
  if (needed) *needed = (int64_t)text.size();
  if (!out || cap < (int64_t)text.size()) return out ? GUAC_E_OOM : GUAC_OK;
  std::memcpy(out, text.data(), text.size());
  return GUAC_OK;
}

}  // namespace

extern "C" {

void guac_host_free(void* p) { std::free(p); }

int guac_format_double(int32_t kind, double x, char* out, int32_t cap) {
  if (!out || cap <= 0) return GUAC_E_INVALID_INPUT;
  const std::string s = kind == 0 ? java_g8(x) : kind == 1 ? format_vcf_double(x) : java_fixed(x, 2);
  if ((int)s.size() + 1 > cap) return GUAC_E_OOM;
  std::memcpy(out, s.c_str(), s.size() + 1);
  return GUAC_OK;
}

namespace {
int pack_reads_impl(guac_ctx* ctx, const guac_sam_records* in, const guac_read_filters* filters, const guac_host_reference* refs, int32_t n_refs,
                    int32_t n_threads, guac_read_batch_compact* out, int64_t* n_kept) {
  if (!in || !out || in->n_records < 0 || in->n_samples <= 0 || in->n_samples > 255 || n_refs < 0 || (n_refs > 0 && !refs))
    return GUAC_E_INVALID_INPUT;
  auto failp = [&](int st, const char* msg) {
    if (ctx) ctx->err = msg;
    return st;
  };
  const int64_t N = in->n_records;
  if (n_threads <= 0) n_threads = (int)std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 64u);
  guac_read_filters F;
  std::memset(&F, 0, sizeof F);
  if (filters) F = *filters;
  std::memset(out, 0, sizeof *out);
  // 1. Read.apply + ReadSets.filterRDD per record: kept <=> a mapped read that passes every filter
  Uninit<uint8_t> keep;
  keep.alloc((size_t)N + 1);
  Uninit<int64_t> ref_len;
  ref_len.alloc((size_t)N + 1);  // padded reference length (MappedRead.scala:75), for overlapsLoci
  host_parallel_for(n_threads, N, [&](int64_t lo, int64_t hi, int) {
    for (int64_t i = lo; i < hi; ++i) {
      const uint16_t fl = in->sam_flags[i];
      int64_t lead_clip = 0, rl = 0;
      bool seen_aligned = false;
      for (uint32_t c = in->cigar_off[i]; c < in->cigar_off[i + 1]; ++c) {
        const int op = in->cigar[c] & 15;
        const int64_t len = in->cigar[c] >> 4;
        if (!seen_aligned && (op == GUAC_CIGAR_S || op == GUAC_CIGAR_H)) lead_clip += len; else seen_aligned = true;
        if (op == GUAC_CIGAR_M || op == GUAC_CIGAR_D || op == GUAC_CIGAR_N || op == GUAC_CIGAR_EQ || op == GUAC_CIGAR_X || op == GUAC_CIGAR_P)
          rl += len;
      }
      ref_len[(size_t)i] = rl;
      // isMapped (Read.scala:65-70): getAlignmentStart / getUnclippedStart are 1-based
      bool k = !(fl & SAM_UNMAPPED) && in->reference_index[i] >= 0 && in->alignment_start[i] >= 0 &&
               in->alignment_start[i] - lead_clip >= 0 && in->alignment_start[i] >= 1 && in->sample_id[i] < in->n_samples;
      if (k && F.non_duplicate && (fl & SAM_DUPLICATE)) k = false;
      if (k && F.passed_vendor_quality_checks && (fl & SAM_FAILED_QC)) k = false;
      if (k && F.is_paired && !(fl & SAM_PAIRED)) k = false;
      if (k && F.min_alignment_quality > 0 && in->mapping_quality[i] < F.min_alignment_quality) k = false;
      if (k && F.n_overlaps_loci > 0) {  // loci.intersects(read): any eligible range of the contig overlaps [start, end)
        const int64_t s = in->alignment_start[i] - 1, e = s + rl;
        bool hit = false;
        for (int r = 0; r < F.n_overlaps_loci && !hit; ++r)
          hit = F.overlaps_loci[r].contig_id == in->reference_index[i] && F.overlaps_loci[r].start < e && s < F.overlaps_loci[r].end;
        k = hit;
      }
      keep[(size_t)i] = k;
    }
  });
  // the kept records' indices, in input order: counts per chunk, their prefix, then every chunk writes its own part
  std::vector<int64_t> idx;
  {
    std::vector<int64_t> kept((size_t)n_threads + 1, 0);
    host_parallel_for(n_threads, N, [&](int64_t lo, int64_t hi, int t) {
      int64_t c = 0;
      for (int64_t i = lo; i < hi; ++i) c += keep[(size_t)i];
      kept[(size_t)t + 1] = c;
    });
    for (int t = 0; t < n_threads; ++t) kept[(size_t)t + 1] += kept[(size_t)t];
    idx.resize((size_t)kept[(size_t)n_threads]);
    host_parallel_for(n_threads, N, [&](int64_t lo, int64_t hi, int t) {
      int64_t w = kept[(size_t)t];
      for (int64_t i = lo; i < hi; ++i)
        if (keep[(size_t)i]) idx[(size_t)w++] = i;
    });
  }
  const int64_t n = (int64_t)idx.size();
  // 2. (contig, start) order, stable (repartitionAndSortWithinPartitions of PartitionedRegions.scala:186-208)
  bool sorted = true;
  {
    std::vector<uint8_t> unsorted((size_t)n_threads, 0);
    host_parallel_for(n_threads, n, [&](int64_t lo, int64_t hi, int t) {
      bool ok = true;
      for (int64_t k = std::max<int64_t>(lo, 1); k < hi && ok; ++k) {  // (k = lo compares across the chunk boundary)
        const int64_t a = idx[(size_t)k - 1], b = idx[(size_t)k];
        ok = in->reference_index[a] < in->reference_index[b] ||
             (in->reference_index[a] == in->reference_index[b] && in->alignment_start[a] <= in->alignment_start[b]);
      }
      unsorted[(size_t)t] = !ok;
    });
    for (uint8_t u : unsorted) sorted = sorted && !u;
  }
  if (!sorted)
    std::stable_sort(idx.begin(), idx.end(), [&](int64_t a, int64_t b) {
      if (in->reference_index[a] != in->reference_index[b]) return in->reference_index[a] < in->reference_index[b];
      return in->alignment_start[a] < in->alignment_start[b];
    });
  // 3. offsets of the kept reads' bases and CIGAR words; which reads are "one M over all their bases"
  Uninit<uint64_t> base_off;
  base_off.alloc((size_t)n + 1);
  Uninit<int64_t> wire_off;
  wire_off.alloc((size_t)n + 1);
  Uninit<uint8_t> implicit;
  implicit.alloc((size_t)n + 1);
  base_off[(size_t)n] = 0;  // (n = 0: nothing below writes it)
  wire_off[(size_t)n] = 0;
  const int64_t len0 = n ? (int64_t)(in->seq_off[idx[0] + 1] - in->seq_off[idx[0]]) : 0;
  // per chunk: lengths, flags and the chunk's totals; then the offsets from the totals' prefix (two passes, all threads)
  struct ChunkTotals {
    uint64_t bases = 0;
    int64_t words = 0, wire_words = 0;
    bool uniform = true, ops_fit8 = true, too_long = false;
  };
  std::vector<ChunkTotals> chunk((size_t)n_threads);
  std::vector<std::pair<int64_t, int64_t>> bounds((size_t)n_threads, {0, 0});
  host_parallel_for(n_threads, n, [&](int64_t lo, int64_t hi, int t) {
    ChunkTotals c;
    for (int64_t k = lo; k < hi; ++k) {
      const int64_t i = idx[(size_t)k];
      const int64_t len = (int64_t)(in->seq_off[i + 1] - in->seq_off[i]);
      const int64_t ops = (int64_t)in->cigar_off[i + 1] - (int64_t)in->cigar_off[i];
      c.too_long = c.too_long || len > 65535 || ops > 65535;
      c.uniform = c.uniform && len == len0;
      c.ops_fit8 = c.ops_fit8 && ops >= 1 && ops <= 255;
      implicit[(size_t)k] = ops == 1 && in->cigar[in->cigar_off[i]] == ((uint32_t)len << 4) && len < (1 << 28);
      c.bases += (uint64_t)len;
      c.words += ops;
      c.wire_words += implicit[(size_t)k] ? 0 : ops;
    }
    chunk[(size_t)t] = c;
    bounds[(size_t)t] = {lo, hi};
  });
  bool uniform = n > 0, ops_fit8 = true;
  for (auto& c : chunk) {
    if (c.too_long) return failp(GUAC_E_UNSUPPORTED, "a read longer than 65535 bases or CIGAR operators (wire-form limit)");
    uniform = uniform && c.uniform;
    ops_fit8 = ops_fit8 && c.ops_fit8;
  }
  const bool narrow_ops = ops_fit8 && n > 0;
  {
    std::vector<uint64_t> b0((size_t)n_threads + 1, 0);
    std::vector<int64_t> w0((size_t)n_threads + 1, 0);
    for (int t = 0; t < n_threads; ++t) {
      b0[(size_t)t + 1] = b0[(size_t)t] + chunk[(size_t)t].bases;
      w0[(size_t)t + 1] = w0[(size_t)t] + (narrow_ops ? chunk[(size_t)t].wire_words : chunk[(size_t)t].words);
    }
    host_parallel_for(n_threads, n, [&](int64_t lo, int64_t hi, int t) {
      // (same chunking as above: host_parallel_for cuts [0, n) the same way for the same n and thread count)
      uint64_t b = b0[(size_t)t];
      int64_t w = w0[(size_t)t];
      for (int64_t k = lo; k < hi; ++k) {
        const int64_t i = idx[(size_t)k];
        base_off[(size_t)k] = b;
        wire_off[(size_t)k] = w;
        b += in->seq_off[i + 1] - in->seq_off[i];
        w += (narrow_ops && implicit[(size_t)k]) ? 0 : (int64_t)in->cigar_off[i + 1] - (int64_t)in->cigar_off[i];
      }
      if (hi == n) {
        base_off[(size_t)n] = b;
        wire_off[(size_t)n] = w;
      }
    });
  }
  const uint64_t n_bases = base_off[(size_t)n];
  // contig runs and the narrow start column
  std::vector<int32_t> run_contig;
  std::vector<int64_t> run_first;
  {
    std::vector<std::vector<int64_t>> starts((size_t)n_threads);  // run starts found by each chunk, in order
    host_parallel_for(n_threads, n, [&](int64_t lo, int64_t hi, int t) {
      for (int64_t k = lo; k < hi; ++k)
        if (k == 0 || in->reference_index[idx[(size_t)k]] != in->reference_index[idx[(size_t)k - 1]]) starts[(size_t)t].push_back(k);
    });
    for (auto& v : starts)
      for (int64_t k : v) {
        run_contig.push_back(in->reference_index[idx[(size_t)k]]);
        run_first.push_back(k);
      }
  }
  run_first.push_back(n);
  bool narrow_start = n > 0;
  for (size_t r = 0; r + 1 < run_first.size() && narrow_start; ++r)
    narrow_start = in->alignment_start[idx[(size_t)run_first[r + 1] - 1]] - in->alignment_start[idx[(size_t)run_first[r]]] < ((int64_t)1 << 32);
  // 4. quality table: at most 4 / 16 distinct values travel as 2- / 4-bit codes
  std::vector<std::vector<uint8_t>> seen((size_t)n_threads, std::vector<uint8_t>(256, 0));
  host_parallel_for(n_threads, n, [&](int64_t lo, int64_t hi, int t) {
    for (int64_t k = lo; k < hi; ++k) {
      const int64_t i = idx[(size_t)k];
      uint8_t* st = seen[(size_t)t].data();
      const uint8_t* q = in->base_qualities + in->seq_off[i];
      const uint8_t* qe = in->base_qualities + in->seq_off[i + 1];
      for (; q + 8 <= qe; q += 8) {
        st[q[0]] = 1, st[q[1]] = 1, st[q[2]] = 1, st[q[3]] = 1;
        st[q[4]] = 1, st[q[5]] = 1, st[q[6]] = 1, st[q[7]] = 1;
      }
      for (; q < qe; ++q) st[*q] = 1;
    }
  });
  int n_values = 0;
  uint8_t code_of[256];
  std::memset(code_of, 0, sizeof code_of);
  std::memset(out->qual_table, 0, sizeof out->qual_table);
  for (int q = 0; q < 256; ++q) {
    bool any = false;
    for (auto& s : seen) any = any || s[(size_t)q];
    if (any) {
      if (n_values < 16) {
        out->qual_table[n_values] = (uint8_t)q;
        code_of[q] = (uint8_t)n_values;
      }
      ++n_values;
    }
  }
  const int qual_bits = (n_values >= 1 && n_values <= 4) ? 2 : (n_values <= 16 && n_values >= 1) ? 4 : 8;
  const int per_byte = 8 / qual_bits;
  if (qual_bits == 8) std::memset(out->qual_table, 0, sizeof out->qual_table);  // (no table: one byte per base)
  // 4b. reference-relative bases (guac_read_batch_compact::mm_count): which reads travel as differences, how many words each
  const bool ref_rel = n_refs > 0 && n > 0;
  Uninit<uint8_t> mm_count;
  Uninit<uint64_t> expl_off;  // explicit reads' bases, concatenated: offsets over the kept reads (0 bases for the others)
  Uninit<int64_t> mm_off;
  uint8_t base_code[256];
  std::memset(base_code, 4, sizeof base_code);
  base_code['A'] = 0, base_code['C'] = 1, base_code['G'] = 2, base_code['T'] = 3;
  auto window_of = [&](int32_t contig) -> const guac_host_reference* {
    for (int r = 0; r < n_refs; ++r)
      if (refs[r].contig_id == contig) return &refs[r];
    return nullptr;
  };
  // one bit per base of every window: not A C G T (or missing).  A read on clean reference is compared eight bases at a time.
  std::vector<std::vector<uint64_t>> ref_bad((size_t)n_refs);
  for (int r = 0; r < n_refs && n > 0; ++r) {
    const int64_t len = refs[r].len;
    ref_bad[(size_t)r].assign((size_t)(len / 64) + 2, 0);
    uint64_t* bits = ref_bad[(size_t)r].data();
    const uint8_t* rb = refs[r].bases;
    host_parallel_for(n_threads, (len + 63) / 64, [&](int64_t lo, int64_t hi, int) {
      for (int64_t w = lo; w < hi; ++w) {
        uint64_t v = 0;
        const int64_t end = std::min<int64_t>(len, 64 * w + 64);
        for (int64_t p = 64 * w; p < end; ++p) v |= (uint64_t)(base_code[rb[p]] >= 4) << (p & 63);
        bits[w] = v;
      }
    });
  }
  // differences of kept read k from its contig's window, or -1 when it has to travel explicitly; `emit` sees each word
  auto differences = [&](int64_t k, auto&& emit) -> int {
    const int64_t i = idx[(size_t)k];
    const int64_t len = (int64_t)(in->seq_off[i + 1] - in->seq_off[i]);
    if ((int64_t)in->cigar_off[i + 1] - (int64_t)in->cigar_off[i] != 1 || len >= 16384 || len == 0) return -1;
    const uint32_t w = in->cigar[in->cigar_off[i]];
    const int op = (int)(w & 15);
    if (!(op == GUAC_CIGAR_M || op == GUAC_CIGAR_EQ || op == GUAC_CIGAR_X) || (int64_t)(w >> 4) != len) return -1;
    const guac_host_reference* win = window_of(in->reference_index[i]);
    if (!win) return -1;
    const int64_t rel = in->alignment_start[i] - 1 - win->offset;
    if (rel < 0 || rel + len > win->len) return -1;
    {  // every reference base under the read is A C G T
      const uint64_t* bits = ref_bad[(size_t)(win - refs)].data();
      const int64_t w0 = rel >> 6, w1 = (rel + len - 1) >> 6;
      for (int64_t q = w0; q <= w1; ++q) {
        uint64_t v = bits[q];
        if (q == w0) v &= ~0ull << (rel & 63);
        if (q == w1 && ((rel + len) & 63)) v &= ~0ull >> (64 - ((rel + len) & 63));
        if (v) return -1;
      }
    }
    const uint8_t* rb = win->bases + rel;
    const uint8_t* sb = in->bases_ascii + in->seq_off[i];
    int cnt = 0;
    auto one = [&](int64_t p) {  // false: the read has to travel explicitly
      const int bc = base_code[sb[p]];
      if (bc >= 4) return false;
      const int rc = base_code[rb[p]];
      if (rc != bc) {
        if (++cnt >= 255) return false;
        emit((uint16_t)((p << 2) | (rc ^ bc)));
      }
      return true;
    };
    int64_t p = 0;
    for (; p + 8 <= len; p += 8) {  // equal to clean reference: nothing to look at
      uint64_t a, b;
      std::memcpy(&a, rb + p, 8);
      std::memcpy(&b, sb + p, 8);
      if (a == b) continue;
      for (int64_t q = p; q < p + 8; ++q)
        if (!one(q)) return -1;
    }
    for (; p < len; ++p)
      if (!one(p)) return -1;
    return cnt;
  };
  std::vector<std::vector<uint16_t>> thread_words((size_t)n_threads);
  std::vector<int64_t> cw((size_t)n_threads + 1, 0);
  if (ref_rel) {
    mm_count.alloc((size_t)n + 1);
    expl_off.alloc((size_t)n + 1);
    mm_off.alloc((size_t)n + 1);
    std::vector<uint64_t> cb((size_t)n_threads + 1, 0);
    host_parallel_for(n_threads, n, [&](int64_t lo, int64_t hi, int t) {
      uint64_t eb = 0;
      std::vector<uint16_t>& words = thread_words[(size_t)t];  // (the reads of a thread are consecutive: so are their words)
      for (int64_t k = lo; k < hi; ++k) {
        const size_t before = words.size();
        const int c = differences(k, [&](uint16_t word) { words.push_back(word); });
        mm_count[(size_t)k] = c < 0 ? (uint8_t)255 : (uint8_t)c;
        if (c < 0) {
          words.resize(before);
          eb += base_off[(size_t)k + 1] - base_off[(size_t)k];
        }
      }
      cb[(size_t)t + 1] = eb;
      cw[(size_t)t + 1] = (int64_t)words.size();
    });
    for (int t = 0; t < n_threads; ++t) {
      cb[(size_t)t + 1] += cb[(size_t)t];
      cw[(size_t)t + 1] += cw[(size_t)t];
    }
    host_parallel_for(n_threads, n, [&](int64_t lo, int64_t hi, int t) {
      uint64_t eb = cb[(size_t)t];
      int64_t ew = cw[(size_t)t];
      for (int64_t k = lo; k < hi; ++k) {
        expl_off[(size_t)k] = eb;
        mm_off[(size_t)k] = ew;
        if (mm_count[(size_t)k] == 255) eb += base_off[(size_t)k + 1] - base_off[(size_t)k];
        else ew += mm_count[(size_t)k];
      }
      if (hi == n) {
        expl_off[(size_t)n] = eb;
        mm_off[(size_t)n] = ew;
      }
    });
  }
  const uint64_t n_explicit = ref_rel ? expl_off[(size_t)n] : n_bases;
  // 5. outputs
  PackAlloc A{ctx};
  auto* o_run_contig = static_cast<int32_t*>(A.get(run_contig.size() * sizeof(int32_t)));
  auto* o_run_first = static_cast<int64_t*>(A.get(run_first.size() * sizeof(int64_t)));
  auto* o_start = narrow_start ? nullptr : static_cast<int64_t*>(A.get((size_t)n * sizeof(int64_t)));
  auto* o_run_base = narrow_start ? static_cast<int64_t*>(A.get(run_contig.size() * sizeof(int64_t))) : nullptr;
  auto* o_start_rel = narrow_start ? static_cast<uint32_t*>(A.get((size_t)n * sizeof(uint32_t))) : nullptr;
  auto* o_read_len = uniform ? nullptr : static_cast<uint16_t*>(A.get((size_t)n * sizeof(uint16_t)));
  auto* o_ops16 = narrow_ops ? nullptr : static_cast<uint16_t*>(A.get((size_t)n * sizeof(uint16_t)));
  auto* o_ops8 = narrow_ops ? static_cast<uint8_t*>(A.get((size_t)n)) : nullptr;
  auto* o_cigar = static_cast<uint32_t*>(A.get((size_t)wire_off[(size_t)n] * sizeof(uint32_t)));
  auto* o_bases2 = static_cast<uint8_t*>(A.get((size_t)((n_explicit + 3) / 4) + 8));
  auto* o_mm_count = ref_rel ? static_cast<uint8_t*>(A.get((size_t)n)) : nullptr;
  auto* o_mm_words = ref_rel ? static_cast<uint16_t*>(A.get((size_t)mm_off[(size_t)n] * sizeof(uint16_t) + 8)) : nullptr;
  auto* o_qual = static_cast<uint8_t*>(A.get((size_t)((n_bases + per_byte - 1) / per_byte) + 8));
  auto* o_mapq = static_cast<uint8_t*>(A.get((size_t)n));
  auto* o_flags = static_cast<uint8_t*>(A.get((size_t)n));
  auto* o_sample = static_cast<uint8_t*>(A.get((size_t)n));
  if (!A.ok) {
    A.release_all();
    return failp(GUAC_E_OOM, "guac_pack_reads: out of host memory");
  }
  for (size_t r = 0; r < run_contig.size(); ++r) {
    o_run_contig[r] = run_contig[r];
    if (o_run_base) o_run_base[r] = in->alignment_start[idx[(size_t)run_first[r]]] - 1;
  }
  for (size_t r = 0; r < run_first.size(); ++r) o_run_first[r] = run_first[r];
  // per-read columns
  host_parallel_for(n_threads, n, [&](int64_t lo, int64_t hi, int) {
    size_t run = (size_t)(std::upper_bound(run_first.begin(), run_first.end(), lo) - run_first.begin()) - 1;
    for (int64_t k = lo; k < hi; ++k) {
      while (run + 1 < run_first.size() && run_first[run + 1] <= k) ++run;
      const int64_t i = idx[(size_t)k];
      const uint16_t fl = in->sam_flags[i];
      const int64_t start = in->alignment_start[i] - 1;  // samtools is 1-based (Read.scala:84)
      if (o_start) o_start[k] = start;
      if (o_start_rel) o_start_rel[k] = (uint32_t)(start - o_run_base[run]);
      if (o_read_len) o_read_len[k] = (uint16_t)(in->seq_off[i + 1] - in->seq_off[i]);
      const int64_t ops = (int64_t)in->cigar_off[i + 1] - (int64_t)in->cigar_off[i];
      if (o_ops16) o_ops16[k] = (uint16_t)ops;
      if (o_ops8) o_ops8[k] = implicit[(size_t)k] ? 0 : (uint8_t)ops;
      if (!(narrow_ops && implicit[(size_t)k]))
        for (int64_t c = 0; c < ops; ++c) o_cigar[wire_off[(size_t)k] + c] = in->cigar[in->cigar_off[i] + c];
      o_mapq[k] = in->mapping_quality[i];
      o_flags[k] = (uint8_t)(((fl & SAM_REVERSE) ? 0 : GUAC_READ_POSITIVE_STRAND) | ((fl & SAM_PAIRED) ? GUAC_READ_PAIRED : 0) |
                             ((fl & SAM_DUPLICATE) ? GUAC_READ_DUPLICATE : 0) | ((fl & SAM_FAILED_QC) ? GUAC_READ_FAILED_QC : 0));
      o_sample[k] = in->sample_id[i];
    }
  });
  // bases and qualities: every thread owns a range of OUTPUT bytes (global base indices [g0, g1), multiples of 8), so
  // no two threads ever touch the same byte
  std::vector<std::vector<std::pair<uint64_t, uint8_t>>> exc((size_t)n_threads);
  uint8_t base_lut[256];
  std::memset(base_lut, 4, sizeof base_lut);
  base_lut['A'] = 0;
  base_lut['C'] = 1;
  base_lut['G'] = 2;
  base_lut['T'] = 3;
  const int64_t n_groups = (int64_t)((n_bases + 7) / 8);
  host_parallel_for(n_threads, n_groups, [&](int64_t glo, int64_t ghi, int t) {
    const uint64_t g0 = (uint64_t)glo * 8, g1 = std::min<uint64_t>(n_bases, (uint64_t)ghi * 8);
    if (g0 >= g1) return;
    if (!ref_rel) std::memset(o_bases2 + g0 / 4, 0, (size_t)((g1 + 3) / 4 - g0 / 4));
    std::memset(o_qual + g0 / per_byte, 0, (size_t)((g1 + per_byte - 1) / per_byte - g0 / per_byte));
    int64_t k = (int64_t)(std::upper_bound(base_off.begin(), base_off.end(), g0) - base_off.begin()) - 1;
    const int qshift = qual_bits == 2 ? 2 : qual_bits == 4 ? 1 : 0;  // log2(bases per quality byte)
    const unsigned qmask = (1u << qshift) - 1;
    for (uint64_t g = g0; g < g1;) {
      while (base_off[(size_t)k + 1] <= g) ++k;
      const int64_t i = idx[(size_t)k];
      const uint64_t end = std::min<uint64_t>(g1, base_off[(size_t)k + 1]);
      const uint8_t* sb = in->bases_ascii + in->seq_off[i] + (g - base_off[(size_t)k]);
      const uint8_t* sq = in->base_qualities + in->seq_off[i] + (g - base_off[(size_t)k]);
      // a read that travels as differences has A C G T bases only (no exception to find) and no bases to pack
      const bool skip_bases = ref_rel && mm_count[(size_t)k] != 255;
      auto one = [&]() {  // one base, read-modify-write of its output bytes
        const int code = skip_bases ? 0 : base_lut[*sb];
        if (code >= 4) exc[(size_t)t].push_back({g, *sb});
        else if (!ref_rel) o_bases2[g >> 2] |= (uint8_t)(code << ((g & 3) * 2));
        if (qual_bits == 8) o_qual[g] = *sq;
        else o_qual[g >> qshift] |= (uint8_t)(code_of[*sq] << (qual_bits * (int)(g & qmask)));
        ++g, ++sb, ++sq;
      };
      while (g < end && (g & 3)) one();
      for (; g + 4 <= end; g += 4, sb += 4, sq += 4) {  // four bases per output byte, plain stores
        const int c0 = skip_bases ? 0 : base_lut[sb[0]], c1 = skip_bases ? 0 : base_lut[sb[1]], c2 = skip_bases ? 0 : base_lut[sb[2]],
                  c3 = skip_bases ? 0 : base_lut[sb[3]];
        if ((c0 | c1 | c2 | c3) < 4) {
          if (!ref_rel) o_bases2[g >> 2] = (uint8_t)(c0 | (c1 << 2) | (c2 << 4) | (c3 << 6));
        } else {
          for (int j = 0; j < 4; ++j) {
            const int c = base_lut[sb[j]];
            if (c >= 4) exc[(size_t)t].push_back({g + (uint64_t)j, sb[j]});
            else if (!ref_rel) o_bases2[g >> 2] |= (uint8_t)(c << (2 * j));
          }
        }
        if (qual_bits == 2) {
          o_qual[g >> 2] = (uint8_t)(code_of[sq[0]] | (code_of[sq[1]] << 2) | (code_of[sq[2]] << 4) | (code_of[sq[3]] << 6));
        } else if (qual_bits == 4) {
          o_qual[g >> 1] = (uint8_t)(code_of[sq[0]] | (code_of[sq[1]] << 4));
          o_qual[(g >> 1) + 1] = (uint8_t)(code_of[sq[2]] | (code_of[sq[3]] << 4));
        } else {
          std::memcpy(o_qual + g, sq, 4);
        }
      }
      while (g < end) one();
    }
  });
  if (ref_rel) {
    // the difference words, and the explicit reads' bases: every thread owns a range of output bytes of the explicit stream
    host_parallel_for(n_threads, n, [&](int64_t lo, int64_t hi, int t) {
      for (int64_t k = lo; k < hi; ++k) o_mm_count[k] = mm_count[(size_t)k];
      const std::vector<uint16_t>& words = thread_words[(size_t)t];
      if (!words.empty()) std::memcpy(o_mm_words + cw[(size_t)t], words.data(), words.size() * sizeof(uint16_t));
    });
    const int64_t e_groups = (int64_t)((n_explicit + 7) / 8);
    host_parallel_for(n_threads, e_groups, [&](int64_t glo, int64_t ghi, int) {
      const uint64_t g0 = (uint64_t)glo * 8, g1 = std::min<uint64_t>(n_explicit, (uint64_t)ghi * 8);
      if (g0 >= g1) return;
      std::memset(o_bases2 + g0 / 4, 0, (size_t)((g1 + 3) / 4 - g0 / 4));
      int64_t k = (int64_t)(std::upper_bound(expl_off.begin(), expl_off.end(), g0) - expl_off.begin()) - 1;
      for (uint64_t g = g0; g < g1;) {
        while (expl_off[(size_t)k + 1] <= g) ++k;
        const int64_t i = idx[(size_t)k];
        const uint64_t end = std::min<uint64_t>(g1, expl_off[(size_t)k + 1]);
        const uint8_t* sb = in->bases_ascii + in->seq_off[i] + (g - expl_off[(size_t)k]);
        for (; g < end; ++g, ++sb) {
          const int code = base_lut[*sb];
          if (code < 4) o_bases2[g >> 2] |= (uint8_t)(code << ((g & 3) * 2));
        }
      }
    });
  }
  size_t n_exc = 0;
  for (auto& v : exc) n_exc += v.size();
  auto* o_exc_pos = static_cast<uint64_t*>(A.get(n_exc * sizeof(uint64_t)));
  auto* o_exc_base = static_cast<uint8_t*>(A.get(n_exc));
  if (!A.ok) {
    A.release_all();
    return failp(GUAC_E_OOM, "guac_pack_reads: out of host memory");
  }
  {
    size_t w = 0;
    for (auto& v : exc)
      for (auto& e : v) {
        o_exc_pos[w] = e.first;
        o_exc_base[w++] = e.second;
      }
  }
  out->n_reads = n;
  out->n_samples = in->n_samples;
  out->n_contig_runs = (int32_t)run_contig.size();
  out->run_contig_id = o_run_contig;
  out->run_first_read = o_run_first;
  out->start = o_start;
  out->read_len = o_read_len;
  out->n_cigar_ops = o_ops16;
  out->cigar = o_cigar;
  out->bases2 = o_bases2;
  out->n_exc = (int64_t)n_exc;
  out->exc_pos = o_exc_pos;
  out->exc_base = o_exc_base;
  out->qual_bits = qual_bits;
  out->qual = o_qual;
  out->mapq = o_mapq;
  out->flags = o_flags;
  out->sample_id = o_sample;
  out->run_start_base = o_run_base;
  out->start_rel = o_start_rel;
  out->n_cigar_ops8 = o_ops8;
  out->uniform_read_len = uniform ? (int32_t)len0 : 0;
  if (ref_rel) {
    out->mm_count = o_mm_count;
    out->mm_words = o_mm_words;
    out->n_mm_words = mm_off[(size_t)n];
    out->n_explicit_bases = (int64_t)n_explicit;
  }
  if (n_kept) *n_kept = n;
  return GUAC_OK;
}
}  // namespace

int guac_pack_reads(guac_ctx* ctx, const guac_sam_records* in, const guac_read_filters* filters, int32_t n_threads,
                    guac_read_batch_compact* out, int64_t* n_kept) {
  return pack_reads_impl(ctx, in, filters, nullptr, 0, n_threads, out, n_kept);
}

int guac_pack_reads_ref(guac_ctx* ctx, const guac_sam_records* in, const guac_read_filters* filters, const guac_host_reference* refs,
                        int32_t n_refs, int32_t n_threads, guac_read_batch_compact* out, int64_t* n_kept) {
  return pack_reads_impl(ctx, in, filters, refs, n_refs, n_threads, out, n_kept);
}

int guac_ss_genotype_filter(const guac_ss_filter_params* P, const guac_ss_call* calls, int64_t n, uint8_t* keep) {
  if (!P || n < 0 || (n > 0 && (!calls || !keep))) return GUAC_E_INVALID_INPUT;
  for (int64_t i = 0; i < n; ++i) {
    const guac_allele_evidence &t = calls[i].tumor_variant_evidence, &nr = calls[i].normal_reference_evidence;
    bool k = t.read_depth >= P->min_tumor_read_depth && t.read_depth < P->max_tumor_read_depth;  // SomaticReadDepthFilter
    k = k && nr.read_depth >= P->min_normal_read_depth;
    if (P->min_tumor_alternate_read_depth >= 0) k = k && t.allele_read_depth >= P->min_tumor_alternate_read_depth;
    if (P->min_likelihood >= 0) k = k && phred_scaled_somatic_likelihood(calls[i]) >= P->min_likelihood;
    const float vaf = (float)t.allele_read_depth / (float)t.read_depth;  // Float division (AlleleEvidence.scala:37)
    k = k && (double)vaf * 100.0 > (double)P->min_vaf;
    // the "average base quality" filter tests meanMappingQuality too (SomaticAverageBaseQualityFilter.scala:11-12)
    for (int thr : {P->min_average_mapping_quality, P->min_average_base_quality})
      k = k && t.mean_mapping_quality >= thr && nr.mean_mapping_quality >= thr;
    if (P->maximum_median_mismatches >= 0) k = k && t.median_mismatches_per_read <= P->maximum_median_mismatches;
    keep[i] = k ? 1 : 0;
  }
  return GUAC_OK;
}

int guac_ss_vcf_rows(const guac_ss_call* calls, int64_t n, const uint8_t* keep, const char* arena, const char* const* contig_names,
                     const char* sample_name, char* out_text, int64_t cap, int64_t* needed) {
  if (n < 0 || (n > 0 && (!calls || !arena || !contig_names)) || !sample_name) return GUAC_E_INVALID_INPUT;
  std::vector<int64_t> order;
  for (int64_t i = 0; i < n; ++i)
    if (!keep || keep[i]) order.push_back(i);
  std::stable_sort(order.begin(), order.end(), [&](int64_t a, int64_t b) { return calls[a].start < calls[b].start; });  // GenotypeOutputArgs.scala:96-98
  std::string text = std::string("#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\t") + sample_name + "\n";
  for (int64_t i : order) {
    const guac_ss_call& c = calls[i];
    const int dp = c.tumor_variant_evidence.read_depth, ad = c.tumor_variant_evidence.allele_read_depth;
    const int gq = std::min(99, phred_scaled_somatic_likelihood(c));
    text += std::string(contig_names[c.contig_id]) + "\t" + std::to_string(c.start + 1) + "\t.\t" + std::string(arena + c.ref_off, c.ref_len) + "\t" +
            std::string(arena + c.alt_off, c.alt_len) + "\t.\t.\t.\tGT:AD:DP:GQ\t0/1:" + std::to_string(dp - ad) + "," + std::to_string(ad) + ":" +
            std::to_string(dp) + ":" + std::to_string(gq) + "\n";
  }
  return copy_text_out(text, out_text, cap, needed);
}

int guac_sj_vcf_rows(const guac_sj_params* P, const guac_sj_allele_evidence* alleles, int64_t n_alleles, const guac_sj_sample_evidence* samples,
                     const char* arena, const char* const* contig_names, const char* const* sample_names, int32_t include_pooled_normal,
                     int32_t include_pooled_tumor, int32_t somatic_genotype_policy, char* out_text, int64_t cap, int64_t* needed) {
  if (!P || n_alleles < 0 || (n_alleles > 0 && (!alleles || !samples || !arena)) || !contig_names || !sample_names) return GUAC_E_INVALID_INPUT;
  const int S = P->n_samples;
  int n_normal = 0, n_tumor = 0;
  for (int s = 0; s < S; ++s) {
    n_normal += P->sample_kind[s] == GUAC_SJ_NORMAL_DNA;
    n_tumor += P->sample_kind[s] == GUAC_SJ_TUMOR_DNA;
  }
  const bool ipn = include_pooled_normal < 0 ? n_normal > 1 : include_pooled_normal != 0;  // SomaticJointCaller.scala:223
  const bool ipt = include_pooled_tumor < 0 ? n_tumor > 1 : include_pooled_tumor != 0;
  struct Column {
    int idx, kind;
  };
  std::vector<Column> columns;
  for (int s = 0; s < S; ++s) columns.push_back({s, P->sample_kind[s]});
  if (ipn) columns.push_back({S, GUAC_SJ_NORMAL_DNA});
  if (ipt) columns.push_back({S + 1, GUAC_SJ_TUMOR_DNA});
  std::string text = "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT";
  for (int s = 0; s < S; ++s) text += std::string("\t") + sample_names[s];
  if (ipn) text += "\tpooled_normal";
  if (ipt) text += "\tpooled_tumor";
  text += "\n";
  std::vector<int64_t> order((size_t)n_alleles);
  for (int64_t i = 0; i < n_alleles; ++i) order[(size_t)i] = i;
  std::stable_sort(order.begin(), order.end(), [&](int64_t a, int64_t b) {  // VCFOutput.scala:97-100: by (contig name, start)
    const int c = std::strcmp(contig_names[alleles[a].contig_id], contig_names[alleles[b].contig_id]);
    return c != 0 ? c < 0 : alleles[a].start < alleles[b].start;
  });
  auto str = [&](uint32_t off, uint32_t len) { return std::string(arena + off, len); };
  auto shown = [](const std::string& a) { return a.empty() ? std::string(".") : a; };
  for (int64_t ai : order) {
    const guac_sj_allele_evidence& ev = alleles[ai];
    const guac_sj_sample_evidence* se_all = samples + ev.first_sample_evidence;
    const std::string ref = str(ev.ref_off, ev.ref_len), alt = str(ev.alt_off, ev.alt_len);
    auto depth_of = [&](const guac_sj_sample_evidence& se, const std::string& a) {
      for (int k = 0; k < se.n_depths; ++k)
        if (str(se.depths[k].seq_off, se.depths[k].seq_len) == a) return se.depths[k].count;
      return 0;
    };
    auto sample_depth = [&](const guac_sj_sample_evidence& se) {  // allelicDepths.values.sum of the truncated map
      int d = 0;
      for (int k = 0; k < se.n_depths; ++k) d += se.depths[k].count;
      return d;
    };
    bool any_ff = false;
    for (auto& col : columns) any_ff = any_ff || se_all[col.idx].sb_filtered;
    std::vector<std::string> keys = {"GT", "AD", "ADP", "DP"};
    if (any_ff) keys.push_back("FF");
    for (const char* k : {"FS", "RL", "TRIGGERED", "VAF"}) keys.push_back(k);
    std::vector<std::string> cells;
    for (auto& col : columns) {
      const guac_sj_sample_evidence& se = se_all[col.idx];
      std::string a1, a2, rl, triggered;
      if (col.kind == GUAC_SJ_NORMAL_DNA) {
        const std::pair<std::string, std::string> names[3] = {{ref, ref}, {ref, alt}, {alt, alt}};
        int best = 0;  // posteriors.maxBy(_._2): the first strictly greater entry wins
        for (int g = 1; g < 3; ++g)
          if (se.posteriors[g] > se.posteriors[best]) best = g;
        a1 = names[best].first;
        a2 = names[best].second;
        for (int g = 0; g < 3; ++g) rl += std::string(g ? "," : "") + "[" + names[g].first + "/" + names[g].second + "]=" + java_g8(se.posteriors[g]);
        triggered = ".";
      } else {
        if (somatic_genotype_policy == 0) {  // presence: the two most frequent alleles seen
          std::vector<std::pair<std::string, int>> present;
          for (int k = 0; k < se.n_depths; ++k)
            if (se.depths[k].count > 0) present.push_back({str(se.depths[k].seq_off, se.depths[k].seq_len), se.depths[k].count});
          std::stable_sort(present.begin(), present.end(), [](const auto& x, const auto& y) { return x.second > y.second; });
          if (present.size() > 2) present.resize(2);
          bool has_ref = false, has_alt = false, other = false;
          for (auto& p : present) {
            if (p.first == ref) has_ref = true;
            else if (p.first == alt) has_alt = true;
            else other = true;
          }
          if (has_ref && has_alt && present.size() == 2) {
            a1 = ref;
            a2 = alt;
          } else if (present.size() == 1 && (has_ref || has_alt) && !other) {
            a1 = a2 = has_ref ? ref : alt;
          } else {
            a1 = a2 = ref;
          }
        } else {
          a1 = ref;
          a2 = se.triggered ? alt : ref;
        }
        rl = "[" + shown(ref) + "->" + java_fixed(1.0, 2) + "]=" + java_g8(se.posteriors[0]);
        if (ref != alt)
          rl += ",[" + shown(alt) + "->" + java_fixed(se.alt_vaf, 2) + "|" + shown(ref) + "->" + java_fixed(1.0 - se.alt_vaf, 2) + "]=" +
                java_g8(se.posteriors[1]);
        triggered = se.triggered ? (col.kind == GUAC_SJ_TUMOR_RNA ? "EXPRESSED" : "YES") : "NO";
      }
      std::string g1 = a1 == ref ? "0" : "1", g2 = a2 == ref ? "0" : "1";
      if (g2 < g1) std::swap(g1, g2);
      // ADP: ref, alt, then the sample's other alleles in depth order, each once
      std::vector<std::string> adp_alleles = {ref};
      if (alt != ref) adp_alleles.push_back(alt);
      for (int k = 0; k < se.n_depths; ++k) {
        const std::string a = str(se.depths[k].seq_off, se.depths[k].seq_len);
        if (std::find(adp_alleles.begin(), adp_alleles.end(), a) == adp_alleles.end()) adp_alleles.push_back(a);
      }
      std::string adp;
      for (size_t k = 0; k < adp_alleles.size(); ++k) adp += std::string(k ? "," : "") + shown(adp_alleles[k]) + "=" + std::to_string(depth_of(se, adp_alleles[k]));
      const int sd = sample_depth(se);
      const double vaf = sd ? (double)depth_of(se, alt) / sd : std::nan("");
      std::string cell;
      for (size_t k = 0; k < keys.size(); ++k) {
        std::string v;
        if (keys[k] == "GT") v = g1 + "/" + g2;
        else if (keys[k] == "AD") v = std::to_string(depth_of(se, ref)) + "," + std::to_string(depth_of(se, alt));
        else if (keys[k] == "ADP") v = adp;
        else if (keys[k] == "DP") v = std::to_string(sd);
        else if (keys[k] == "FF") v = se.sb_filtered ? "STRAND_BIAS" : ".";
        else if (keys[k] == "FS") v = std::to_string((long long)std::floor(se.sb_phred + 0.5));
        else if (keys[k] == "RL") v = rl;
        else if (keys[k] == "TRIGGERED") v = triggered;
        else v = format_vcf_double(vaf);
        cell += (k ? ":" : "") + v;
      }
      cells.push_back(cell);
    }
    std::string triggers;
    if (ev.is_germline_call) {
      triggers = "GERMLINE_POOLED";
    } else {
      if (se_all[S + 1].triggered) triggers = "SOMATIC_POOLED";
      bool individual = false;
      for (int s = 0; s < S; ++s) individual = individual || (P->sample_kind[s] == GUAC_SJ_TUMOR_DNA && se_all[s].triggered);
      if (individual) triggers += std::string(triggers.empty() ? "" : ",") + "SOMATIC_INDIVIDUAL";
    }
    std::string flt;
    if (ev.failing_filters & 2u) flt = "INSUFFICIENT_NORMAL";
    if (ev.failing_filters & 1u) flt += std::string(flt.empty() ? "" : ";") + "STRAND_BIAS";
    if (flt.empty()) flt = (ev.is_germline_call || ev.is_somatic_call) ? "PASS" : ".";
    std::string format;
    for (size_t k = 0; k < keys.size(); ++k) format += (k ? ":" : "") + keys[k];
    text += std::string(contig_names[ev.contig_id]) + "\t" + std::to_string(ev.start + 1) + "\t.\t" + ref + "\t" + alt + "\t.\t" + flt + "\tTRIGGER=" +
            (triggers.empty() ? "NONE" : triggers) + ";TUMOR_EXPRESSION=" + (ev.tumor_rna_expressed ? "YES" : "NO") + "\t" + format;
    for (auto& c : cells) text += "\t" + c;
    text += "\n";
  }
  return copy_text_out(text, out_text, cap, needed);
}

}  // extern "C"
