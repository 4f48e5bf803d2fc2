// support_entry.inl — guac_variant_support and guac_vaf_histogram (part of the guac_b200.cu translation unit).

extern "C" {

int guac_variant_support(guac_ctx* ctx, const guac_range* loci, int32_t n_ranges, guac_allele_count** out_counts, int64_t* n_out,
                         char** out_arena, int64_t* arena_len) {
  if (!ctx || !out_counts || !n_out || !out_arena || !arena_len) return GUAC_E_INVALID_INPUT;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  Plan plan;
  GUAC_TRY(make_plan(ctx, loci, n_ranges, plan));
  *n_out = 0;
  *arena_len = 0;
  if (plan.n_tiles == 0) {
    GUAC_TRY(host_out(ctx, out_counts, nullptr, 0));
    GUAC_TRY(host_out(ctx, out_arena, nullptr, 0));
    return GUAC_OK;
  }
  const int S = ctx->n_samples;
  GUAC_TRY(reset_error(ctx));
  GUAC_TRY(ensure(ctx, ctx->b_visited, (size_t)plan.n_slots));
  GUAC_TRY(ensure(ctx, ctx->b_ncalls, sizeof(int)));
  GUAC_TRY(pinned_block(ctx));
  int& n = *static_cast<int*>(ctx->h_counters);
  int cap = (int)std::max<size_t>(ctx->g_evs.cap / sizeof(AlleleCountDev), (size_t)std::min<int64_t>(plan.n_slots * 4 + 1024, 1 << 26));
  for (int attempt = 0; attempt < 2; ++attempt) {
    GUAC_TRY(ensure(ctx, ctx->g_evs, (size_t)cap * sizeof(AlleleCountDev)));
    CUDA_TRY(ctx, cudaMemsetAsync(ctx->b_ncalls.p, 0, sizeof(int), ctx->stream));
    allele_counts_kernel<<<(unsigned)plan.n_tiles, CTA_THREADS, 0, ctx->stream>>>(ctx->R, ctx->d_refs.as<RefView>(), ctx->b_tiles.as<Tile>(), S,
                                                                                 ctx->b_visited.as<uint8_t>(), ctx->g_evs.as<AlleleCountDev>(),
                                                                                 ctx->b_ncalls.as<int>(), cap, ctx->b_err.as<DeviceError>());
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    CUDA_TRY(ctx, cudaMemcpyAsync(&n, ctx->b_ncalls.p, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    GUAC_TRY(check_device_error(ctx));  // synchronises
    if (n <= cap) break;
    cap = n;
  }
  const int n_rec = n;
  std::vector<AlleleCountDev> rec((size_t)n_rec);
  if (n_rec) {
    CUDA_TRY(ctx, cudaMemcpyAsync(rec.data(), ctx->g_evs.p, (size_t)n_rec * sizeof(AlleleCountDev), cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  }
  // strings: Allele(referenceBases, sequencedBases) of pileup/Alignment.scala:14-75 per kind
  std::vector<SJStringJob> jobs;
  std::vector<guac_allele_count> res((size_t)n_rec);
  uint32_t arena_off = 0;
  auto add = [&](int kind, int64_t a, int32_t b, int32_t len, uint32_t& off, uint32_t& out_len) {
    off = arena_off;
    out_len = (uint32_t)len;
    if (len <= 0) return;
    SJStringJob j;
    j.a = a;
    j.b = b;
    j.len = len;
    j.arena_off = arena_off;
    j.kind = kind;
    jobs.push_back(j);
    arena_off += (uint32_t)len;
  };
  std::vector<int32_t> read_len_of;  // (insertion alleles are clipped to the read: fetched lazily below)
  for (int i = 0; i < n_rec; ++i) {
    const AlleleCountDev& a = rec[(size_t)i];
    guac_allele_count& g = res[(size_t)i];
    std::memset(&g, 0, sizeof g);
    g.contig_id = a.contig_id;
    g.sample_id = a.sample;
    g.locus = a.locus;
    g.count = a.count;
    switch (a.kind) {
      case GUAC_ALN_MATCH:
        add(1, a.locus, a.ref_index, 1, g.ref_off, g.ref_len);
        add(1, a.locus, a.ref_index, 1, g.alt_off, g.alt_len);
        break;
      case GUAC_ALN_MISMATCH:
        add(1, a.locus, a.ref_index, 1, g.ref_off, g.ref_len);
        add(3, 0, a.base, 1, g.alt_off, g.alt_len);
        break;
      case GUAC_ALN_INSERTION: {  // (sequenced head, anchor + inserted bases); the exact key tells the clipped length
        const int alt_len = (a.key >> 60) == 2 ? (int)((a.key >> 54) & 63) : a.indel_len + 1;
        add(0, a.rep_read, a.rep_rp, 1, g.ref_off, g.ref_len);
        add(0, a.rep_read, a.rep_rp, alt_len, g.alt_off, g.alt_len);
        break;
      }
      case GUAC_ALN_DELETION:
        add(1, a.locus, a.ref_index, a.indel_len + 1, g.ref_off, g.ref_len);
        add(1, a.locus, a.ref_index, 1, g.alt_off, g.alt_len);
        break;
      case GUAC_ALN_MID_DELETION:
        add(1, a.locus, a.ref_index, 1, g.ref_off, g.ref_len);
        add(1, a.locus, a.ref_index, 0, g.alt_off, g.alt_len);
        break;
      default:  // Clipped: Allele("", "")
        add(1, a.locus, a.ref_index, 0, g.ref_off, g.ref_len);
        add(1, a.locus, a.ref_index, 0, g.alt_off, g.alt_len);
    }
  }
  std::vector<char> arena((size_t)arena_off + 1);
  if (!jobs.empty()) {
    GUAC_TRY(upload(ctx, ctx->g_strings, jobs.data(), jobs.size()));
    GUAC_TRY(ensure(ctx, ctx->b_arena, (size_t)arena_off + 16));
    sj_strings_kernel<<<blocks_for((int64_t)jobs.size(), 128), 128, 0, ctx->stream>>>(ctx->R, ctx->d_refs.as<RefView>(),
                                                                                     ctx->g_strings.as<SJStringJob>(), (int)jobs.size(),
                                                                                     ctx->b_arena.as<char>());
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    CUDA_TRY(ctx, cudaMemcpyAsync(arena.data(), ctx->b_arena.p, arena_off, cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  }
  // (range order, locus), sample, then the allele strings: the reference's order within a pileup is a hash map's
  std::vector<int> order((size_t)n_rec);
  for (int i = 0; i < n_rec; ++i) order[(size_t)i] = i;
  auto str = [&](uint32_t off, uint32_t len) { return std::string(arena.data() + off, len); };
  std::sort(order.begin(), order.end(), [&](int x, int y) {
    const AlleleCountDev &a = rec[(size_t)x], &b = rec[(size_t)y];
    if (a.slot != b.slot) return a.slot < b.slot;
    if (a.sample != b.sample) return a.sample < b.sample;
    const guac_allele_count &ga = res[(size_t)x], &gb = res[(size_t)y];
    const std::string ra = str(ga.ref_off, ga.ref_len), rb = str(gb.ref_off, gb.ref_len);
    if (ra != rb) return ra < rb;
    return str(ga.alt_off, ga.alt_len) < str(gb.alt_off, gb.alt_len);
  });
  void* ho = nullptr;
  void* ha = nullptr;
  GUAC_TRY(pinned_alloc(ctx, &ho, std::max<size_t>((size_t)n_rec * sizeof(guac_allele_count), 16)));
  {
    const int st = pinned_alloc(ctx, &ha, std::max<size_t>(arena_off, 16));
    if (st != GUAC_OK) {
      guac_free(ctx, ho);
      return st;
    }
  }
  for (int i = 0; i < n_rec; ++i) static_cast<guac_allele_count*>(ho)[i] = res[(size_t)order[(size_t)i]];
  if (arena_off) std::memcpy(ha, arena.data(), arena_off);
  *out_counts = static_cast<guac_allele_count*>(ho);
  *n_out = n_rec;
  *out_arena = static_cast<char*>(ha);
  *arena_len = arena_off;
  return GUAC_OK;
}

int guac_vaf_histogram(guac_ctx* ctx, const guac_range* loci, int32_t n_ranges, int32_t bins, int32_t min_read_depth,
                       int32_t min_vaf_percent, int64_t* out_hist, int64_t* out_n_variant_loci, guac_variant_locus** out_loci,
                       int64_t* n_loci) {
  if (!ctx || !out_hist || !out_n_variant_loci) return GUAC_E_INVALID_INPUT;
  if (bins < 1 || bins > 100) return fail(ctx, GUAC_E_INVALID_INPUT, "assumption failed: Bins should be between 1 and 100");  // :204
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  Plan plan;
  GUAC_TRY(make_plan(ctx, loci, n_ranges, plan));
  const int S = ctx->n_samples;
  std::memset(out_hist, 0, (size_t)S * 101 * sizeof(int64_t));
  std::memset(out_n_variant_loci, 0, (size_t)S * sizeof(int64_t));
  if (n_loci) *n_loci = 0;
  if (plan.n_tiles == 0) {
    if (out_loci) GUAC_TRY(host_out(ctx, out_loci, nullptr, 0));
    return GUAC_OK;
  }
  GUAC_TRY(reset_error(ctx));
  GUAC_TRY(ensure(ctx, ctx->b_visited, (size_t)plan.n_slots));
  GUAC_TRY(ensure(ctx, ctx->b_counts, (size_t)plan.n_slots * S * sizeof(guac_locus_counts)));
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->b_counts.p, 0, (size_t)plan.n_slots * S * sizeof(guac_locus_counts), ctx->stream));
  pileup_counts_kernel<<<(unsigned)plan.n_tiles, CTA_THREADS, 0, ctx->stream>>>(ctx->R, ctx->d_refs.as<RefView>(), ctx->b_tiles.as<Tile>(), S,
                                                                               ctx->b_visited.as<uint8_t>(), ctx->b_counts.as<guac_locus_counts>(),
                                                                               ctx->b_err.as<DeviceError>());
  ctx->launches += 1;
  CUDA_TRY(ctx, cudaGetLastError());
  const int64_t n_cells = plan.n_slots * S;
  const size_t hist_bytes = ((size_t)S * 101 + S) * sizeof(unsigned long long);
  GUAC_TRY(ensure(ctx, ctx->u_flag, (size_t)n_cells + 16));
  GUAC_TRY(ensure(ctx, ctx->u_tmp, (size_t)n_cells * sizeof(float) + hist_bytes + 64));
  GUAC_TRY(ensure(ctx, ctx->u_off, (size_t)(n_cells + 2) * sizeof(int64_t)));
  float* d_vaf = ctx->u_tmp.as<float>();
  unsigned long long* d_hist = reinterpret_cast<unsigned long long*>(ctx->u_tmp.as<unsigned char>() + (((size_t)n_cells * sizeof(float) + 63) / 64) * 64);
  CUDA_TRY(ctx, cudaMemsetAsync(d_hist, 0, hist_bytes, ctx->stream));
  VafParams VP{S, bins, min_read_depth, min_vaf_percent};
  vaf_loci_kernel<<<blocks_for(n_cells, 256), 256, 0, ctx->stream>>>(ctx->b_counts.as<guac_locus_counts>(), ctx->b_visited.as<uint8_t>(),
                                                                     plan.n_slots, VP, ctx->u_flag.as<uint8_t>(), d_vaf, d_hist, d_hist + (size_t)S * 101);
  ctx->launches += 1;
  CUDA_TRY(ctx, cudaGetLastError());
  std::vector<unsigned long long> h((size_t)S * 101 + S);
  CUDA_TRY(ctx, cudaMemcpyAsync(h.data(), d_hist, hist_bytes, cudaMemcpyDeviceToHost, ctx->stream));
  GUAC_TRY(check_device_error(ctx));  // synchronises
  for (size_t i = 0; i < (size_t)S * 101; ++i) out_hist[i] = (int64_t)h[i];
  int64_t total = 0;
  for (int s = 0; s < S; ++s) {
    out_n_variant_loci[s] = (int64_t)h[(size_t)S * 101 + s];
    total += out_n_variant_loci[s];
  }
  if (out_loci) {  // the VariantLocus records in (range order, locus, sample) order
    int64_t* off = ctx->u_off.as<int64_t>();
    GUAC_TRY((device_scan<true, uint8_t>(ctx, ctx->u_flag.as<uint8_t>(), off, n_cells, OpSum())));
    TmpDev d_out;
    GUAC_TRY(ensure(ctx, d_out, (size_t)std::max<int64_t>(total, 1) * sizeof(guac_variant_locus)));
    vaf_records_kernel<<<(unsigned)plan.n_tiles, 128, 0, ctx->stream>>>(ctx->u_flag.as<uint8_t>(), off, d_vaf, ctx->b_tiles.as<Tile>(), plan.n_tiles,
                                                                        TILE_LOCI, S, d_out.as<guac_variant_locus>());
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    GUAC_TRY(host_out(ctx, out_loci, d_out.p, (size_t)total));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    if (n_loci) *n_loci = total;
  }
  return GUAC_OK;
}

}  // extern "C"
