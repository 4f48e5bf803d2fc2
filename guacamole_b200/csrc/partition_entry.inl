// partition_entry.inl — guac_take_loci and guac_capped_partitioning (part of the guac_b200.cu translation unit).

namespace {

// runs take_loci_kernel and copies the intervals out
int run_take_loci(guac_ctx* ctx, TakeLociArgs A, guac_range** out_ranges, int32_t** out_partition, int64_t* n_out, int32_t* n_partitions) {
  GUAC_TRY(pinned_block(ctx));
  int* h = static_cast<int*>(ctx->h_counters);
  GUAC_TRY(ensure(ctx, ctx->b_ncalls, 2 * sizeof(int)));
  int cap = (int)std::max<size_t>(ctx->g_evs.cap / sizeof(TakeLociOut), 1 << 16);
  for (int attempt = 0; attempt < 2; ++attempt) {
    GUAC_TRY(ensure(ctx, ctx->g_evs, (size_t)cap * sizeof(TakeLociOut)));
    CUDA_TRY(ctx, cudaMemsetAsync(ctx->b_ncalls.p, 0, 2 * sizeof(int), ctx->stream));
    A.out = ctx->g_evs.as<TakeLociOut>();
    A.max_out = cap;
    A.n_out = ctx->b_ncalls.as<int>();
    A.n_partitions = ctx->b_ncalls.as<int>() + 1;
    take_loci_kernel<<<1, 32, 0, ctx->stream>>>(A);
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    CUDA_TRY(ctx, cudaMemcpyAsync(h, ctx->b_ncalls.p, 2 * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    if (h[0] <= cap) break;
    cap = h[0];
  }
  const int n = h[0];
  std::vector<TakeLociOut> iv((size_t)n);
  if (n) {
    CUDA_TRY(ctx, cudaMemcpyAsync(iv.data(), ctx->g_evs.p, (size_t)n * sizeof(TakeLociOut), cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  }
  void* hr = nullptr;
  void* hp = nullptr;
  GUAC_TRY(pinned_alloc(ctx, &hr, std::max<size_t>((size_t)n * sizeof(guac_range), 16)));
  {
    const int st = pinned_alloc(ctx, &hp, std::max<size_t>((size_t)n * sizeof(int32_t), 16));
    if (st != GUAC_OK) {
      guac_free(ctx, hr);
      return st;
    }
  }
  for (int i = 0; i < n; ++i) {
    guac_range& g = static_cast<guac_range*>(hr)[i];
    g.contig_id = iv[(size_t)i].contig_id;
    g._pad = 0;
    g.start = iv[(size_t)i].start;
    g.end = iv[(size_t)i].end;
    static_cast<int32_t*>(hp)[i] = iv[(size_t)i].partition;
  }
  *out_ranges = static_cast<guac_range*>(hr);
  *out_partition = static_cast<int32_t*>(hp);
  *n_out = n;
  if (n_partitions) *n_partitions = h[1];
  return GUAC_OK;
}

}  // namespace

extern "C" {

int guac_take_loci(guac_ctx* ctx, const int32_t* contig_id, const int64_t* locus, const int32_t* depth, const int32_t* starts, int64_t n,
                   int32_t max_regions_per_partition, int32_t trim_ranges, guac_range** out_ranges, int32_t** out_partition, int64_t* n_out,
                   int32_t* n_partitions) {
  if (!ctx || !out_ranges || !out_partition || !n_out || n < 0 || (n > 0 && (!contig_id || !locus || !depth || !starts))) return GUAC_E_INVALID_INPUT;
  if (max_regions_per_partition < 1) return fail(ctx, GUAC_E_INVALID_INPUT, "max_regions_per_partition must be >= 1");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  TmpDev d_contig, d_locus, d_depth32, d_depth, d_starts, d_incl;
  GUAC_TRY(upload(ctx, d_contig, contig_id, (size_t)n));
  GUAC_TRY(upload(ctx, d_locus, locus, (size_t)n));
  GUAC_TRY(upload(ctx, d_depth32, depth, (size_t)n));
  GUAC_TRY(upload(ctx, d_starts, starts, (size_t)n));
  GUAC_TRY(ensure(ctx, d_depth, (size_t)(n + 1) * sizeof(int64_t)));
  GUAC_TRY(ensure(ctx, d_incl, (size_t)(n + 1) * sizeof(int64_t)));
  if (n > 0) {
    widen_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(d_depth32.as<int32_t>(), n, d_depth.as<int64_t>());
    ctx->launches += 1;
    GUAC_TRY((device_scan<false, int32_t>(ctx, d_starts.as<int32_t>(), d_incl.as<int64_t>(), n, OpSum())));
  }
  TakeLociArgs A;
  std::memset(&A, 0, sizeof A);
  A.depth = d_depth.as<int64_t>();
  A.starts = d_starts.as<int>();
  A.starts_incl = d_incl.as<int64_t>();
  A.locus = d_locus.as<int64_t>();
  A.contig = d_contig.as<int32_t>();
  A.n = n;
  A.max_regions = max_regions_per_partition;
  A.trim_ranges = trim_ranges;
  return run_take_loci(ctx, A, out_ranges, out_partition, n_out, n_partitions);
}

int guac_capped_partitioning(guac_ctx* ctx, const guac_range* loci, int32_t n_ranges, int32_t half_window, int32_t max_reads_per_partition,
                             int32_t trim_ranges, guac_range** out_ranges, int32_t** out_partition, int64_t* n_out, int32_t* n_partitions) {
  if (!ctx || !out_ranges || !out_partition || !n_out) return GUAC_E_INVALID_INPUT;
  if (max_reads_per_partition < 1 || half_window < 0) return fail(ctx, GUAC_E_INVALID_INPUT, "max_reads_per_partition must be >= 1, half_window >= 0");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  for (int i = 1; i < n_ranges; ++i)
    if (loci[i].contig_id < loci[i - 1].contig_id) return fail(ctx, GUAC_E_INVALID_INPUT, "loci ranges must be in contig order");
  Plan plan;
  GUAC_TRY(make_plan(ctx, loci, n_ranges, plan));
  *n_out = 0;
  if (n_partitions) *n_partitions = 0;
  if (plan.n_tiles == 0) {
    GUAC_TRY(host_out(ctx, out_ranges, nullptr, 0));
    GUAC_TRY(host_out(ctx, out_partition, nullptr, 0));
    return GUAC_OK;
  }
  const int64_t n = plan.n_slots;
  TmpDev d_diff, d_depth, d_starts, d_incl;
  GUAC_TRY(ensure(ctx, d_diff, (size_t)(n + 2) * sizeof(int)));
  GUAC_TRY(ensure(ctx, d_starts, (size_t)(n + 2) * sizeof(int)));
  GUAC_TRY(ensure(ctx, d_depth, (size_t)(n + 2) * sizeof(int64_t)));
  GUAC_TRY(ensure(ctx, d_incl, (size_t)(n + 2) * sizeof(int64_t)));
  CUDA_TRY(ctx, cudaMemsetAsync(d_diff.p, 0, (size_t)(n + 2) * sizeof(int), ctx->stream));
  CUDA_TRY(ctx, cudaMemsetAsync(d_starts.p, 0, (size_t)(n + 2) * sizeof(int), ctx->stream));
  if (ctx->n_reads > 0) {
    coverage_kernel<<<blocks_for(ctx->n_reads, 256), 256, 0, ctx->stream>>>(ctx->R, ctx->b_contig.as<int32_t>(), ctx->b_ranges.as<RangeDev>(),
                                                                            (int)ctx->plan_ranges.size(), half_window, d_diff.as<int>(),
                                                                            d_starts.as<int>());
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
  }
  GUAC_TRY((device_scan<false, int32_t>(ctx, d_diff.as<int32_t>(), d_depth.as<int64_t>(), n, OpSum())));
  GUAC_TRY((device_scan<false, int32_t>(ctx, d_starts.as<int32_t>(), d_incl.as<int64_t>(), n, OpSum())));
  TakeLociArgs A;
  std::memset(&A, 0, sizeof A);
  A.depth = d_depth.as<int64_t>();
  A.starts = d_starts.as<int>();
  A.starts_incl = d_incl.as<int64_t>();
  A.ranges = ctx->b_ranges.as<RangeDev>();
  A.n_ranges = (int)ctx->plan_ranges.size();
  A.n = n;
  A.max_regions = max_reads_per_partition;
  A.trim_ranges = trim_ranges;
  return run_take_loci(ctx, A, out_ranges, out_partition, n_out, n_partitions);
}

}  // extern "C"
