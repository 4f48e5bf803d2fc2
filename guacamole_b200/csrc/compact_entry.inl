// compact_entry.inl — guac_reads_upload_compact / guac_reads_export_compact (included at the end of guac_b200.cu).

namespace {
// guac_read_batch_compact::mm_count given: the full 2-bit base column from the explicit reads' stream, the contigs' 2-bit
// reference windows and the difference words.  Runs on the context's stream behind the copy of the explicit bases.
int expand_ref_relative(guac_ctx* ctx, const guac_read_batch_compact* b, int64_t n, uint64_t n_bases) {
  GUAC_TRY(upload(ctx, ctx->u_mm_count, b->mm_count, (size_t)n));
  GUAC_TRY(upload(ctx, ctx->u_mm_words, b->mm_words, (size_t)b->n_mm_words, 16));
  GUAC_TRY(ensure(ctx, ctx->u_expl_len, (size_t)n * sizeof(uint16_t) + 16));
  GUAC_TRY(ensure(ctx, ctx->u_mm_n, (size_t)n + 16));
  GUAC_TRY(ensure(ctx, ctx->u_expl_off, (size_t)(n + 1) * sizeof(int64_t)));
  GUAC_TRY(ensure(ctx, ctx->u_mm_off, (size_t)(n + 1) * sizeof(int64_t)));
  compact::refrel_lengths_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(n, ctx->u_mm_count.as<uint8_t>(), ctx->b_seq_off.as<uint64_t>(),
                                                                             ctx->u_expl_len.as<uint16_t>(), ctx->u_mm_n.as<uint8_t>(),
                                                                             ctx->u_bad.as<int>());
  ctx->launches += 1;
  GUAC_TRY((device_scan<true, uint16_t>(ctx, ctx->u_expl_len.as<uint16_t>(), ctx->u_expl_off.as<int64_t>(), n, OpSum(),
                                        ctx->u_expl_off.as<int64_t>() + n)));
  GUAC_TRY((device_scan<true, uint8_t>(ctx, ctx->u_mm_n.as<uint8_t>(), ctx->u_mm_off.as<int64_t>(), n, OpSum(), ctx->u_mm_off.as<int64_t>() + n)));
  int64_t totals[2] = {0, 0};
  CUDA_TRY(ctx, cudaMemcpyAsync(&totals[0], ctx->u_expl_off.as<int64_t>() + n, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(&totals[1], ctx->u_mm_off.as<int64_t>() + n, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  if (totals[0] != b->n_explicit_bases || totals[1] != b->n_mm_words)
    return fail(ctx, GUAC_E_INVALID_INPUT, "reference-relative bases: n_explicit_bases / n_mm_words do not match mm_count");
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->b_bases2.p, 0, (size_t)((n_bases + 3) / 4) + 64, ctx->stream));
  CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream, ctx->join, 0));  // the explicit bases have arrived
  for (auto& sp : ctx->contig_spans) {  // one launch per contig of the batch: its reference rides in the arguments
    const ContigDev& c = ctx->contigs.at(sp.first);
    const int64_t lo = sp.second.first, hi = sp.second.second;
    if (hi <= lo) continue;
    compact::refrel_expand_kernel<<<blocks_for(hi - lo, 256), 256, 0, ctx->stream>>>(
        lo, hi, ctx->u_mm_count.as<uint8_t>(), ctx->u_mm_words.as<uint16_t>(), ctx->u_mm_off.as<int64_t>(), ctx->u_expl.as<uint32_t>(),
        ctx->u_expl_off.as<int64_t>(), ctx->b_start.as<int64_t>(), ctx->b_seq_off.as<uint64_t>(), c.ref2, c.offset, c.len,
        ctx->b_bases2.as<uint32_t>(), ctx->u_bad.as<int>());
    ctx->launches += 1;
  }
  CUDA_TRY(ctx, cudaGetLastError());
  return GUAC_OK;
}
}  // namespace

int guac_reads_upload_compact(guac_ctx* ctx, const guac_read_batch_compact* b) {
  if (!ctx || !b) return GUAC_E_INVALID_INPUT;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  ctx->have_reads = false;
  const int64_t n = b->n_reads;
  if (n < 0 || b->n_samples <= 0 || b->n_samples > 255 || b->n_contig_runs < 0 || (b->qual_bits != 2 && b->qual_bits != 4 && b->qual_bits != 8))
    return fail(ctx, GUAC_E_INVALID_INPUT, "bad compact read batch header");
  // contig runs: distinct contigs, increasing read indices (RegionsByContig.scala:22 "Regions are not sorted by contig")
  ctx->contig_spans.clear();
  for (int i = 0; i < b->n_contig_runs; ++i) {
    const int32_t c = b->run_contig_id[i];
    const int64_t lo = b->run_first_read[i], hi = b->run_first_read[i + 1];
    if (c < 0 || c >= (1 << 22)) return fail(ctx, GUAC_E_INVALID_INPUT, "contig_id out of range");
    if (lo > hi || (i == 0 && lo != 0) || hi > n) return fail(ctx, GUAC_E_INVALID_INPUT, "bad contig runs");
    for (auto& sp : ctx->contig_spans)
      if (sp.first == c) return fail(ctx, GUAC_E_INVALID_INPUT, "Regions are not sorted by contig");
    if (hi > lo) ctx->contig_spans.push_back({c, {lo, hi}});
  }
  if (n > 0 && (b->n_contig_runs == 0 || b->run_first_read[b->n_contig_runs] != n))
    return fail(ctx, GUAC_E_INVALID_INPUT, "contig runs do not cover the batch");
  // reference-relative bases: the 2-bit copies of the contigs' reference windows are the source of most reads' bases
  const bool ref_rel = b->mm_count != nullptr;
  if (ref_rel) {
    if (b->n_mm_words < 0 || b->n_explicit_bases < 0 || (b->n_mm_words > 0 && !b->mm_words))
      return fail(ctx, GUAC_E_INVALID_INPUT, "bad reference-relative columns");
    for (auto& sp : ctx->contig_spans)
      if (!ctx->contigs.count(sp.first))
        return fail(ctx, GUAC_E_STATE, "reference-relative bases: upload the reference window of contig " + std::to_string(sp.first) +
                                           " before the reads");
    GUAC_TRY(sync_ref2(ctx));
  }
  // scratch kept in the context: a cudaMalloc / cudaFree pair per column and call costs more than the kernels here
  DevBuf &d_len = ctx->u_len, &d_ops = ctx->u_ops, &d_runs_c = ctx->u_runs_c, &d_runs_f = ctx->u_runs_f, &d_tmp = ctx->u_tmp,
         &d_qpacked = ctx->u_qpacked, &d_bad = ctx->u_bad;
  GUAC_TRY(side_stream(ctx));
  // which of the three narrow columns the caller sent (each replaces its wide column)
  const bool narrow_start = b->start_rel != nullptr, narrow_ops = b->n_cigar_ops8 != nullptr;
  const int64_t uniform_len = b->uniform_read_len > 0 ? b->uniform_read_len : 0;
  if (narrow_start && b->n_contig_runs > 0 && !b->run_start_base) return fail(ctx, GUAC_E_INVALID_INPUT, "start_rel without run_start_base");
  if (uniform_len > 65535) return fail(ctx, GUAC_E_INVALID_INPUT, "uniform_read_len beyond 65535");
  // The qualities are the largest column.  When the number of bases is known up front (one read length for the whole batch)
  // their copy goes first, on the second stream, with their expansion behind it there: the expansion then runs under the
  // copies of the other columns instead of after the last of them.
  const bool early_qual = uniform_len > 0 && n > 0 && b->qual_bits != 8;
  if (early_qual) {
    const uint64_t nb = (uint64_t)n * (uint64_t)uniform_len;
    const int per_byte = 8 / b->qual_bits;
    const size_t packed = (size_t)((nb + per_byte - 1) / per_byte);
    GUAC_TRY(ensure(ctx, d_qpacked, packed + 32 + 16));
    GUAC_TRY(ensure(ctx, ctx->b_qual, (size_t)nb + 32));
    CUDA_TRY(ctx, cudaEventRecord(ctx->fork, ctx->stream));  // (behind whatever still reads the old batch's buffers)
    CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream2, ctx->fork, 0));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_qpacked.p, b->qual, packed, cudaMemcpyHostToDevice, ctx->stream2));
    compact::QualTable T;
    std::memcpy(T.q, b->qual_table, 16);
    const unsigned blocks = blocks_for((int64_t)((nb + 15) / 16), 256);
    if (b->qual_bits == 4)
      compact::expand_qual_kernel<4><<<blocks, 256, 0, ctx->stream2>>>(nb, d_qpacked.as<uint8_t>(), T, ctx->b_qual.as<uint8_t>());
    else
      compact::expand_qual_kernel<2><<<blocks, 256, 0, ctx->stream2>>>(nb, d_qpacked.as<uint8_t>(), T, ctx->b_qual.as<uint8_t>());
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaEventRecord(ctx->qual_ready, ctx->stream2));
  }
  // (the copy is in flight out of the caller's buffer from here on: whichever way the function returns, it waits for it)
  struct WaitOnExit {
    cudaEvent_t ev;
    bool armed;
    ~WaitOnExit() {
      if (armed) cudaEventSynchronize(ev);
    }
  } qual_copy_guard{ctx->qual_ready, early_qual};
  if (narrow_start) {
    GUAC_TRY(upload(ctx, ctx->u_start_rel, b->start_rel, (size_t)n));
    GUAC_TRY(upload(ctx, ctx->u_runs_base, b->run_start_base, (size_t)b->n_contig_runs));
    GUAC_TRY(ensure(ctx, ctx->b_start, (size_t)std::max<int64_t>(n, 1) * sizeof(int64_t)));
  } else {
    GUAC_TRY(upload(ctx, ctx->b_start, b->start, (size_t)n));
  }
  if (!uniform_len) GUAC_TRY(upload(ctx, d_len, b->read_len, (size_t)n));
  if (narrow_ops) {
    GUAC_TRY(upload(ctx, ctx->u_ops8, b->n_cigar_ops8, (size_t)n));
    GUAC_TRY(ensure(ctx, ctx->u_ops_eff, (size_t)n + 16));
    GUAC_TRY(ensure(ctx, ctx->u_wire_off, (size_t)(n + 1) * sizeof(int64_t)));
  } else {
    GUAC_TRY(upload(ctx, d_ops, b->n_cigar_ops, (size_t)n));
  }
  GUAC_TRY(upload(ctx, d_runs_c, b->run_contig_id, (size_t)b->n_contig_runs));
  GUAC_TRY(upload(ctx, d_runs_f, b->run_first_read, (size_t)b->n_contig_runs + 1));
  GUAC_TRY(upload(ctx, ctx->b_mapq, b->mapq, (size_t)n));
  GUAC_TRY(upload(ctx, ctx->b_flags, b->flags, (size_t)n));
  GUAC_TRY(upload(ctx, ctx->b_sample, b->sample_id, (size_t)n));
  GUAC_TRY(upload(ctx, ctx->b_exc_pos, b->exc_pos, (size_t)b->n_exc));
  GUAC_TRY(upload(ctx, ctx->b_exc_base, b->exc_base, (size_t)b->n_exc));
  // offsets = prefix sums of the lengths (the last entry is the total)
  GUAC_TRY(ensure(ctx, ctx->b_seq_off, (size_t)(n + 1) * sizeof(uint64_t)));
  GUAC_TRY(ensure(ctx, ctx->b_cigar_off, (size_t)(n + 1) * sizeof(uint32_t)));
  GUAC_TRY(ensure(ctx, ctx->b_contig, (size_t)std::max<int64_t>(n, 1) * sizeof(int32_t)));
  GUAC_TRY(ensure(ctx, d_tmp, (size_t)(n + 1) * sizeof(int64_t)));
  GUAC_TRY(ensure(ctx, d_bad, sizeof(int)));
  CUDA_TRY(ctx, cudaMemsetAsync(d_bad.p, 0, sizeof(int), ctx->stream));
  uint64_t n_bases = 0;
  int64_t n_cigar = 0, n_wire = 0;  // operators of the batch, and how many of them travel (all, without the narrow column)
  if (n > 0) {
    int64_t* seq_off = ctx->b_seq_off.as<int64_t>();
    if (uniform_len) {
      compact::uniform_offsets_kernel<<<blocks_for(n + 1, 256), 256, 0, ctx->stream>>>(n + 1, (uint64_t)uniform_len, ctx->b_seq_off.as<uint64_t>());
      ctx->launches += 1;
    } else {
      GUAC_TRY((device_scan<true, uint16_t>(ctx, d_len.as<uint16_t>(), seq_off, n, OpSum(), seq_off + n)));
    }
    if (narrow_ops) {
      compact::effective_ops_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(n, ctx->u_ops8.as<uint8_t>(), ctx->u_ops_eff.as<uint8_t>());
      ctx->launches += 1;
      GUAC_TRY((device_scan<true, uint8_t>(ctx, ctx->u_ops_eff.as<uint8_t>(), d_tmp.as<int64_t>(), n, OpSum(), d_tmp.as<int64_t>() + n)));
      GUAC_TRY((device_scan<true, uint8_t>(ctx, ctx->u_ops8.as<uint8_t>(), ctx->u_wire_off.as<int64_t>(), n, OpSum(),
                                           ctx->u_wire_off.as<int64_t>() + n)));
    } else {
      GUAC_TRY((device_scan<true, uint16_t>(ctx, d_ops.as<uint16_t>(), d_tmp.as<int64_t>(), n, OpSum(), d_tmp.as<int64_t>() + n)));
    }
    compact::narrow_offsets_kernel<<<blocks_for(n + 1, 256), 256, 0, ctx->stream>>>(n + 1, d_tmp.as<int64_t>(), ctx->b_cigar_off.as<uint32_t>());
    compact::contig_from_runs_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(n, d_runs_c.as<int32_t>(), d_runs_f.as<int64_t>(),
                                                                                 b->n_contig_runs, ctx->b_contig.as<int32_t>());
    if (narrow_start) {
      compact::starts_from_rel_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(n, ctx->u_runs_base.as<int64_t>(), d_runs_f.as<int64_t>(),
                                                                                  b->n_contig_runs, ctx->u_start_rel.as<uint32_t>(),
                                                                                  ctx->b_start.as<int64_t>());
      ctx->launches += 1;
    }
    compact::validate_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(n, ctx->b_start.as<int64_t>(), ctx->b_contig.as<int32_t>(),
                                                                         ctx->b_sample.as<uint8_t>(), b->n_samples, d_bad.as<int>());
    ctx->launches += 3;
    CUDA_TRY(ctx, cudaGetLastError());
    CUDA_TRY(ctx, cudaMemcpyAsync(&n_bases, seq_off + n, sizeof n_bases, cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(&n_cigar, d_tmp.as<int64_t>() + n, sizeof n_cigar, cudaMemcpyDeviceToHost, ctx->stream));
    if (narrow_ops) CUDA_TRY(ctx, cudaMemcpyAsync(&n_wire, ctx->u_wire_off.as<int64_t>() + n, sizeof n_wire, cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    if (!narrow_ops) n_wire = n_cigar;
    if (n_cigar >= ((int64_t)1 << 32)) return fail(ctx, GUAC_E_UNSUPPORTED, "more than 2^32-1 CIGAR operators in one batch");
  } else {
    CUDA_TRY(ctx, cudaMemsetAsync(ctx->b_seq_off.p, 0, sizeof(uint64_t), ctx->stream));
    CUDA_TRY(ctx, cudaMemsetAsync(ctx->b_cigar_off.p, 0, sizeof(uint32_t), ctx->stream));
  }
  if (narrow_ops) {
    GUAC_TRY(upload(ctx, ctx->u_wire, b->cigar, (size_t)n_wire));
    GUAC_TRY(ensure(ctx, ctx->b_cigar, (size_t)n_cigar * sizeof(uint32_t) + 16));
    if (n > 0) {
      compact::cigar_from_ops8_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(
          n, ctx->u_ops8.as<uint8_t>(), ctx->b_cigar_off.as<uint32_t>(), ctx->u_wire_off.as<int64_t>(), ctx->u_wire.as<uint32_t>(),
          ctx->b_seq_off.as<uint64_t>(), ctx->b_cigar.as<uint32_t>());
      ctx->launches += 1;
    }
  } else {
    GUAC_TRY(upload(ctx, ctx->b_cigar, b->cigar, (size_t)n_cigar));
  }
  GUAC_TRY(ensure(ctx, ctx->b_bases2, (size_t)((n_bases + 3) / 4) + 64 + 16));
  GUAC_TRY(ensure(ctx, ctx->b_nmask, (size_t)((n_bases + 7) / 8) + 16));
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->b_nmask.p, 0, (size_t)((n_bases + 7) / 8) + 16, ctx->stream));
  if (b->n_exc > 0) {
    compact::nmask_from_exceptions_kernel<<<blocks_for(b->n_exc, 256), 256, 0, ctx->stream>>>(b->n_exc, ctx->b_exc_pos.as<uint64_t>(),
                                                                                             ctx->b_nmask.as<uint8_t>());
    ctx->launches += 1;
  }
  // The qualities travel before the bases: nothing of the on-device preparation below (quality expansion, read ends,
  // flags, headers, the multi-operator list) reads a base, so the copy of the bases, queued on the second stream
  // behind the copy of the qualities, runs under those kernels.
  if (early_qual) {
    if (n_bases != (uint64_t)n * (uint64_t)uniform_len) {  // (cannot happen: the offsets are n x uniform_len)
      return fail(ctx, GUAC_E_INVALID_INPUT, "uniform_read_len does not match the batch");
    }
  } else if (b->qual_bits == 8) {
    GUAC_TRY(upload(ctx, ctx->b_qual, b->qual, (size_t)n_bases));
  } else {
    const int per_byte = 8 / b->qual_bits;
    GUAC_TRY(upload(ctx, d_qpacked, b->qual, (size_t)((n_bases + per_byte - 1) / per_byte), 32));
    GUAC_TRY(ensure(ctx, ctx->b_qual, (size_t)n_bases + 32));
  }
  CUDA_TRY(ctx, cudaEventRecord(ctx->fork, ctx->stream));
  CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->stream2, ctx->fork, 0));
  if (ref_rel) {  // the explicit reads' bases only; the rest is built below
    GUAC_TRY(ensure(ctx, ctx->u_expl, (size_t)((b->n_explicit_bases + 3) / 4) + 64));
    if (b->n_explicit_bases > 0)
      CUDA_TRY(ctx, cudaMemcpyAsync(ctx->u_expl.p, b->bases2, (size_t)((b->n_explicit_bases + 3) / 4), cudaMemcpyHostToDevice, ctx->stream2));
  } else if (n_bases > 0) {
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->b_bases2.p, b->bases2, (size_t)((n_bases + 3) / 4), cudaMemcpyHostToDevice, ctx->stream2));
  }
  // From here on the copy of the bases is in flight out of the caller's buffer: every exit goes through the wait on
  // `join` at the end, so no early return below.
  int st = GUAC_OK;
  auto cuda_ok = [&](cudaError_t e, const char* what) {
    if (e != cudaSuccess && st == GUAC_OK) {
      ctx->err = std::string("CUDA error: ") + cudaGetErrorString(e) + " at " + what;
      st = GUAC_E_CUDA;
    }
  };
  cuda_ok(cudaEventRecord(ctx->join, ctx->stream2), "cudaEventRecord(join)");
  if (early_qual) {
    cuda_ok(cudaStreamWaitEvent(ctx->stream, ctx->qual_ready, 0), "cudaStreamWaitEvent(qual_ready)");  // what follows reads the qualities
  } else if (b->qual_bits != 8 && n_bases > 0) {
    compact::QualTable T;
    std::memcpy(T.q, b->qual_table, 16);
    const unsigned blocks = blocks_for((int64_t)((n_bases + 15) / 16), 256);
    if (b->qual_bits == 4)
      compact::expand_qual_kernel<4><<<blocks, 256, 0, ctx->stream>>>(n_bases, d_qpacked.as<uint8_t>(), T, ctx->b_qual.as<uint8_t>());
    else
      compact::expand_qual_kernel<2><<<blocks, 256, 0, ctx->stream>>>(n_bases, d_qpacked.as<uint8_t>(), T, ctx->b_qual.as<uint8_t>());
    ctx->launches += 1;
  }
  cuda_ok(cudaGetLastError(), "the quality expansion");
  if (ref_rel && n > 0 && st == GUAC_OK) st = expand_ref_relative(ctx, b, n, n_bases);
  int bad = 0;
  cuda_ok(cudaMemcpyAsync(&bad, d_bad.p, sizeof bad, cudaMemcpyDeviceToHost, ctx->stream), "cudaMemcpyAsync(bad)");
  cuda_ok(cudaStreamSynchronize(ctx->stream), "cudaStreamSynchronize");
  if (st == GUAC_OK && (bad & 1)) st = fail(ctx, GUAC_E_INVALID_INPUT, "Regions must be sorted by start locus");
  if (st == GUAC_OK && (bad & 2)) st = fail(ctx, GUAC_E_INVALID_INPUT, "sample_id >= n_samples");
  if (st == GUAC_OK && (bad & 4)) st = fail(ctx, GUAC_E_INVALID_INPUT, "read start out of range");
  if (st == GUAC_OK && (bad & (8 | 16 | 32)))
    st = fail(ctx, GUAC_E_INVALID_INPUT, "reference-relative bases: a read of 16384 bases or more, a read outside its contig's reference window, "
                                         "or a difference word out of range");
  if (st == GUAC_OK) st = finish_reads(ctx, n, b->n_samples, b->n_exc);
  // the caller's buffers may go away after return, whatever the outcome: the copy of the bases must have ended
  const cudaError_t e2 = cudaEventSynchronize(ctx->join);
  if (st == GUAC_OK && e2 != cudaSuccess) {
    ctx->err = std::string("CUDA error: ") + cudaGetErrorString(e2);
    ctx->have_reads = false;
    st = GUAC_E_CUDA;
  }
  return st;
}

int guac_reads_export_compact(guac_ctx* ctx, guac_read_batch_compact* out) {
  if (!ctx || !out) return GUAC_E_INVALID_INPUT;
  if (!ctx->have_reads) return fail(ctx, GUAC_E_STATE, "no reads uploaded");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const int64_t n = ctx->n_reads;
  std::memset(out, 0, sizeof *out);
  out->n_reads = n;
  out->n_samples = ctx->n_samples;
  uint64_t n_bases = 0;
  uint32_t n_cigar = 0;
  if (n) {
    CUDA_TRY(ctx, cudaMemcpyAsync(&n_bases, ctx->R.seq_off + n, 8, cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(&n_cigar, ctx->R.cigar_off + n, 4, cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  }
  // contig runs from the spans recorded at upload
  auto spans = ctx->contig_spans;
  std::sort(spans.begin(), spans.end(), [](const auto& a, const auto& b) { return a.second.first < b.second.first; });
  out->n_contig_runs = (int32_t)spans.size();
  void* hc = nullptr;
  void* hf = nullptr;
  GUAC_TRY(pinned_alloc(ctx, &hc, std::max<size_t>(spans.size() * sizeof(int32_t), 16)));
  GUAC_TRY(pinned_alloc(ctx, &hf, (spans.size() + 1) * sizeof(int64_t)));
  for (size_t i = 0; i < spans.size(); ++i) {
    static_cast<int32_t*>(hc)[i] = spans[i].first;
    static_cast<int64_t*>(hf)[i] = spans[i].second.first;
  }
  static_cast<int64_t*>(hf)[spans.size()] = n;
  out->run_contig_id = static_cast<int32_t*>(hc);
  out->run_first_read = static_cast<int64_t*>(hf);
  DevBuf d_len, d_ops, d_flag, d_present, d_qpacked;
  struct Free {
    std::vector<DevBuf*> v;
    ~Free() {
      for (DevBuf* d : v)
        if (d->p) cudaFree(d->p);
    }
  } fr{{&d_len, &d_ops, &d_flag, &d_present, &d_qpacked}};
  GUAC_TRY(ensure(ctx, d_len, (size_t)std::max<int64_t>(n, 1) * sizeof(uint16_t)));
  GUAC_TRY(ensure(ctx, d_ops, (size_t)std::max<int64_t>(n, 1) * sizeof(uint16_t)));
  GUAC_TRY(ensure(ctx, d_flag, sizeof(int)));
  GUAC_TRY(ensure(ctx, d_present, 256 * sizeof(unsigned int)));
  CUDA_TRY(ctx, cudaMemsetAsync(d_flag.p, 0, sizeof(int), ctx->stream));
  CUDA_TRY(ctx, cudaMemsetAsync(d_present.p, 0, 256 * sizeof(unsigned int), ctx->stream));
  unsigned int present[256] = {0};
  int too_long = 0;
  if (n) {
    compact::lengths_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(n, ctx->R.seq_off, ctx->R.cigar_off, d_len.as<uint16_t>(),
                                                                        d_ops.as<uint16_t>(), d_flag.as<int>());
    compact::qual_histogram_kernel<<<148 * 8, 256, 0, ctx->stream>>>(n_bases, ctx->R.qual, d_present.as<unsigned int>());
    ctx->launches += 2;
    CUDA_TRY(ctx, cudaGetLastError());
    CUDA_TRY(ctx, cudaMemcpyAsync(present, d_present.p, sizeof present, cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(&too_long, d_flag.p, sizeof too_long, cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    if (too_long) return fail(ctx, GUAC_E_UNSUPPORTED, "a read has more than 65535 bases or CIGAR operators: use guac_reads_export");
  }
  // the narrow columns, where the batch allows them
  bool narrow_start = false, narrow_ops = false;
  int64_t uniform_len = 0, n_wire = 0;
  DevBuf d_rel, d_ops8, d_woff, d_wire, d_nflags;
  Free fr2{{&d_rel, &d_ops8, &d_woff, &d_wire, &d_nflags}};
  if (n && ctx->opt_compact_narrow) {
    GUAC_TRY(upload(ctx, ctx->u_runs_f, static_cast<const int64_t*>(hf), spans.size() + 1));
    GUAC_TRY(ensure(ctx, d_rel, (size_t)n * sizeof(uint32_t)));
    GUAC_TRY(ensure(ctx, d_ops8, (size_t)n + 16));
    GUAC_TRY(ensure(ctx, d_nflags, sizeof(int)));
    CUDA_TRY(ctx, cudaMemsetAsync(d_nflags.p, 0, sizeof(int), ctx->stream));
    compact::narrow_columns_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(
        n, ctx->R.start, ctx->u_runs_f.as<int64_t>(), (int)spans.size(), ctx->R.seq_off, ctx->R.cigar_off, ctx->R.cigar,
        d_rel.as<uint32_t>(), d_ops8.as<uint8_t>(), d_nflags.as<int>());
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    int nflags = 0;
    uint64_t first_offsets[2] = {0, 0};
    CUDA_TRY(ctx, cudaMemcpyAsync(&nflags, d_nflags.p, sizeof nflags, cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(first_offsets, ctx->R.seq_off, sizeof first_offsets, cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    narrow_start = !(nflags & 1);
    narrow_ops = !(nflags & 2);
    if (!(nflags & 4) && first_offsets[1] - first_offsets[0] > 0) uniform_len = (int64_t)(first_offsets[1] - first_offsets[0]);
    if (narrow_ops) {
      GUAC_TRY(ensure(ctx, d_woff, (size_t)(n + 1) * sizeof(int64_t)));
      GUAC_TRY((device_scan<true, uint8_t>(ctx, d_ops8.as<uint8_t>(), d_woff.as<int64_t>(), n, OpSum(), d_woff.as<int64_t>() + n)));
      CUDA_TRY(ctx, cudaMemcpyAsync(&n_wire, d_woff.as<int64_t>() + n, sizeof n_wire, cudaMemcpyDeviceToHost, ctx->stream));
      CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
      GUAC_TRY(ensure(ctx, d_wire, (size_t)n_wire * sizeof(uint32_t) + 16));
      compact::wire_cigar_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(n, d_ops8.as<uint8_t>(), ctx->R.cigar_off, ctx->R.cigar,
                                                                            d_woff.as<int64_t>(), d_wire.as<uint32_t>());
      ctx->launches += 1;
      CUDA_TRY(ctx, cudaGetLastError());
    }
  }
  if (narrow_start) {
    GUAC_TRY(host_out(ctx, const_cast<uint32_t**>(&out->start_rel), d_rel.p, (size_t)n));
    void* hb = nullptr;
    GUAC_TRY(pinned_alloc(ctx, &hb, std::max<size_t>(spans.size() * sizeof(int64_t), 16)));
    for (size_t i = 0; i < spans.size(); ++i)  // the run's base = the start of its first read
      CUDA_TRY(ctx, cudaMemcpyAsync(static_cast<int64_t*>(hb) + i, ctx->R.start + spans[i].second.first, sizeof(int64_t),
                                    cudaMemcpyDeviceToHost, ctx->stream));
    out->run_start_base = static_cast<int64_t*>(hb);
  } else {
    GUAC_TRY(host_out(ctx, const_cast<int64_t**>(&out->start), ctx->b_start.p, (size_t)n));
  }
  if (uniform_len) out->uniform_read_len = (int32_t)uniform_len;
  else GUAC_TRY(host_out(ctx, const_cast<uint16_t**>(&out->read_len), d_len.p, (size_t)n));
  if (narrow_ops) {
    GUAC_TRY(host_out(ctx, const_cast<uint8_t**>(&out->n_cigar_ops8), d_ops8.p, (size_t)n));
    GUAC_TRY(host_out(ctx, const_cast<uint32_t**>(&out->cigar), d_wire.p, (size_t)n_wire));
  } else {
    GUAC_TRY(host_out(ctx, const_cast<uint16_t**>(&out->n_cigar_ops), d_ops.p, (size_t)n));
    GUAC_TRY(host_out(ctx, const_cast<uint32_t**>(&out->cigar), ctx->b_cigar.p, (size_t)n_cigar));
  }
  bool ref_rel = ctx->opt_compact_ref_relative && n > 0;
  for (auto& sp : ctx->contig_spans) ref_rel = ref_rel && ctx->contigs.count(sp.first) != 0;
  DevBuf d_mm_count, d_mm_words, d_expl, d_expl_len, d_mm_n, d_expl_off, d_mm_off;
  Free fr3{{&d_mm_count, &d_mm_words, &d_expl, &d_expl_len, &d_mm_n, &d_expl_off, &d_mm_off}};
  if (ref_rel) {
    // the read pass over the whole batch: a single-operator read's records are its differences from the reference
    ctx->plan_ranges.clear();
    for (auto& sp : ctx->contig_spans) {
      RangeDev rg;
      std::memset(&rg, 0, sizeof rg);
      rg.contig_id = sp.first;
      rg.start = 0;
      rg.end = (int64_t)1 << 40;
      ctx->plan_ranges.push_back(rg);
    }
    GUAC_TRY(run_read_pass(ctx));
    GUAC_TRY(ensure(ctx, d_mm_count, (size_t)n + 16));
    GUAC_TRY(ensure(ctx, d_expl_len, (size_t)n * sizeof(uint16_t) + 16));
    GUAC_TRY(ensure(ctx, d_mm_n, (size_t)n + 16));
    GUAC_TRY(ensure(ctx, d_expl_off, (size_t)(n + 1) * sizeof(int64_t)));
    GUAC_TRY(ensure(ctx, d_mm_off, (size_t)(n + 1) * sizeof(int64_t)));
    for (auto& sp : ctx->contig_spans) {
      const ContigDev& c = ctx->contigs.at(sp.first);
      const int64_t lo = sp.second.first, hi = sp.second.second;
      if (hi <= lo) continue;
      compact::refrel_classify_kernel<<<blocks_for(hi - lo, 256), 256, 0, ctx->stream>>>(
          lo, hi, ctx->b_hdr.as<uint4>(), ctx->b_bin_count.as<uint8_t>(), STREAM_EV_CAP, ctx->R.start, ctx->R.end, ctx->R.seq_off, c.offset, c.len,
          c.refbad, d_mm_count.as<uint8_t>(), d_expl_len.as<uint16_t>(), d_mm_n.as<uint8_t>());
      ctx->launches += 1;
    }
    CUDA_TRY(ctx, cudaGetLastError());
    GUAC_TRY((device_scan<true, uint16_t>(ctx, d_expl_len.as<uint16_t>(), d_expl_off.as<int64_t>(), n, OpSum(), d_expl_off.as<int64_t>() + n)));
    GUAC_TRY((device_scan<true, uint8_t>(ctx, d_mm_n.as<uint8_t>(), d_mm_off.as<int64_t>(), n, OpSum(), d_mm_off.as<int64_t>() + n)));
    int64_t totals[2] = {0, 0};
    CUDA_TRY(ctx, cudaMemcpyAsync(&totals[0], d_expl_off.as<int64_t>() + n, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(&totals[1], d_mm_off.as<int64_t>() + n, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    out->n_explicit_bases = totals[0];
    out->n_mm_words = totals[1];
    GUAC_TRY(ensure(ctx, d_mm_words, (size_t)totals[1] * sizeof(uint16_t) + 16));
    GUAC_TRY(ensure(ctx, d_expl, (size_t)((totals[0] + 3) / 4) + 64));
    CUDA_TRY(ctx, cudaMemsetAsync(d_expl.p, 0, (size_t)((totals[0] + 3) / 4) + 64, ctx->stream));
    for (auto& sp : ctx->contig_spans) {
      const ContigDev& c = ctx->contigs.at(sp.first);
      const int64_t lo = sp.second.first, hi = sp.second.second;
      if (hi <= lo) continue;
      compact::refrel_words_kernel<<<blocks_for(hi - lo, 256), 256, 0, ctx->stream>>>(
          lo, hi, d_mm_count.as<uint8_t>(), d_mm_off.as<int64_t>(), ctx->b_bins.as<uint2>(), STREAM_EV_CAP, ctx->R.start, c.ref2, c.offset,
          d_mm_words.as<uint16_t>());
      ctx->launches += 1;
    }
    compact::refrel_explicit_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(n, d_mm_count.as<uint8_t>(), d_expl_off.as<int64_t>(), ctx->R.seq_off,
                                                                               reinterpret_cast<const uint32_t*>(ctx->R.bases2), d_expl.as<uint32_t>());
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    GUAC_TRY(host_out(ctx, const_cast<uint8_t**>(&out->mm_count), d_mm_count.p, (size_t)n));
    GUAC_TRY(host_out(ctx, const_cast<uint16_t**>(&out->mm_words), d_mm_words.p, (size_t)totals[1]));
    GUAC_TRY(host_out(ctx, const_cast<uint8_t**>(&out->bases2), d_expl.p, (size_t)((totals[0] + 3) / 4)));
  } else {
    GUAC_TRY(host_out(ctx, const_cast<uint8_t**>(&out->bases2), ctx->b_bases2.p, (size_t)((n_bases + 3) / 4)));
  }
  out->n_exc = ctx->R.n_exc;
  GUAC_TRY(host_out(ctx, const_cast<uint64_t**>(&out->exc_pos), ctx->b_exc_pos.p, (size_t)ctx->R.n_exc));
  GUAC_TRY(host_out(ctx, const_cast<uint8_t**>(&out->exc_base), ctx->b_exc_base.p, (size_t)ctx->R.n_exc));
  int distinct = 0;
  compact::QualCodes codes;
  std::memset(&codes, 0, sizeof codes);
  for (int q = 0; q < 256; ++q)
    if (present[q]) {
      if (distinct < 16) {
        out->qual_table[distinct] = (uint8_t)q;
        codes.code[q] = (uint8_t)distinct;
      }
      ++distinct;
    }
  // (opt_compact_min_qual_bits: measurement hook — the 4-bit and 8-bit forms of a batch whose qualities would fit 2 bits)
  if (distinct <= 16 && n_bases > 0 && ctx->opt_compact_min_qual_bits < 8) {
    out->qual_bits = (distinct <= 4 && ctx->opt_compact_min_qual_bits <= 2) ? 2 : 4;
    const int per_byte = 8 / out->qual_bits;
    const size_t packed = (size_t)((n_bases + per_byte - 1) / per_byte);
    GUAC_TRY(ensure(ctx, d_qpacked, packed + 16));
    if (out->qual_bits == 4)
      compact::pack_qual_kernel<4><<<blocks_for((int64_t)packed, 256), 256, 0, ctx->stream>>>(n_bases, ctx->R.qual, codes, d_qpacked.as<uint8_t>());
    else
      compact::pack_qual_kernel<2><<<blocks_for((int64_t)packed, 256), 256, 0, ctx->stream>>>(n_bases, ctx->R.qual, codes, d_qpacked.as<uint8_t>());
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    GUAC_TRY(host_out(ctx, const_cast<uint8_t**>(&out->qual), d_qpacked.p, packed));
  } else {
    out->qual_bits = 8;
    std::memset(out->qual_table, 0, sizeof out->qual_table);
    GUAC_TRY(host_out(ctx, const_cast<uint8_t**>(&out->qual), ctx->b_qual.p, (size_t)n_bases));
  }
  GUAC_TRY(host_out(ctx, const_cast<uint8_t**>(&out->mapq), ctx->b_mapq.p, (size_t)n));
  GUAC_TRY(host_out(ctx, const_cast<uint8_t**>(&out->flags), ctx->b_flags.p, (size_t)n));
  GUAC_TRY(host_out(ctx, const_cast<uint8_t**>(&out->sample_id), ctx->b_sample.p, (size_t)n));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return GUAC_OK;
}
