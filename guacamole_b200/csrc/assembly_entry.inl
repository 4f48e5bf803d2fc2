// assembly_entry.inl — guac_germline_assembly: host orchestration of the window-parallel assembler.
// Included at the end of guac_b200.cu (uses that file's context and helpers).
//
// The reference's closure (M/commands/GermlineAssemblyCaller.scala:103-219) carries two pieces of state from
// locus to locus — lastCalledLocus and the window jump after a successful assembly — so which windows it
// assembles depends on earlier results.  Each window's result, however, is a pure function of its locus.
// This path therefore assembles every candidate window of the call on the device at once, and replays the
// reference's walk over the finished results on the host (a few comparisons per candidate locus).

namespace {

// Device scratch of one call: bump allocation out of blocks the context keeps between calls (cudaMalloc / cudaFree
// per stage cost more than the kernels of a window batch).  Strict stack discipline: one DevTmp alive at a time.
struct DevTmp {
  guac_ctx* owner = nullptr;
  ~DevTmp() {
    if (owner) {
      owner->asm_block = 0;
      owner->asm_off = 0;
    }
  }
  template <class T>
  int alloc(guac_ctx* ctx, T** out, size_t n, int fill = -1) {
    owner = ctx;
    const size_t bytes = (std::max<size_t>(n * sizeof(T), 256) + 255) & ~(size_t)255;
    while (ctx->asm_block < ctx->asm_blocks.size() && ctx->asm_off + bytes > ctx->asm_blocks[ctx->asm_block].second) {
      ctx->asm_block += 1;
      ctx->asm_off = 0;
    }
    if (ctx->asm_block == ctx->asm_blocks.size()) {
      void* p = nullptr;
      const size_t want = std::max<size_t>(bytes, (size_t)64 << 20);
      cudaError_t e = cudaMalloc(&p, want);
      if (e != cudaSuccess) {
        (void)cudaGetLastError();
        return fail(ctx, GUAC_E_OOM, std::string("cudaMalloc of ") + std::to_string(want) + " bytes failed: " + cudaGetErrorString(e));
      }
      ctx->asm_blocks.push_back({p, want});
      ctx->asm_off = 0;
    }
    void* p = static_cast<char*>(ctx->asm_blocks[ctx->asm_block].first) + ctx->asm_off;
    ctx->asm_off += bytes;
    if (fill >= 0) CUDA_TRY(ctx, cudaMemsetAsync(p, fill, bytes, ctx->stream));
    *out = reinterpret_cast<T*>(p);
    return GUAC_OK;
  }
};

struct AsmVariant {
  int64_t locus;
  std::string ref, alt;
  bool operator<(const AsmVariant& o) const { return std::tie(locus, ref, alt) < std::tie(o.locus, o.ref, o.alt); }
};

struct AsmWindowResult {
  int32_t contig_id = 0;
  int64_t locus = 0;
  int32_t n_nodes = 0;
  int32_t n_paths = 0;  // after discoverHaplotypes
  std::vector<AsmVariant> variants;
  AsmEvidence evidence{};
};

// AssemblyUtils.buildVariantsFromPath (M/assembly/AssemblyUtils.scala:127-174) over the device-made CIGAR.
// `ref1` is the window's reference slice preceded by one extra base (0 when that base is not in the reference).
int variants_from_cigar(guac_ctx* ctx, const AsmCigar& c, const std::string& path, const std::string& ref1, int64_t reference_start,
                        int32_t contig_id, std::set<AsmVariant>& out) {
  int reference_index = c.ref_start, path_index = 0;
  for (int ci = 0; ci < c.n_elements; ++ci) {
    const int op = c.op[ci];
    const int reference_length = op == 2 ? 0 : c.len[ci];
    const int path_length = op == 3 ? 0 : c.len[ci];
    const int64_t locus = reference_start + reference_index;
    if (op == 1) {
      const std::string r = ref1.substr(1 + reference_index, reference_length), a = path.substr(path_index, path_length);
      if (!r.empty() && !a.empty() && r != a) out.insert({locus, r, a});
    } else if ((op == 2 || op == 3) && ci != 0 && ci != c.n_elements - 1) {
      if (ref1[reference_index] == 0) {
        char buf[96];
        snprintf(buf, sizeof buf, "Position %d:%lld missing from reference", contig_id, (long long)(locus - 1));
        return fail(ctx, GUAC_E_INVALID_INPUT, buf);
      }
      const std::string r = ref1.substr(reference_index, reference_length + 1);
      const std::string a = path.substr(std::max(0, path_index - 1), path_length + 1);
      if (!r.empty() && !a.empty() && r != a) out.insert({locus - 1, r, a});
    }
    reference_index += reference_length;
    path_index += path_length;
  }
  return GUAC_OK;
}

// One thread per window that needs scoring: the (read, path) alignment jobs of AssemblyUtils.scala:92-104.
__global__ void asm_score_jobs_kernel(ReadsView R, const AsmWindow* __restrict__ wins, const int32_t* __restrict__ which,
                                      const int64_t* __restrict__ job_base, const int32_t* __restrict__ sum_base, int n, AsmParams P,
                                      AsmAlignJob* __restrict__ jobs) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const AsmWindow& w = wins[which[i]];
  int64_t j = job_base[i];
  for (int64_t r = w.read_lo; r < w.read_hi; ++r) {
    if (!asm_in_window(R, r, P, w.locus)) continue;
    for (int p = 0; p < w.n_paths; ++p) {
      AsmAlignJob job;
      job.seq = r;
      job.ref = w.path_off + (int64_t)p * w.path_cap;
      job.seq_len = asm_read_len(R, r);
      job.ref_len = w.path_len[p];
      job.seq_kind = 0;
      job.sum_slot = sum_base[i] + p;
      job.trace = -1;
      job._pad = 0;
      job.head_off = 0;
      jobs[j++] = job;
    }
  }
}

size_t asm_align_smem(int S) { return (size_t)3 * (S + 1) * (sizeof(double) + sizeof(int32_t) + 1) + S + 16; }

struct AsmStage {  // device and host state of one batch of windows once their graphs have been walked
  DevTmp tmp;
  AsmParams P{};
  AsmWindow* d_win = nullptr;
  uint8_t* d_strings = nullptr;  // [path strings | window reference slices, each preceded by one base]
  int64_t path_bytes = 0;
  int ref_stride = 0;
  uint8_t* strings = nullptr;    // pinned host copy of d_strings
  guac_ctx* ctx = nullptr;
  ~AsmStage() {
    if (strings) guac_free(ctx, strings);
  }
};

// k-mer tables -> ranked graph -> unipath merge + DFS for the n windows at `h` (updated in place with the path lengths).
// `custom_ref` (P.ref_len bytes) replaces the contig's reference slice for a single window given by the caller.
int build_paths(guac_ctx* ctx, AsmStage& st, AsmWindow* h, int n, const uint8_t* custom_ref) {
  DevTmp& tmp = st.tmp;
  const AsmParams& P = st.P;
  const int K = P.K;
  AsmWindow*& d_win = st.d_win;
  GUAC_TRY(tmp.alloc(ctx, &d_win, (size_t)n));
  const size_t win_bytes = (size_t)n * sizeof(AsmWindow);
  DeviceError* d_err = ctx->b_err.as<DeviceError>();
  CUDA_TRY(ctx, cudaMemcpyAsync(d_win, h, win_bytes, cudaMemcpyHostToDevice, ctx->stream));
  asm_window_kernel<<<blocks_for((int64_t)n * 32, 128), 128, 0, ctx->stream>>>(ctx->R, d_win, n, P);
  ctx->launches += 1;
  CUDA_TRY(ctx, cudaGetLastError());
  CUDA_TRY(ctx, cudaMemcpyAsync(h, d_win, win_bytes, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  // hash tables
  int64_t slots = 0;
  for (int i = 0; i < n; ++i) {
    int cap = 16;
    while (cap < h[i].n_instances + 1) cap <<= 1;  // distinct k-mers <= instances: probing always terminates
    h[i].table_off = slots;
    h[i].table_cap = cap;
    slots += cap;
  }
  AsmTable T;
  GUAC_TRY(tmp.alloc(ctx, &T.tag, (size_t)slots, 0));
  GUAC_TRY(tmp.alloc(ctx, &T.key, (size_t)slots));
  GUAC_TRY(tmp.alloc(ctx, &T.count, (size_t)slots, 0));
  GUAC_TRY(tmp.alloc(ctx, &T.node, (size_t)slots, 0xFF));
  CUDA_TRY(ctx, cudaMemcpyAsync(d_win, h, win_bytes, cudaMemcpyHostToDevice, ctx->stream));
  asm_count_kernel<<<(unsigned)n, 256, 0, ctx->stream>>>(ctx->R, d_win, T, P, d_err);
  ctx->launches += 1;
  CUDA_TRY(ctx, cudaGetLastError());
  CUDA_TRY(ctx, cudaMemcpyAsync(h, d_win, win_bytes, cudaMemcpyDeviceToHost, ctx->stream));
  GUAC_TRY(check_device_error(ctx));
  // graph arrays
  int64_t nodes = 0, path_bytes = 0;
  for (int i = 0; i < n; ++i) {
    h[i].node_off = nodes;
    nodes += h[i].n_nodes;
    h[i].path_cap = (h[i].n_nodes + K + 64 + 7) & ~7;
    h[i].path_off = path_bytes;
    path_bytes += (int64_t)ASM_MAX_PATHS * h[i].path_cap;
    h[i].n_paths = 0;
  }
  st.path_bytes = path_bytes;
  const int ref_stride = st.ref_stride = P.ref_len + 1;
  uint8_t*& d_strings = st.d_strings;
  GUAC_TRY(tmp.alloc(ctx, &d_strings, (size_t)path_bytes + (size_t)n * ref_stride + 16));
  uint8_t* d_ref1 = d_strings + path_bytes;
  AsmGraph G;
  AsmWalk X;
  const size_t N = (size_t)nodes + 1, NW = (size_t)n;
  GUAC_TRY(tmp.alloc(ctx, &G.key, N));
  GUAC_TRY(tmp.alloc(ctx, &G.tmp_key, N));
  GUAC_TRY(tmp.alloc(ctx, &G.tmp_slot, N));
  GUAC_TRY(tmp.alloc(ctx, &G.cgroup, N));
  GUAC_TRY(tmp.alloc(ctx, &G.pargroup, N));
  GUAC_TRY(tmp.alloc(ctx, &G.pgroup, N));
  GUAC_TRY(tmp.alloc(ctx, &G.sgroup, N));
  GUAC_TRY(tmp.alloc(ctx, &G.plist, N * 6));
  GUAC_TRY(tmp.alloc(ctx, &G.slist, N * 6));
  GUAC_TRY(tmp.alloc(ctx, &G.pcount, N));
  GUAC_TRY(tmp.alloc(ctx, &G.scount, N));
  const size_t M2 = 2 * N + 2 * NW + 8;
  GUAC_TRY(tmp.alloc(ctx, &X.first, M2));
  GUAC_TRY(tmp.alloc(ctx, &X.last, M2));
  GUAC_TRY(tmp.alloc(ctx, &X.coff, M2));
  GUAC_TRY(tmp.alloc(ctx, &X.ccnt, M2));
  GUAC_TRY(tmp.alloc(ctx, &X.alive, M2));
  GUAC_TRY(tmp.alloc(ctx, &X.visited, M2));
  GUAC_TRY(tmp.alloc(ctx, &X.stamp, M2));
  GUAC_TRY(tmp.alloc(ctx, &X.mi_node, N));
  GUAC_TRY(tmp.alloc(ctx, &X.mi_idx, N));
  GUAC_TRY(tmp.alloc(ctx, &X.in_all, N));
  GUAC_TRY(tmp.alloc(ctx, &X.pool, 4 * N + 64 * NW + 8));
  GUAC_TRY(tmp.alloc(ctx, &X.chain, 4 * N + 64 * NW + 8));
  GUAC_TRY(tmp.alloc(ctx, &X.frontier, 10 * N + 64 * NW + 8));
  GUAC_TRY(tmp.alloc(ctx, &X.path, 2 * N + 8 * NW + 8));
  GUAC_TRY(tmp.alloc(ctx, &X.branch, 2 * (2 * N + 8 * NW + 8)));
  X.paths = d_strings;
  CUDA_TRY(ctx, cudaMemcpyAsync(d_win, h, win_bytes, cudaMemcpyHostToDevice, ctx->stream));
  if (custom_ref) {
    CUDA_TRY(ctx, cudaMemsetAsync(d_ref1, 0, 1, ctx->stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_ref1 + 1, custom_ref, (size_t)P.ref_len, cudaMemcpyHostToDevice, ctx->stream));
  } else {
    asm_refwin_kernel<<<blocks_for((int64_t)n * ref_stride, 256), 256, 0, ctx->stream>>>(ctx->d_refs.as<RefView>(), d_win, n, P.W, d_ref1, d_err);
    ctx->launches += 1;
  }
  asm_nodes_kernel<<<(unsigned)n, 256, 0, ctx->stream>>>(ctx->R, d_win, T, G, P);
  asm_walk_kernel<<<blocks_for((int64_t)n * 32, 128), 128, 0, ctx->stream>>>(d_win, n, T, G, X, P, d_ref1, d_err);
  ctx->launches += 2;
  CUDA_TRY(ctx, cudaGetLastError());
  st.ctx = ctx;
  {
    void* hp = nullptr;
    GUAC_TRY(pinned_alloc(ctx, &hp, (size_t)path_bytes + (size_t)n * ref_stride + 16));
    st.strings = static_cast<uint8_t*>(hp);
  }
  CUDA_TRY(ctx, cudaMemcpyAsync(h, d_win, win_bytes, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(st.strings, d_strings, (size_t)path_bytes + (size_t)n * ref_stride, cudaMemcpyDeviceToHost, ctx->stream));
  GUAC_TRY(check_device_error(ctx));
  return GUAC_OK;
}

AsmPenalties asm_penalties(double mismatch_probability, double open_gap_probability, double close_gap_probability) {  // ReadAlignment.scala:98-104
  AsmPenalties Q;
  Q.mismatch = -std::log(mismatch_probability);
  Q.open_gap = -std::log(open_gap_probability);
  Q.no_gap = -std::log(1 - open_gap_probability);
  Q.close_gap = -std::log(close_gap_probability);
  Q.continue_gap = -std::log(1 - close_gap_probability);
  return Q;
}

int asm_align_prepare(guac_ctx* ctx) {
  (void)ctx;  // the kernel's shared-memory attribute is set per device in guac_ctx_create
  return GUAC_OK;
}

// Assembles windows [w0, w1) of `windows`; results appended to `results`.
int assemble_chunk(guac_ctx* ctx, const guac_ga_params* params, std::vector<AsmWindow>& windows, size_t w0, size_t w1,
                   std::vector<AsmWindowResult>& results) {
  AsmStage st;
  const int n = (int)(w1 - w0);
  const int K = params->kmer_size, W = params->assembly_window_range;
  st.P = AsmParams{params->sample_id, K, W, std::max(params->min_occurrence, 0) /* counts are >= 1: a threshold below 0 is 0 */, params->min_mean_kmer_quality, ASM_MAX_PATHS, 2 * W, 0};
  AsmWindow* h = windows.data() + w0;
  GUAC_TRY(build_paths(ctx, st, h, n, nullptr));
  DevTmp& tmp = st.tmp;
  const AsmParams P = st.P;
  AsmWindow* d_win = st.d_win;
  uint8_t* d_strings = st.d_strings;
  const int64_t path_bytes = st.path_bytes;
  const int ref_stride = st.ref_stride;
  const uint8_t* strings = st.strings;
  DeviceError* d_err = ctx->b_err.as<DeviceError>();
  // discoverHaplotypes (AssemblyUtils.scala:79-114): score when there are more paths than the ploidy
  const int expected_ploidy = 2, max_paths_to_score = ASM_MAX_PATHS - 1;
  std::vector<int32_t> which, sum_base;
  std::vector<int64_t> job_base;
  int64_t n_jobs = 0;
  int32_t n_sums = 0;
  int max_S = 0;
  for (int i = 0; i < n; ++i)
    if (h[i].n_paths > expected_ploidy && h[i].n_paths <= max_paths_to_score) {
      which.push_back(i);
      job_base.push_back(n_jobs);
      sum_base.push_back(n_sums);
      n_jobs += (int64_t)h[i].n_reads * h[i].n_paths;
      n_sums += h[i].n_paths;
    }
  const AsmPenalties Q = asm_penalties(std::exp(-4.0), std::exp(-6.0), 1 - std::exp(-1.0));  // ReadAlignment.scala:64-70 defaults
  GUAC_TRY(asm_align_prepare(ctx));
  std::vector<long long> sums((size_t)n_sums, 0);
  if (n_jobs > 0) {
    int32_t* d_which = nullptr;
    int32_t* d_sum_base = nullptr;
    int64_t* d_job_base = nullptr;
    AsmAlignJob* d_jobs = nullptr;
    long long* d_sums = nullptr;
    GUAC_TRY(tmp.alloc(ctx, &d_which, which.size()));
    GUAC_TRY(tmp.alloc(ctx, &d_sum_base, which.size()));
    GUAC_TRY(tmp.alloc(ctx, &d_job_base, which.size()));
    GUAC_TRY(tmp.alloc(ctx, &d_jobs, (size_t)n_jobs));
    GUAC_TRY(tmp.alloc(ctx, &d_sums, (size_t)n_sums, 0));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_which, which.data(), which.size() * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_sum_base, sum_base.data(), which.size() * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_job_base, job_base.data(), which.size() * sizeof(int64_t), cudaMemcpyHostToDevice, ctx->stream));
    asm_score_jobs_kernel<<<blocks_for((int64_t)which.size(), 64), 64, 0, ctx->stream>>>(ctx->R, d_win, d_which, d_job_base, d_sum_base,
                                                                                          (int)which.size(), P, d_jobs);
    for (int i : which) max_S = std::max(max_S, h[i].max_read_len);  // reads are the sequences here
    if (max_S > ASM_ALIGN_MAX_S)
      return fail(ctx, GUAC_E_UNSUPPORTED, "a read in an assembly window is longer than " + std::to_string(ASM_ALIGN_MAX_S) + " bases (library limit)");
    asm_align_kernel<<<(unsigned)n_jobs, ASM_ALIGN_THREADS, asm_align_smem(max_S), ctx->stream>>>(ctx->R, d_jobs, Q, d_strings, nullptr, d_sums,
                                                                                                  nullptr, d_err);
    ctx->launches += 2;
    CUDA_TRY(ctx, cudaGetLastError());
    CUDA_TRY(ctx, cudaMemcpyAsync(sums.data(), d_sums, (size_t)n_sums * sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
    GUAC_TRY(check_device_error(ctx));
  }
  // chosen haplotypes -> alignment against the window's reference with trace-back
  struct Chosen {
    int window, path;
  };
  std::vector<Chosen> chosen;
  std::vector<int> n_final((size_t)n, 0);
  {
    size_t si = 0;
    for (int i = 0; i < n; ++i) {
      const int np = h[i].n_paths;
      if (np == 0 || np > max_paths_to_score) continue;
      // the reference lists paths most recent first (`paths ::= ...`); the device emits them in discovery order
      std::vector<int> order;
      for (int p = np - 1; p >= 0; --p) order.push_back(p);
      if (np > expected_ploidy) {
        const long long* s = sums.data() + sum_base[si++];
        // sortBy(-_._1) over (sum of -alignmentScore, path): stable, largest sum of -score first (:97-103)
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return -s[a] < -s[b]; });
        order.resize(expected_ploidy);
      }
      for (int p : order) chosen.push_back({i, p});
      n_final[i] = (int)order.size();
    }
  }
  std::vector<AsmCigar> cigars(chosen.size());
  if (!chosen.empty()) {
    std::vector<AsmAlignJob> jobs(chosen.size());
    int64_t head_bytes = 0;
    int longest = 0;
    for (size_t c = 0; c < chosen.size(); ++c) {
      const AsmWindow& w = h[chosen[c].window];
      AsmAlignJob& j = jobs[c];
      j.seq = w.path_off + (int64_t)chosen[c].path * w.path_cap;
      j.ref = path_bytes + (int64_t)chosen[c].window * ref_stride + 1;
      j.seq_len = w.path_len[chosen[c].path];
      j.ref_len = 2 * W;
      j.seq_kind = 1;
      j.sum_slot = -1;
      j.trace = (int32_t)c;
      j._pad = 0;
      j.head_off = head_bytes;
      head_bytes += (int64_t)j.seq_len * (j.ref_len + 1);
      longest = std::max(longest, j.seq_len);
    }
    if (longest > ASM_ALIGN_MAX_S)
      return fail(ctx, GUAC_E_UNSUPPORTED, "an assembled haplotype is longer than " + std::to_string(ASM_ALIGN_MAX_S) + " bases (library limit)");
    AsmAlignJob* d_jobs = nullptr;
    uint8_t* d_heads = nullptr;
    AsmCigar* d_cigars = nullptr;
    GUAC_TRY(tmp.alloc(ctx, &d_jobs, jobs.size()));
    GUAC_TRY(tmp.alloc(ctx, &d_heads, (size_t)head_bytes));
    GUAC_TRY(tmp.alloc(ctx, &d_cigars, jobs.size()));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_jobs, jobs.data(), jobs.size() * sizeof(AsmAlignJob), cudaMemcpyHostToDevice, ctx->stream));
    asm_align_kernel<<<(unsigned)jobs.size(), ASM_ALIGN_THREADS, asm_align_smem(longest), ctx->stream>>>(ctx->R, d_jobs, Q, d_strings, d_heads,
                                                                                                         nullptr, d_cigars, d_err);
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    CUDA_TRY(ctx, cudaMemcpyAsync(cigars.data(), d_cigars, jobs.size() * sizeof(AsmCigar), cudaMemcpyDeviceToHost, ctx->stream));
    GUAC_TRY(check_device_error(ctx));
  }
  std::vector<std::set<AsmVariant>> sets((size_t)n);
  for (size_t c = 0; c < chosen.size(); ++c) {
    const int i = chosen[c].window;
    const AsmWindow& w = h[i];
    const std::string path((const char*)strings + w.path_off + (int64_t)chosen[c].path * w.path_cap, (size_t)w.path_len[chosen[c].path]);
    const std::string ref1((const char*)strings + path_bytes + (int64_t)i * ref_stride, (size_t)ref_stride);
    GUAC_TRY(variants_from_cigar(ctx, cigars[c], path, ref1, w.locus - W, w.contig_id, sets[(size_t)i]));
  }
  // window-wide evidence, only where variants came out
  std::vector<int32_t> with_variants;
  for (int i = 0; i < n; ++i)
    if (!sets[(size_t)i].empty()) with_variants.push_back(i);
  std::vector<AsmEvidence> ev((size_t)n);
  if (!with_variants.empty()) {
    int32_t* d_which = nullptr;
    AsmEvidence* d_ev = nullptr;
    const size_t m = with_variants.size();
    std::vector<AsmEvidence> packed(m);
    GUAC_TRY(tmp.alloc(ctx, &d_which, m));
    GUAC_TRY(tmp.alloc(ctx, &d_ev, m));
    CUDA_TRY(ctx, cudaMemcpyAsync(d_which, with_variants.data(), m * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
    asm_evidence_kernel<<<blocks_for((int64_t)m, 4), 128, 0, ctx->stream>>>(ctx->R, d_win, d_which, (int)m, P, d_ev);
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    CUDA_TRY(ctx, cudaMemcpyAsync(packed.data(), d_ev, m * sizeof(AsmEvidence), cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    for (size_t j = 0; j < m; ++j) ev[(size_t)with_variants[j]] = packed[j];
  }
  for (int i = 0; i < n; ++i) {
    AsmWindowResult r;
    r.contig_id = h[i].contig_id;
    r.locus = h[i].locus;
    r.n_nodes = h[i].n_nodes;
    r.n_paths = n_final[(size_t)i];
    r.variants.assign(sets[(size_t)i].begin(), sets[(size_t)i].end());
    r.evidence = ev[(size_t)i];
    results.push_back(std::move(r));
  }
  return GUAC_OK;
}

}  // namespace

int guac_germline_assembly(guac_ctx* ctx, const guac_ga_params* params, const guac_range* loci, int32_t n_ranges,
                           guac_called_allele** out_calls, int64_t* n_calls, char** out_arena, int64_t* arena_len,
                           guac_assembled_window** out_windows, int64_t* n_windows, int64_t* out_n_visited) {
  if (!ctx || !params || !out_calls || !n_calls || !out_arena || !arena_len || !out_windows || !n_windows) return GUAC_E_INVALID_INPUT;
  const int K = params->kmer_size, W = params->assembly_window_range;
  if (K < 2 || K > ASM_MAX_K)
    return fail(ctx, GUAC_E_UNSUPPORTED, "kmer_size must be in [2, " + std::to_string(ASM_MAX_K) + "] (library limit: 3-bit packed k-mers in four words)");
  if (W < K)
    return fail(ctx, GUAC_E_INVALID_INPUT,
                "assembly_window_range must be >= kmer_size: after an assembly the reference moves its window to locus + range - kmerSize, "
                "and \"Pileup window can only move forward in locus\"");
  // 1. candidate loci: the dense pass and the toy caller of the shortcut branch (:135-153)
  guac_gs_params gs;
  gs.sample_id = params->sample_id;
  gs.min_alt_reads = params->min_alt_reads;
  gs.min_phred_scaled_likelihood = params->min_phred_scaled_likelihood;
  gs.respect_active_regions = 1;
  gs.min_area_vaf = params->shortcut_assembly ? params->min_area_vaf : -1.0f;  // without the shortcut every candidate locus is assembled
  gs._pad = 0;
  guac_called_allele* toy = nullptr;
  char* toy_arena = nullptr;
  int64_t* active = nullptr;
  int64_t n_toy = 0, toy_arena_len = 0, n_active = 0, n_visited = 0;
  GUAC_TRY(guac_germline_standard(ctx, &gs, loci, n_ranges, &toy, &n_toy, &toy_arena, &toy_arena_len, &active, &n_active, &n_visited));
  std::vector<guac_called_allele> toy_calls(toy, toy + n_toy);
  const std::string toy_strings(toy_arena, toy_arena + toy_arena_len);
  std::vector<AsmWindow> windows((size_t)n_active);
  for (int64_t i = 0; i < n_active; ++i) {
    AsmWindow& w = windows[(size_t)i];
    std::memset(&w, 0, sizeof w);
    w.contig_id = (int32_t)active[2 * i];
    w.locus = active[2 * i + 1];
    for (auto& sp : ctx->contig_spans)
      if (sp.first == w.contig_id) {
        w.read_clo = sp.second.first;
        w.read_chi = sp.second.second;
      }
    w.ref_index = ctx->contigs[w.contig_id].ref_index;
  }
  guac_free(ctx, toy);
  guac_free(ctx, toy_arena);
  guac_free(ctx, active);
  // 2. every candidate window assembled on the device
  std::vector<AsmWindowResult> results;
  const size_t chunk = 4096;
  cudaEventRecord(ctx->ev0, ctx->stream);
  for (size_t w0 = 0; w0 < windows.size(); w0 += chunk)
    GUAC_TRY(assemble_chunk(ctx, params, windows, w0, std::min(windows.size(), w0 + chunk), results));
  cudaEventRecord(ctx->ev1, ctx->stream);
  ctx->have_timing = true;
  ctx->timed_loci = (int64_t)windows.size();
  // 3. the reference's walk: lastCalledLocus filter and the window jump (:195-213), contig by contig
  std::vector<int32_t> contig_order;
  for (int i = 0; i < n_ranges; ++i)
    if (std::find(contig_order.begin(), contig_order.end(), loci[i].contig_id) == contig_order.end()) contig_order.push_back(loci[i].contig_id);
  std::vector<guac_called_allele> out;
  std::vector<guac_assembled_window> out_w;
  std::string arena;
  auto put = [&](const std::string& s, uint32_t& off, uint32_t& len) {
    off = (uint32_t)arena.size();
    len = (uint32_t)s.size();
    arena += s;
  };
  for (int32_t contig : contig_order) {
    size_t ti = 0, wi = 0;
    while (ti < toy_calls.size() && toy_calls[ti].contig_id != contig) ++ti;
    while (wi < results.size() && results[wi].contig_id != contig) ++wi;
    bool have_last = false;
    int64_t last_called = 0, skip_until = -1;
    while (true) {
      const bool has_t = ti < toy_calls.size() && toy_calls[ti].contig_id == contig;
      const bool has_w = wi < results.size() && results[wi].contig_id == contig;
      if (!has_t && !has_w) break;
      if (has_t && (!has_w || toy_calls[ti].start < results[wi].locus)) {  // toy-called locus
        const int64_t locus = toy_calls[ti].start;
        const bool visited = locus > skip_until;
        for (; ti < toy_calls.size() && toy_calls[ti].contig_id == contig && toy_calls[ti].start == locus; ++ti) {
          if (!visited) continue;
          guac_called_allele g = toy_calls[ti];
          put(toy_strings.substr(g.ref_off, g.ref_len), g.ref_off, g.ref_len);
          put(toy_strings.substr(g.alt_off, g.alt_len), g.alt_off, g.alt_len);
          out.push_back(g);
          have_last = true;
          last_called = locus;
        }
        continue;
      }
      const AsmWindowResult& r = results[wi++];
      guac_assembled_window aw;
      aw.contig_id = r.contig_id;
      aw.n_paths = r.n_paths;
      aw.locus = r.locus;
      aw.n_graph_nodes = r.n_nodes;
      aw.consumed = r.locus > skip_until;
      out_w.push_back(aw);
      if (!aw.consumed || r.n_paths == 0) continue;
      int64_t new_last = last_called;
      bool any = false;
      for (const AsmVariant& v : r.variants) {
        if (have_last && !(last_called < v.locus)) continue;
        guac_called_allele g;
        std::memset(&g, 0, sizeof g);
        g.contig_id = contig;
        g.n_genotypes = 0;
        g.start = v.locus;
        put(v.ref, g.ref_off, g.ref_len);
        put(v.alt, g.alt_off, g.alt_len);
        g.genotype_log_probability = 0.0;
        g.evidence.probability = 1.0;
        g.evidence.read_depth = g.evidence.allele_read_depth = g.evidence.forward_depth = g.evidence.allele_forward_depth = r.evidence.depth;
        g.evidence.mean_mapping_quality = r.evidence.mean_mapq;
        g.evidence.median_mapping_quality = r.evidence.median_mapq;
        g.evidence.mean_base_quality = r.evidence.mean_baseq;
        g.evidence.median_base_quality = r.evidence.median_baseq;
        g.evidence.median_mismatches_per_read = 0.0;
        out.push_back(g);
        new_last = any ? std::max(new_last, v.locus) : v.locus;
        any = true;
      }
      if (any) {
        have_last = true;
        last_called = new_last;
      }
      skip_until = r.locus + W - K;
    }
  }
  void* hc = nullptr;
  void* hw = nullptr;
  void* ha = nullptr;
  GUAC_TRY(pinned_alloc(ctx, &hc, std::max<size_t>(out.size() * sizeof(guac_called_allele), 16)));
  GUAC_TRY(pinned_alloc(ctx, &hw, std::max<size_t>(out_w.size() * sizeof(guac_assembled_window), 16)));
  GUAC_TRY(pinned_alloc(ctx, &ha, std::max<size_t>(arena.size(), 16)));
  if (!out.empty()) std::memcpy(hc, out.data(), out.size() * sizeof(guac_called_allele));
  if (!out_w.empty()) std::memcpy(hw, out_w.data(), out_w.size() * sizeof(guac_assembled_window));
  if (!arena.empty()) std::memcpy(ha, arena.data(), arena.size());
  *out_calls = (guac_called_allele*)hc;
  *n_calls = (int64_t)out.size();
  *out_windows = (guac_assembled_window*)hw;
  *n_windows = (int64_t)out_w.size();
  *out_arena = (char*)ha;
  *arena_len = (int64_t)arena.size();
  if (out_n_visited) *out_n_visited = n_visited;
  return GUAC_OK;
}

int guac_discover_paths(guac_ctx* ctx, int32_t sample_id, const char* reference_sequence, int32_t reference_len, int32_t kmer_size,
                        int32_t min_occurrence, int32_t max_paths, int32_t min_mean_kmer_quality, char** out_paths, int64_t* out_len,
                        int32_t* n_paths) {
  if (!ctx || !reference_sequence || !out_paths || !out_len || !n_paths) return GUAC_E_INVALID_INPUT;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  if (!ctx->have_reads) return fail(ctx, GUAC_E_STATE, "no reads uploaded (guac_reads_upload)");
  if (kmer_size < 2 || kmer_size > ASM_MAX_K)
    return fail(ctx, GUAC_E_UNSUPPORTED, "kmer_size must be in [2, " + std::to_string(ASM_MAX_K) + "] (library limit)");
  if (max_paths < 1 || max_paths > ASM_MAX_PATHS)
    return fail(ctx, GUAC_E_UNSUPPORTED, "max_paths must be in [1, " + std::to_string(ASM_MAX_PATHS) + "] (library limit)");
  if (reference_len < kmer_size) return fail(ctx, GUAC_E_INVALID_INPUT, "Source kmer has size != kmerSize");  // DeBruijnGraph.scala:190-191
  GUAC_TRY(reset_error(ctx));
  AsmStage st;
  st.P = AsmParams{sample_id, kmer_size, 0, std::max(min_occurrence, 0), min_mean_kmer_quality, max_paths, reference_len, 1};
  AsmWindow w;
  std::memset(&w, 0, sizeof w);
  w.read_clo = 0;
  w.read_chi = ctx->n_reads;
  GUAC_TRY(build_paths(ctx, st, &w, 1, reinterpret_cast<const uint8_t*>(reference_sequence)));
  std::string text;
  for (int p = w.n_paths - 1; p >= 0; --p) {  // the reference prepends each new path to its list
    text.append((const char*)st.strings + w.path_off + (int64_t)p * w.path_cap, (size_t)w.path_len[p]);
    text.push_back('\0');
  }
  void* h = nullptr;
  GUAC_TRY(pinned_alloc(ctx, &h, std::max<size_t>(text.size(), 16)));
  std::memcpy(h, text.data(), text.size());
  *out_paths = (char*)h;
  *out_len = (int64_t)text.size();
  *n_paths = w.n_paths;
  return GUAC_OK;
}

int guac_read_alignment(guac_ctx* ctx, const char* sequence, int32_t sequence_len, const char* reference, int32_t reference_len,
                        double mismatch_probability, double open_gap_probability, double close_gap_probability, char* out_cigar,
                        int32_t cigar_cap, int32_t* ref_start, int32_t* ref_end, int32_t* alignment_score, double* raw_score) {
  if (!ctx || !sequence || !reference || !out_cigar || cigar_cap < 1 || sequence_len < 0 || reference_len < 0) return GUAC_E_INVALID_INPUT;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  if (sequence_len > ASM_ALIGN_MAX_S || sequence_len > 65535 || reference_len > 65535)
    return fail(ctx, GUAC_E_UNSUPPORTED, "sequence longer than " + std::to_string(ASM_ALIGN_MAX_S) + " or reference longer than 65535 bases (library limit)");
  out_cigar[0] = 0;
  GUAC_TRY(reset_error(ctx));
  GUAC_TRY(asm_align_prepare(ctx));
  const AsmPenalties Q = asm_penalties(mismatch_probability > 0 ? mismatch_probability : std::exp(-4.0),
                                       open_gap_probability > 0 ? open_gap_probability : std::exp(-6.0),
                                       close_gap_probability > 0 ? close_gap_probability : 1 - std::exp(-1.0));
  DevTmp tmp;
  uint8_t* d_strings = nullptr;
  uint8_t* d_heads = nullptr;
  AsmAlignJob* d_job = nullptr;
  AsmCigar* d_cigar = nullptr;
  GUAC_TRY(tmp.alloc(ctx, &d_strings, (size_t)sequence_len + reference_len + 16));
  GUAC_TRY(tmp.alloc(ctx, &d_heads, (size_t)sequence_len * (reference_len + 1) + 16));
  GUAC_TRY(tmp.alloc(ctx, &d_job, 1));
  GUAC_TRY(tmp.alloc(ctx, &d_cigar, 1, 0));
  AsmAlignJob job;
  std::memset(&job, 0, sizeof job);
  job.seq = 0;
  job.ref = sequence_len;
  job.seq_len = sequence_len;
  job.ref_len = reference_len;
  job.seq_kind = 1;
  job.sum_slot = -1;
  job.trace = 0;
  job.head_off = 0;
  if (sequence_len) CUDA_TRY(ctx, cudaMemcpyAsync(d_strings, sequence, (size_t)sequence_len, cudaMemcpyHostToDevice, ctx->stream));
  if (reference_len) CUDA_TRY(ctx, cudaMemcpyAsync(d_strings + sequence_len, reference, (size_t)reference_len, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(d_job, &job, sizeof job, cudaMemcpyHostToDevice, ctx->stream));
  AsmCigar c;
  std::memset(&c, 0, sizeof c);
  if (sequence_len > 0) {
    asm_align_kernel<<<1, ASM_ALIGN_THREADS, asm_align_smem(sequence_len), ctx->stream>>>(ctx->R, d_job, Q, d_strings, d_heads, nullptr, d_cigar,
                                                                                        ctx->b_err.as<DeviceError>());
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    CUDA_TRY(ctx, cudaMemcpyAsync(&c, d_cigar, sizeof c, cudaMemcpyDeviceToHost, ctx->stream));
    GUAC_TRY(check_device_error(ctx));
  }
  std::string text;  // toCigarString (:22-43)
  for (int i = 0; i < c.n_elements; ++i) text += std::to_string(c.len[i]) + "=XID"[c.op[i]];
  if ((int)text.size() + 1 > cigar_cap) return fail(ctx, GUAC_E_INVALID_INPUT, "cigar buffer too small");
  std::memcpy(out_cigar, text.c_str(), text.size() + 1);
  if (ref_start) *ref_start = c.ref_start;
  if (ref_end) *ref_end = c.ref_end;
  if (alignment_score) *alignment_score = c.score;
  if (raw_score) *raw_score = c.raw_score;
  return GUAC_OK;
}
