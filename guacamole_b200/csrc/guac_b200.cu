// guac_b200.cu — C ABI (include/guac_b200.h) and host-side planning for libguac_b200.so.
// Everything that touches per-locus data runs in the CUDA kernels of kernels.cuh /
// tile_kernels.cuh; this file owns device memory, the stream, error translation and the
// sparse result assembly.  There is no CPU fallback.
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <set>
#include <string>
#include <tuple>
#include <thread>
#include <vector>

#include "../../include/guac_b200.h"
#include "kernels.cuh"
#include "scan.cuh"
#include "synth_kernels.cuh"
#include "tile_kernels.cuh"
#include "stream_kernels.cuh"
#include "joint_stream_kernels.cuh"
#include "germline_kernels.cuh"
#include "assembly_kernels.cuh"
#include "compact_kernels.cuh"
#include "support_kernels.cuh"
#include "partition_kernels.cuh"

using namespace guac;

namespace {

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  template <class T>
  T* as() const { return reinterpret_cast<T*>(p); }
};

// A device buffer that lives for one call: released on every return path (cudaFree waits for the device).
struct TmpDev : DevBuf {
  TmpDev() = default;
  TmpDev(const TmpDev&) = delete;
  TmpDev& operator=(const TmpDev&) = delete;
  ~TmpDev() {
    if (p) cudaFree(p);
  }
};

struct ContigDev {
  int64_t offset = 0, len = 0;
  uint8_t* bases = nullptr;
  int ref_index = -1;
  uint32_t* ref2 = nullptr;  // 2-bit copy for the stream path (16 loci per word), rebuilt when the window changes
  uint32_t* refbad = nullptr;  // one bit per 16 loci: the group holds a base that is not A C G T (or is missing)
  size_t ref2_words = 0;
  bool ref2_dirty = true;
};

}  // namespace

struct guac_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  std::string err;
  int64_t launches = 0;
  std::map<int32_t, ContigDev> contigs;
  DevBuf d_refs;  // RefView table
  bool refs_dirty = true;
  // reads
  bool have_reads = false;
  int64_t n_reads = 0;
  int32_t n_samples = 0;
  ReadsView R{};
  DevBuf b_start, b_end, b_key, b_runmax, b_contig, b_cigar_off, b_cigar, b_seq_off, b_bases2, b_nmask, b_exc_pos, b_exc_base,
      b_qual, b_mapq, b_flags, b_sample, b_scan_tmp;
  std::vector<std::pair<int32_t, std::pair<int64_t, int64_t>>> contig_spans;  // contig -> [lo, hi) read indices
  // plan / scratch
  DevBuf b_ranges, b_tiles, b_visited, b_vis_idx, b_counts, b_totals, b_elem_off, b_summary, b_calls, b_ncalls, b_err, b_arena;
  std::map<void*, size_t> host_allocs;            // live pinned output buffers -> capacity
  std::multimap<size_t, void*> pinned_free;        // released buffers kept for reuse
  size_t pinned_free_bytes = 0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  bool have_timing = false;
  int64_t timed_loci = 0;
  DevBuf b_nops, b_rflags, b_logtab, b_fallback;
  int64_t opt_ss_impl = 2;            // 2 = stream (read pass + locus pass; the default), 0 = single tile kernel, 1 = gather kernel only
  int64_t opt_force_fallback_mod = 0; // test hook: tiles with index % k == 0 are sent to the fallback path
  int64_t opt_ss_rank_max = SS_RANK_MAX;  // test hook: more calls than this are put in order by the host instead of the device
  int64_t last_fallback_tiles = 0, last_heavy_loci = 0;
  DevBuf b_heavy, b_boundtab, b_vcount, b_sj_sites, b_arena_used, b_hdr;
  double bound_tab_vaf = -1;
  cudaStream_t stream2 = nullptr;  // second stream: the copy of the bases in guac_reads_upload_compact
  cudaEvent_t fork = nullptr, join = nullptr, qual_ready = nullptr;
  DevBuf b_ev_overflow;  // stream path: overflow arena of the read pass (first 64 bytes: its counter)
  DevBuf b_bins, b_bin_count, b_medium;  // stream path: per-read event slots and counts; list of multi-operator reads
  int64_t n_medium = 0;
  // every READ_INDEX_STRIDE-th entry of the start column, of the running-max end keys and of the multi-operator list, on the
  // host: a call over part of the batch's loci runs the read pass over the reads that can overlap them only
  std::vector<int64_t> h_idx_start, h_idx_runmax, h_idx_medium;
  DevBuf u_len, u_ops, u_runs_c, u_runs_f, u_tmp, u_qpacked, u_bad, u_flag, u_off, u_idx, u_expl, u_mm_count, u_mm_words, u_expl_len, u_mm_n, u_expl_off, u_mm_off;  // upload scratch, kept between calls
  DevBuf u_start_rel, u_runs_base, u_ops8, u_ops_eff, u_wire_off, u_wire;          // ... of the narrow columns of the wire form
  int64_t opt_compact_narrow = 1;
  int64_t opt_compact_ref_relative = 0;  // guac_reads_export_compact: reads as differences from the reference where they can
  int64_t opt_compact_min_qual_bits = 2;  // guac_reads_export_compact: 2 (default) the narrowest form the qualities allow; 4 / 8 force wider ones
  int64_t opt_sj_host_threads = 0;  // 0: as many as the host has cores, at most 32
  std::vector<RangeDev> plan_ranges;            // host copy of the current call's range table (make_plan)
  DevBuf r_bitmap, r_counts, r_prefix, r_rank;  // rank_by_slot
  DevBuf g_active, g_jobs, g_evs, g_strings;  // scratch of guac_germline_standard, kept between calls
  // somatic-joint: force-call ranges, simple candidates, the fast kernel's sites / raw records / hand-backs, the generic
  // kernels' raw records, emitted sites and their order, offsets, finished records, ln n! table
  DevBuf j_force, j_simple, j_sites_fast, j_raw_fast, j_bail, j_raw_gen, j_emits, j_order, j_cnt, j_off, j_out_alleles, j_out_samples, j_logfact, j_logtab;
  int64_t opt_synth_profile = 0;  // guac_synth_generate: 0 the SURVEY 8(d) workload, 1 its "realistic" variant (synth_model.h)
  int64_t opt_sj_prune = 1;  // locus pass drops simple sites that certainly hold no call (test hook: 0 sends them all to the site kernel)
  int64_t opt_sj_impl = 2;  // 2: read pass + classifying locus pass + fast site kernel (default); 1: single-kernel locus pass; 0: no fast site kernel
  int64_t stat_sj_simple = 0, stat_sj_bail = 0, stat_sj_sites = 0, stat_sj_emitted_sites = 0;
  void* h_counters = nullptr;  // pinned: the per-call counter block comes back without a staging copy
  std::vector<std::pair<void*, size_t>> asm_blocks;  // assembly scratch, bump-allocated per call (assembly_entry.inl)
  size_t asm_block = 0, asm_off = 0;
};

namespace {

#define CUDA_TRY(ctx, expr)                                                                       \
  do {                                                                                            \
    cudaError_t _e = (expr);                                                                      \
    if (_e != cudaSuccess) {                                                                      \
      (ctx)->err = std::string("CUDA error: ") + cudaGetErrorString(_e) + " at " #expr;           \
      return GUAC_E_CUDA;                                                                         \
    }                                                                                             \
  } while (0)

#define GUAC_TRY(expr)          \
  do {                          \
    int _s = (expr);            \
    if (_s != GUAC_OK) return _s; \
  } while (0)

// GUAC_TRACE=1: host-side timestamps of the phases of a call, on stderr
struct Trace {
  bool on;
  std::chrono::steady_clock::time_point t0, last;
  Trace() : on(std::getenv("GUAC_TRACE") != nullptr), t0(std::chrono::steady_clock::now()), last(t0) {}
  void mark(const char* what) {
    if (!on) return;
    auto now = std::chrono::steady_clock::now();
    std::fprintf(stderr, "[guac trace] %-28s +%8.3f ms (total %8.3f)\n", what,
                 std::chrono::duration<double, std::milli>(now - last).count(),
                 std::chrono::duration<double, std::milli>(now - t0).count());
    last = now;
  }
};

int fail(guac_ctx* ctx, int status, const std::string& msg) {
  ctx->err = msg;
  return status;
}

int ensure(guac_ctx* ctx, DevBuf& b, size_t bytes) {
  if (bytes <= b.cap && b.p) return GUAC_OK;
  if (b.p) CUDA_TRY(ctx, cudaFree(b.p));
  b.p = nullptr;
  b.cap = 0;
  size_t want = std::max<size_t>(bytes, 256);
  cudaError_t e = cudaMalloc(&b.p, want);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    return fail(ctx, GUAC_E_OOM, std::string("cudaMalloc of ") + std::to_string(want) + " bytes failed: " + cudaGetErrorString(e));
  }
  b.cap = want;
  return GUAC_OK;
}

template <class T>
int upload(guac_ctx* ctx, DevBuf& b, const T* host, size_t n, size_t extra_elems = 0) {
  GUAC_TRY(ensure(ctx, b, (n + extra_elems) * sizeof(T) + 16));
  if (n) CUDA_TRY(ctx, cudaMemcpyAsync(b.p, host, n * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
  return GUAC_OK;
}

// Pinned host memory for outputs comes from a per-context pool: cudaMallocHost / cudaFreeHost cost
// about a millisecond each, more than a whole kernel pass here.
int pinned_alloc(guac_ctx* ctx, void** out, size_t bytes) {
  size_t want = 4096;
  while (want < bytes) want <<= 1;
  auto it = ctx->pinned_free.lower_bound(want);
  if (it != ctx->pinned_free.end() && it->first <= want * 2) {
    *out = it->second;
    ctx->host_allocs[*out] = it->first;
    ctx->pinned_free_bytes -= it->first;
    ctx->pinned_free.erase(it);
    return GUAC_OK;
  }
  void* h = nullptr;
  CUDA_TRY(ctx, cudaMallocHost(&h, want));
  ctx->host_allocs[h] = want;
  *out = h;
  return GUAC_OK;
}

template <class T>
int host_out(guac_ctx* ctx, T** out, const void* dev, size_t n) {
  void* h = nullptr;
  GUAC_TRY(pinned_alloc(ctx, &h, std::max<size_t>(n * sizeof(T), 16)));
  if (n) CUDA_TRY(ctx, cudaMemcpyAsync(h, dev, n * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
  *out = reinterpret_cast<T*>(h);
  return GUAC_OK;
}

inline unsigned blocks_for(int64_t n, int threads) { return (unsigned)((n + threads - 1) / threads); }

// device-wide scan (three phases), T = int64_t
template <bool EXCLUSIVE, class TIn, class Op>
int device_scan(guac_ctx* ctx, const TIn* in, int64_t* out, int64_t n, Op op, int64_t* grand_total_dev = nullptr) {
  if (n == 0) return GUAC_OK;
  int64_t n_tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  GUAC_TRY(ensure(ctx, ctx->b_scan_tmp, (size_t)(n_tiles + 1) * sizeof(int64_t)));
  int64_t* totals = ctx->b_scan_tmp.as<int64_t>();
  scan_reduce_kernel<TIn, int64_t, Op><<<(unsigned)n_tiles, SCAN_THREADS, 0, ctx->stream>>>(in, totals, n, op);
  scan_totals_kernel<int64_t, Op><<<1, SCAN_THREADS, 0, ctx->stream>>>(totals, n_tiles, op);
  scan_apply_kernel<EXCLUSIVE, TIn, int64_t, Op><<<(unsigned)n_tiles, SCAN_THREADS, 0, ctx->stream>>>(in, out, totals, n, op);
  ctx->launches += 3;
  CUDA_TRY(ctx, cudaGetLastError());
  if (grand_total_dev)
    CUDA_TRY(ctx, cudaMemcpyAsync(grand_total_dev, totals + n_tiles, sizeof(int64_t), cudaMemcpyDeviceToDevice, ctx->stream));
  return GUAC_OK;
}

// rank[i] = position of record i among the n records in slot order (slots distinct, < n_slots); see slot_bitmap_*_kernel
int rank_by_slot(guac_ctx* ctx, const void* d_records, size_t stride, size_t slot_offset, int n, int64_t n_slots, int* d_rank) {
  if (n == 0) return GUAC_OK;
  const int64_t n_words = (n_slots + 31) / 32;
  GUAC_TRY(ensure(ctx, ctx->r_bitmap, (size_t)n_words * sizeof(unsigned int)));
  GUAC_TRY(ensure(ctx, ctx->r_counts, (size_t)n_words + 16));
  GUAC_TRY(ensure(ctx, ctx->r_prefix, (size_t)(n_words + 1) * sizeof(int64_t)));
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->r_bitmap.p, 0, (size_t)n_words * sizeof(unsigned int), ctx->stream));
  const unsigned char* rec = static_cast<const unsigned char*>(d_records);
  slot_bitmap_set_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(rec, stride, slot_offset, n, ctx->r_bitmap.as<unsigned int>());
  slot_bitmap_popc_kernel<<<blocks_for(n_words, 256), 256, 0, ctx->stream>>>(ctx->r_bitmap.as<unsigned int>(), n_words, ctx->r_counts.as<uint8_t>());
  GUAC_TRY((device_scan<true, uint8_t>(ctx, ctx->r_counts.as<uint8_t>(), ctx->r_prefix.as<int64_t>(), n_words, OpSum())));
  slot_bitmap_rank_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(rec, stride, slot_offset, n, ctx->r_bitmap.as<unsigned int>(),
                                                                       ctx->r_prefix.as<int64_t>(), d_rank);
  ctx->launches += 6;
  CUDA_TRY(ctx, cudaGetLastError());
  return GUAC_OK;
}

int sync_ref2(guac_ctx* ctx);
cudaError_t configure_device_functions();  // defined at the end of this file, after every kernel is in view

int sync_refs(guac_ctx* ctx) {
  GUAC_TRY(sync_ref2(ctx));  // 2-bit copies and bad-group bitmaps of the windows that changed
  if (!ctx->refs_dirty) return GUAC_OK;
  std::vector<RefView> v;
  int idx = 0;
  for (auto& kv : ctx->contigs) {
    kv.second.ref_index = idx++;
    RefView r;
    r.bases = kv.second.bases;
    r.offset = kv.second.offset;
    r.len = kv.second.len;
    r.bad = kv.second.refbad;
    r.ref2 = kv.second.ref2;
    v.push_back(r);
  }
  GUAC_TRY(upload(ctx, ctx->d_refs, v.data(), v.size()));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));  // `v` is a temporary
  ctx->refs_dirty = false;
  return GUAC_OK;
}

struct Plan {
  int64_t n_slots = 0;
  int64_t n_tiles = 0;
};

// Validates the ranges, builds the device range table and the tile descriptors.
int make_plan(guac_ctx* ctx, const guac_range* loci, int32_t n_ranges, Plan& plan, int tile_loci = TILE_LOCI) {
  if (!ctx->have_reads) return fail(ctx, GUAC_E_STATE, "no reads uploaded (guac_reads_upload)");
  if (n_ranges < 0 || (n_ranges > 0 && !loci)) return fail(ctx, GUAC_E_INVALID_INPUT, "bad loci ranges");
  GUAC_TRY(sync_refs(ctx));
  std::vector<RangeDev>& rs = ctx->plan_ranges;  // kept in the context, so nothing below has to wait for the copy
  rs.clear();
  int64_t slot = 0, tile = 0;
  for (int i = 0; i < n_ranges; ++i) {
    const guac_range& g = loci[i];
    if (g.end < g.start || g.start < 0) return fail(ctx, GUAC_E_INVALID_INPUT, "loci range with end < start or negative start");
    if (i > 0 && (g.contig_id == loci[i - 1].contig_id) && g.start < loci[i - 1].end)
      return fail(ctx, GUAC_E_INVALID_INPUT, "loci ranges must be sorted and disjoint within a contig");
    if (g.end == g.start) continue;
    RangeDev r;
    r.start = g.start;
    r.end = g.end;
    r.slot0 = slot;
    r.tile0 = tile;
    r.contig_id = g.contig_id;
    r.read_clo = r.read_chi = 0;
    for (auto& sp : ctx->contig_spans)
      if (sp.first == g.contig_id) {
        r.read_clo = sp.second.first;
        r.read_chi = sp.second.second;
      }
    auto it = ctx->contigs.find(g.contig_id);
    if (it == ctx->contigs.end()) {
      if (r.read_chi > r.read_clo)
        return fail(ctx, GUAC_E_INVALID_INPUT, "Contig " + std::to_string(g.contig_id) + " does not exist in the current reference");
      continue;  // no reads and no reference: nothing can be visited
    }
    r.ref_index = it->second.ref_index;
    slot += g.end - g.start;
    tile += (g.end - g.start + tile_loci - 1) / tile_loci;
    rs.push_back(r);
  }
  if (slot >= (int64_t)1 << 31) return fail(ctx, GUAC_E_UNSUPPORTED, "more than 2^31-1 loci in one call: split the LociPartitioning");
  plan.n_slots = slot;
  plan.n_tiles = tile;
  if (tile == 0) return GUAC_OK;
  GUAC_TRY(upload(ctx, ctx->b_ranges, rs.data(), rs.size()));
  GUAC_TRY(ensure(ctx, ctx->b_tiles, (size_t)tile * sizeof(Tile)));
  build_tiles_kernel<<<blocks_for(tile, 256), 256, 0, ctx->stream>>>(ctx->b_ranges.as<RangeDev>(), (int)rs.size(), tile, tile_loci,
                                                                      ctx->R, ctx->b_tiles.as<Tile>(), ctx->d_refs.as<RefView>());
  ctx->launches += 1;
  CUDA_TRY(ctx, cudaGetLastError());
  return GUAC_OK;  // (no wait: the kernels that read the tiles follow on the same stream)
}

// Stream path: 2-bit copies of the reference windows, rebuilt for the windows that changed.
int sync_ref2(guac_ctx* ctx) {
  for (auto& kv : ctx->contigs) {
    ContigDev& c = kv.second;
    const size_t words = (size_t)(c.len / 16) + 4;  // read-ahead of the compare loop
    if (!c.ref2_dirty && c.ref2 && c.ref2_words == words) continue;
    if (!c.ref2 || c.ref2_words != words) {  // the allocation is kept while the window keeps its size: RefView tables stay valid
      if (c.ref2) CUDA_TRY(ctx, cudaFree(c.ref2));
      if (c.refbad) CUDA_TRY(ctx, cudaFree(c.refbad));
      c.ref2 = c.refbad = nullptr;
      if (cudaMalloc(&c.ref2, words * sizeof(uint32_t)) != cudaSuccess || cudaMalloc(&c.refbad, (words / 32 + 4) * sizeof(uint32_t)) != cudaSuccess) {
        (void)cudaGetLastError();
        return fail(ctx, GUAC_E_OOM, "cudaMalloc failed for the 2-bit reference");
      }
      c.ref2_words = words;
      ctx->refs_dirty = true;
    }
    CUDA_TRY(ctx, cudaMemsetAsync(c.ref2, 0, words * sizeof(uint32_t), ctx->stream));
    CUDA_TRY(ctx, cudaMemsetAsync(c.refbad, 0, (words / 32 + 4) * sizeof(uint32_t), ctx->stream));
    if (c.len > 0) {
      ref2_pack_kernel<<<blocks_for((c.len + 15) / 16, 256), 256, 0, ctx->stream>>>(c.bases, c.len, c.ref2, c.refbad);
      ctx->launches += 1;
      CUDA_TRY(ctx, cudaGetLastError());
    }
    c.ref2_dirty = false;
  }
  return GUAC_OK;
}

int reset_error(guac_ctx* ctx) {
  GUAC_TRY(ensure(ctx, ctx->b_err, sizeof(DeviceError)));
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->b_err.p, 0, sizeof(DeviceError), ctx->stream));
  return GUAC_OK;
}

// Small results (counter blocks, the error record) come back through one pinned block per context: a copy into
// pageable memory is staged by the driver and costs a second synchronisation.  Layout: [0, 256) counters of the
// call, [256, 512) the DeviceError.
int pinned_block(guac_ctx* ctx) {
  if (!ctx->h_counters) CUDA_TRY(ctx, cudaMallocHost(&ctx->h_counters, 512));
  return GUAC_OK;
}

// Translates a device-side error into the reference's wording.
int check_device_error(guac_ctx* ctx) {
  GUAC_TRY(pinned_block(ctx));
  DeviceError& e = *reinterpret_cast<DeviceError*>(static_cast<unsigned char*>(ctx->h_counters) + 256);
  CUDA_TRY(ctx, cudaMemcpyAsync(&e, ctx->b_err.p, sizeof e, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  if (e.code == DERR_NONE) return GUAC_OK;
  char buf[256];
  switch (e.code) {
    case DERR_REF_MISSING:
      snprintf(buf, sizeof buf, "Position %d:%lld missing from reference", e.contig_id, e.locus);
      return fail(ctx, GUAC_E_INVALID_INPUT, buf);
    case DERR_INSERTION_ELEMENT:
      snprintf(buf, sizeof buf, "Should not have a PileupElement at non-reference-consuming cigar-operator I. Locus: %lld (read %lld)",
               e.locus, e.read_index);
      return fail(ctx, GUAC_E_INVALID_INPUT, buf);
    case DERR_DELETION_PRECEDED:
      snprintf(buf, sizeof buf, "Found deletion preceded by a non-aligned cigar operator at PileupElement for read %lld at locus %lld",
               e.read_index, e.locus);
      return fail(ctx, GUAC_E_INVALID_INPUT, buf);
    case DERR_PADDING:
      return fail(ctx, GUAC_E_INVALID_INPUT, "`P` CIGAR-ops should have been ignored earlier in `findNextCigarElement`");
    case DERR_EMPTY_MAXBY:
      snprintf(buf, sizeof buf, "empty.maxBy: every non-match tumor element at %d:%lld is Clipped", e.contig_id, e.locus);
      return fail(ctx, GUAC_E_INVALID_INPUT, buf);
    case DERR_CURSOR:
      snprintf(buf, sizeof buf, "CIGAR of read %lld does not reach locus %lld", e.read_index, e.locus);
      return fail(ctx, GUAC_E_INVALID_INPUT, buf);
    case DERR_ALLELE_TABLE_FULL:
      snprintf(buf, sizeof buf, "more than 32 distinct variant alleles at %d:%lld (library limit)", e.contig_id, e.locus);
      return fail(ctx, GUAC_E_UNSUPPORTED, buf);
    case DERR_MISMATCH_HIST:
      snprintf(buf, sizeof buf, "read %lld has more than %d mismatches (library limit)", e.read_index, MM_BINS - 1);
      return fail(ctx, GUAC_E_UNSUPPORTED, buf);
    case DERR_ASM_BASE:
      snprintf(buf, sizeof buf, "read %lld in the assembly window at %d:%lld has a base other than A, C, G, T, N (library limit)", e.read_index,
               e.contig_id, e.locus);
      return fail(ctx, GUAC_E_UNSUPPORTED, buf);
    case DERR_ASM_COLLISION:
      snprintf(buf, sizeof buf, "two k-mers of the assembly window at %d:%lld share a 64-bit hash (library limit)", e.contig_id, e.locus);
      return fail(ctx, GUAC_E_UNSUPPORTED, buf);
    case DERR_ASM_CAPACITY:
      snprintf(buf, sizeof buf, "the assembly graph of the window at %d:%lld exceeds the per-window scratch (library limit)", e.contig_id,
               e.locus);
      return fail(ctx, GUAC_E_UNSUPPORTED, buf);
    case DERR_ASM_CIGAR:
      return fail(ctx, GUAC_E_UNSUPPORTED, "a haplotype alignment has more than 64 CIGAR elements (library limit)");
  }
  return fail(ctx, GUAC_E_CUDA, "unknown device error");
}

// visited flags -> compact indices; returns the number of visited loci.
int compact_visited(guac_ctx* ctx, const Plan& plan, int skip_empty, int64_t& n_visited) {
  GUAC_TRY(ensure(ctx, ctx->b_vis_idx, (size_t)(plan.n_slots + 1) * sizeof(int64_t)));
  if (!skip_empty) CUDA_TRY(ctx, cudaMemsetAsync(ctx->b_visited.p, 1, (size_t)plan.n_slots, ctx->stream));
  int64_t* total_dev = ctx->b_vis_idx.as<int64_t>() + plan.n_slots;
  GUAC_TRY((device_scan<true, uint8_t>(ctx, ctx->b_visited.as<uint8_t>(), ctx->b_vis_idx.as<int64_t>(), plan.n_slots, OpSum(), total_dev)));
  CUDA_TRY(ctx, cudaMemcpyAsync(&n_visited, total_dev, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return GUAC_OK;
}

// Derives end / running-max keys and publishes the device view (shared by upload and synth).
// The context's second stream with its fork / join events, made on first use.
int side_stream(guac_ctx* ctx) {
  if (ctx->stream2) return GUAC_OK;
  CUDA_TRY(ctx, cudaStreamCreateWithFlags(&ctx->stream2, cudaStreamNonBlocking));
  CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->fork, cudaEventDisableTiming));
  CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->join, cudaEventDisableTiming));
  CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->qual_ready, cudaEventDisableTiming));
  return GUAC_OK;
}

constexpr int64_t READ_INDEX_STRIDE = 1024;
__global__ void sample_columns_kernel(const int64_t* a, int64_t na, const int64_t* b, int64_t nb, const int64_t* c, int64_t nc,
                                      int64_t* out) {
  const int64_t ka = (na + READ_INDEX_STRIDE - 1) / READ_INDEX_STRIDE, kb = (nb + READ_INDEX_STRIDE - 1) / READ_INDEX_STRIDE,
                kc = (nc + READ_INDEX_STRIDE - 1) / READ_INDEX_STRIDE;
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < ka) out[i] = a[i * READ_INDEX_STRIDE];
  else if (i < ka + kb) out[i] = b[(i - ka) * READ_INDEX_STRIDE];
  else if (i < ka + kb + kc) out[i] = c[(i - ka - kb) * READ_INDEX_STRIDE];
}
// The host's sampled index of the batch (see guac_ctx::h_idx_start): one small kernel, one copy.
int sample_read_index(guac_ctx* ctx, int64_t n) {
  const int64_t kr = (n + READ_INDEX_STRIDE - 1) / READ_INDEX_STRIDE, km = (ctx->n_medium + READ_INDEX_STRIDE - 1) / READ_INDEX_STRIDE;
  ctx->h_idx_start.assign((size_t)kr, 0);
  ctx->h_idx_runmax.assign((size_t)kr, 0);
  ctx->h_idx_medium.assign((size_t)km, 0);
  const int64_t total = 2 * kr + km;
  if (total == 0) return GUAC_OK;
  GUAC_TRY(ensure(ctx, ctx->u_idx, (size_t)total * sizeof(int64_t)));
  sample_columns_kernel<<<blocks_for(total, 256), 256, 0, ctx->stream>>>(ctx->b_start.as<int64_t>(), n, ctx->b_runmax.as<int64_t>(), n,
                                                                        ctx->b_medium.as<int64_t>(), ctx->n_medium, ctx->u_idx.as<int64_t>());
  ctx->launches += 1;
  CUDA_TRY(ctx, cudaGetLastError());
  std::vector<int64_t> h((size_t)total);
  CUDA_TRY(ctx, cudaMemcpyAsync(h.data(), ctx->u_idx.p, (size_t)total * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  std::copy(h.begin(), h.begin() + kr, ctx->h_idx_start.begin());
  std::copy(h.begin() + kr, h.begin() + 2 * kr, ctx->h_idx_runmax.begin());
  std::copy(h.begin() + 2 * kr, h.end(), ctx->h_idx_medium.begin());
  return GUAC_OK;
}

int finish_reads(guac_ctx* ctx, int64_t n, int32_t n_samples, int64_t n_exc) {
  GUAC_TRY(ensure(ctx, ctx->b_end, (size_t)(n + 1) * sizeof(int64_t)));
  GUAC_TRY(ensure(ctx, ctx->b_key, (size_t)(n + 1) * sizeof(int64_t)));
  GUAC_TRY(ensure(ctx, ctx->b_runmax, (size_t)(n + 1) * sizeof(int64_t)));
  if (n > 0) {
    read_ends_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(n, ctx->b_start.as<int64_t>(), ctx->b_contig.as<int32_t>(),
                                                                  ctx->b_cigar_off.as<uint32_t>(), ctx->b_cigar.as<uint32_t>(),
                                                                  ctx->b_end.as<int64_t>(), ctx->b_key.as<int64_t>());
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    GUAC_TRY((device_scan<false, int64_t>(ctx, ctx->b_key.as<int64_t>(), ctx->b_runmax.as<int64_t>(), n, OpMax())));
  }
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));  // caller's buffers may go away after return
  ReadsView& R = ctx->R;
  R.n_reads = n;
  R.start = ctx->b_start.as<int64_t>();
  R.end = ctx->b_end.as<int64_t>();
  R.runmax_key = ctx->b_runmax.as<int64_t>();
  R.cigar_off = ctx->b_cigar_off.as<uint32_t>();
  R.cigar = ctx->b_cigar.as<uint32_t>();
  R.seq_off = ctx->b_seq_off.as<uint64_t>();
  R.bases2 = ctx->b_bases2.as<uint8_t>();
  R.nmask = ctx->b_nmask.as<uint8_t>();
  R.n_exc = n_exc;
  R.exc_pos = ctx->b_exc_pos.as<uint64_t>();
  R.exc_base = ctx->b_exc_base.as<uint8_t>();
  R.qual = ctx->b_qual.as<uint8_t>();
  R.mapq = ctx->b_mapq.as<uint8_t>();
  R.flags = ctx->b_flags.as<uint8_t>();
  R.sample_id = ctx->b_sample.as<uint8_t>();
  GUAC_TRY(ensure(ctx, ctx->b_rflags, (size_t)n + 16));
  GUAC_TRY(ensure(ctx, ctx->b_hdr, ((size_t)n + 1) * sizeof(uint4)));
  R.hdr = nullptr;  // (the kernels below build it)
  if (n > 0) {
    read_fast_flags_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(R, ctx->b_rflags.as<uint8_t>());
    read_headers_pack_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(R, ctx->b_rflags.as<uint8_t>(), ctx->b_hdr.as<uint4>());
    {  // the multi-operator reads, listed in read order: flags -> exclusive scan -> scatter
      DevBuf& d_flag = ctx->u_flag;
      DevBuf& d_off = ctx->u_off;
      GUAC_TRY(ensure(ctx, d_flag, (size_t)n + 16));
      GUAC_TRY(ensure(ctx, d_off, ((size_t)n + 2) * sizeof(int64_t)));
      medium_reads_flag_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(ctx->b_rflags.as<uint8_t>(), n, d_flag.as<uint8_t>());
      GUAC_TRY((device_scan<true, uint8_t>(ctx, d_flag.as<uint8_t>(), d_off.as<int64_t>(), n, OpSum(), d_off.as<int64_t>() + n)));
      CUDA_TRY(ctx, cudaMemcpyAsync(&ctx->n_medium, d_off.as<int64_t>() + n, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
      CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
      GUAC_TRY(ensure(ctx, ctx->b_medium, ((size_t)ctx->n_medium + 2) * sizeof(int64_t)));
      medium_reads_scatter_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(d_flag.as<uint8_t>(), d_off.as<int64_t>(), n,
                                                                            ctx->b_medium.as<int64_t>());
    }
    ctx->launches += 7;
    CUDA_TRY(ctx, cudaGetLastError());
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  } else {
    ctx->n_medium = 0;
  }
  GUAC_TRY(sample_read_index(ctx, n));
  R.hdr = n > 0 ? ctx->b_hdr.as<uint4>() : nullptr;
  ctx->n_reads = n;
  ctx->n_samples = n_samples;
  ctx->have_reads = true;
  return GUAC_OK;
}

// ln n! for n < SJ_LOGFACT_N, summed in extended precision and rounded once (Fisher's exact test of the joint caller's
// StrandBias annotation, filters/FishersExactTest.scala:12-19)
constexpr int SJ_LOGFACT_N = 1 << 16;
int sj_logfact_table(guac_ctx* ctx) {
  if (ctx->j_logfact.p) return GUAC_OK;
  std::vector<double> t((size_t)SJ_LOGFACT_N);
  long double acc = 0.0L;
  t[0] = 0.0;
  for (int n = 1; n < SJ_LOGFACT_N; ++n) {
    acc += std::log((long double)n);
    t[(size_t)n] = (double)acc;
  }
  GUAC_TRY(upload(ctx, ctx->j_logfact, t.data(), t.size()));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));  // `t` is a temporary
  GUAC_TRY(ensure(ctx, ctx->j_logtab, 65536 * sizeof(double2)));
  sj_log_table_kernel<<<256, 256, 0, ctx->stream>>>(ctx->j_logtab.as<double2>());
  ctx->launches += 1;
  CUDA_TRY(ctx, cudaGetLastError());
  return GUAC_OK;
}

// Read pass of the stream form (stream_kernels.cuh): every read compared once with the 2-bit reference of its contig, its
// non-Match elements left in its slot of b_bins / b_bin_count.  One launch per contig of the batch.
int run_read_pass(guac_ctx* ctx) {
  GUAC_TRY(sync_ref2(ctx));
  GUAC_TRY(ensure(ctx, ctx->b_bins, (size_t)(ctx->n_reads + 1) * STREAM_EV_CAP * sizeof(uint2)));
  GUAC_TRY(ensure(ctx, ctx->b_bin_count, (size_t)ctx->n_reads + 16));
  // overflow arena: room for one read in 32 (a read beyond it, or with more than STREAM_OVF_CAP events, is walked in the
  // locus pass as before)
  const int max_overflow = (int)std::min<int64_t>(ctx->n_reads / 32 + 1024, 1 << 24);
  GUAC_TRY(ensure(ctx, ctx->b_ev_overflow, 64 + (size_t)max_overflow * (STREAM_OVF_CAP * sizeof(uint2) + sizeof(int64_t))));
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->b_ev_overflow.p, 0, 64, ctx->stream));
  {  // reads of a contig without a reference get no launch of the read pass: their counts must read 0
    bool every_contig_has_a_reference = true;
    for (auto& sp : ctx->contig_spans) every_contig_has_a_reference = every_contig_has_a_reference && ctx->contigs.count(sp.first) != 0;
    if (!every_contig_has_a_reference) CUDA_TRY(ctx, cudaMemsetAsync(ctx->b_bin_count.p, 0, (size_t)ctx->n_reads + 16, ctx->stream));
  }
  for (auto& sp : ctx->contig_spans) {  // one launch per contig of the batch: its reference rides in the arguments
    auto it = ctx->contigs.find(sp.first);
    if (it == ctx->contigs.end() || sp.second.second <= sp.second.first) continue;
    // the reads that can overlap the loci of this call (make_plan's ranges), to the grain of the host's sampled index: from
    // the last sampled read whose running-max end is still at or before the first locus, to the first sampled read that
    // starts at or after the last one
    int64_t first_locus = INT64_MAX, end_locus = INT64_MIN;
    for (const RangeDev& r : ctx->plan_ranges)
      if (r.contig_id == sp.first) {
        first_locus = std::min<int64_t>(first_locus, r.start);
        end_locus = std::max<int64_t>(end_locus, r.end);
      }
    if (end_locus <= first_locus) continue;
    int64_t read_lo = sp.second.first, read_hi = sp.second.second;
    {
      const int64_t k0 = (read_lo + READ_INDEX_STRIDE - 1) / READ_INDEX_STRIDE, k1 = (read_hi + READ_INDEX_STRIDE - 1) / READ_INDEX_STRIDE;
      const int64_t key = ((int64_t)sp.first << 40) | first_locus;  // (read_ends_kernel's key)
      const int64_t* rm = ctx->h_idx_runmax.data();
      const int64_t* st = ctx->h_idx_start.data();
      const int64_t ka = std::upper_bound(rm + k0, rm + k1, key) - rm;           // first sample with running max > key
      if (ka > k0) read_lo = (ka - 1) * READ_INDEX_STRIDE;
      const int64_t kb = std::lower_bound(st + k0, st + k1, end_locus) - st;      // first sample starting at or after the end
      if (kb < k1) read_hi = kb * READ_INDEX_STRIDE;
    }
    if (read_hi <= read_lo) continue;
    StreamArgs SA;
    SA.hdr = ctx->b_hdr.as<uint4>();
    SA.read_lo = read_lo;
    SA.read_hi = read_hi;
    SA.ref.bases = it->second.bases;
    SA.ref.offset = it->second.offset;
    SA.ref.len = it->second.len;
    SA.ref.bad = it->second.refbad;
    SA.ref.ref2 = it->second.ref2;
    SA.ref2 = it->second.ref2;
    SA.events = ctx->b_bins.as<uint2>();
    SA.n_events = ctx->b_bin_count.as<uint8_t>();
    SA.overflow = reinterpret_cast<uint2*>(static_cast<unsigned char*>(ctx->b_ev_overflow.p) + 64);
    SA.n_overflow = ctx->b_ev_overflow.as<int>();
    SA.max_overflow = max_overflow;
    SA.overflow_reads = reinterpret_cast<int64_t*>(static_cast<unsigned char*>(ctx->b_ev_overflow.p) + 64 +
                                                   (size_t)max_overflow * STREAM_OVF_CAP * sizeof(uint2));
    SA.list = nullptr;
    SA.n_list = 0;
    SA.list_blocks = SA.list_period = 0;
    SA.inline_medium = ctx->contig_spans.size() == 1 ? 0 : 1;
    unsigned blocks = blocks_for(SA.read_hi - SA.read_lo, 256);
    if (ctx->contig_spans.size() == 1 && ctx->n_medium > 0) {  // the multi-operator reads ride in the same launch
      // (the part of the list inside [read_lo, read_hi), again to the grain of the sampled index)
      const int64_t* md = ctx->h_idx_medium.data();
      const int64_t km = (int64_t)ctx->h_idx_medium.size();
      const int64_t ma = std::lower_bound(md, md + km, read_lo) - md, mb = std::lower_bound(md, md + km, read_hi) - md;
      const int64_t list_lo = std::max<int64_t>(0, ma - 1) * READ_INDEX_STRIDE;
      const int64_t list_hi = std::min<int64_t>(ctx->n_medium, mb * READ_INDEX_STRIDE);
      if (list_hi > list_lo) {
        SA.list = ctx->b_medium.as<int64_t>() + list_lo;
        SA.n_list = list_hi - list_lo;
        SA.list_blocks = blocks_for(SA.n_list, 256);
        SA.list_period = std::max(1u, blocks / SA.list_blocks) + 1;
        blocks += SA.list_blocks;
      } else {
        SA.inline_medium = 1;  // (no multi-operator read in range)
      }
    }
    ss_read_events_kernel<<<blocks, 256, 0, ctx->stream>>>(ctx->R, SA);
    ss_read_overflow_kernel<<<(unsigned)std::min<int64_t>(blocks_for(max_overflow, 128), 148 * 8), 128, 0, ctx->stream>>>(ctx->R, SA);
    ctx->launches += 2;
  }
  CUDA_TRY(ctx, cudaGetLastError());
  return GUAC_OK;
}

}  // namespace

extern "C" {

int guac_ctx_create(int device, guac_ctx** out) {
  if (!out) return GUAC_E_INVALID_INPUT;
  *out = nullptr;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) {
    (void)cudaGetLastError();
    return GUAC_E_NO_DEVICE;
  }
  if (cudaSetDevice(device) != cudaSuccess) return GUAC_E_CUDA;
  guac_ctx* ctx = new guac_ctx();
  ctx->device = device;
  if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
    delete ctx;
    return GUAC_E_CUDA;
  }
  // ADAM PhredUtils table: 1 - 10^(-q/10)
  double table[256];
  for (int q = 0; q < 256; ++q) table[q] = 1.0 - std::pow(10.0, -q / 10.0);
  float odds_m1[256];
  for (int q = 0; q < 256; ++q) {
    const double o = table[q] / (1.0 - table[q]) - 1.0;  // +inf once 1 - p underflows: such a locus goes to the exact path
    odds_m1[q] = std::nextafter((float)o, INFINITY);      // never below the exact value: the sum stays an upper bound
  }
  if (cudaMemcpyToSymbol(c_success_probability, table, sizeof table) != cudaSuccess ||
      cudaMemcpyToSymbol(c_odds_minus_one, odds_m1, sizeof odds_m1) != cudaSuccess) {
    cudaStreamDestroy(ctx->stream);
    delete ctx;
    return GUAC_E_CUDA;
  }
  cudaEventCreate(&ctx->ev0);
  cudaEventCreate(&ctx->ev1);
  // Function attributes are per device: every context sets them for its own device (a process may hold contexts on
  // several GPUs), for every instantiation that is ever launched with dynamic shared memory.
  if (configure_device_functions() != cudaSuccess) {
    (void)cudaGetLastError();
    guac_ctx_destroy(ctx);
    return GUAC_E_CUDA;
  }
  *out = ctx;
  return GUAC_OK;
}

void guac_ctx_destroy(guac_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  DevBuf* bufs[] = {&ctx->d_refs, &ctx->b_start, &ctx->b_end, &ctx->b_key, &ctx->b_runmax, &ctx->b_contig, &ctx->b_cigar_off,
                    &ctx->b_cigar, &ctx->b_seq_off, &ctx->b_bases2, &ctx->b_nmask, &ctx->b_exc_pos, &ctx->b_exc_base, &ctx->b_qual,
                    &ctx->b_mapq, &ctx->b_flags, &ctx->b_sample, &ctx->b_scan_tmp, &ctx->b_ranges, &ctx->b_tiles, &ctx->b_visited,
                    &ctx->b_vis_idx, &ctx->b_counts, &ctx->b_totals, &ctx->b_elem_off, &ctx->b_summary, &ctx->b_calls,
                    &ctx->b_ncalls, &ctx->b_err, &ctx->b_arena};
  for (DevBuf* b : bufs)
    if (b->p) cudaFree(b->p);
  for (auto& kv : ctx->contigs)
    if (kv.second.bases) cudaFree(kv.second.bases);
  for (auto& kv : ctx->host_allocs) cudaFreeHost(kv.first);
  for (auto& kv : ctx->pinned_free) cudaFreeHost(kv.second);
  if (ctx->b_nops.p) cudaFree(ctx->b_nops.p);
  if (ctx->b_rflags.p) cudaFree(ctx->b_rflags.p);
  if (ctx->b_logtab.p) cudaFree(ctx->b_logtab.p);
  if (ctx->b_fallback.p) cudaFree(ctx->b_fallback.p);
  if (ctx->b_heavy.p) cudaFree(ctx->b_heavy.p);
  if (ctx->b_vcount.p) cudaFree(ctx->b_vcount.p);
  if (ctx->b_sj_sites.p) cudaFree(ctx->b_sj_sites.p);
  if (ctx->b_boundtab.p) cudaFree(ctx->b_boundtab.p);
  if (ctx->b_arena_used.p) cudaFree(ctx->b_arena_used.p);
  if (ctx->b_hdr.p) cudaFree(ctx->b_hdr.p);
  for (auto& b : ctx->asm_blocks) cudaFree(b.first);
  if (ctx->stream2) cudaStreamDestroy(ctx->stream2);
  if (ctx->fork) cudaEventDestroy(ctx->fork);
  if (ctx->join) cudaEventDestroy(ctx->join);
  if (ctx->qual_ready) cudaEventDestroy(ctx->qual_ready);
  for (DevBuf* b : {&ctx->b_ev_overflow, &ctx->b_bins, &ctx->b_bin_count, &ctx->b_medium, &ctx->u_len, &ctx->u_ops, &ctx->u_runs_c, &ctx->u_runs_f, &ctx->u_tmp,
                    &ctx->u_qpacked, &ctx->u_bad, &ctx->u_flag, &ctx->u_off, &ctx->u_idx, &ctx->u_expl, &ctx->u_mm_count, &ctx->u_mm_words, &ctx->u_expl_len, &ctx->u_mm_n,
                    &ctx->u_expl_off, &ctx->u_mm_off, &ctx->u_start_rel, &ctx->u_runs_base, &ctx->u_ops8,
                    &ctx->u_ops_eff, &ctx->u_wire_off, &ctx->u_wire, &ctx->g_active, &ctx->g_jobs, &ctx->g_evs, &ctx->g_strings,
                    &ctx->r_bitmap, &ctx->r_counts, &ctx->r_prefix, &ctx->r_rank, &ctx->j_force, &ctx->j_simple, &ctx->j_sites_fast,
                    &ctx->j_raw_fast, &ctx->j_bail, &ctx->j_raw_gen, &ctx->j_emits, &ctx->j_order, &ctx->j_cnt, &ctx->j_off,
                    &ctx->j_out_alleles, &ctx->j_out_samples, &ctx->j_logfact, &ctx->j_logtab})
    if (b->p) cudaFree(b->p);
  if (ctx->h_counters) cudaFreeHost(ctx->h_counters);
  for (auto& kv : ctx->contigs)
    if (kv.second.ref2) cudaFree(kv.second.ref2);
  for (auto& kv : ctx->contigs)
    if (kv.second.refbad) cudaFree(kv.second.refbad);
  if (ctx->ev0) cudaEventDestroy(ctx->ev0);
  if (ctx->ev1) cudaEventDestroy(ctx->ev1);
  cudaStreamDestroy(ctx->stream);
  delete ctx;
}

const char* guac_last_error(guac_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

void guac_free(guac_ctx* ctx, void* p) {
  if (!ctx || !p) return;
  auto it = ctx->host_allocs.find(p);
  if (it == ctx->host_allocs.end()) return;
  const size_t cap = it->second;
  ctx->host_allocs.erase(it);
  if (ctx->pinned_free_bytes + cap <= ((size_t)2 << 30)) {  // keep up to 2 GiB of pinned blocks for reuse
    ctx->pinned_free.insert({cap, p});
    ctx->pinned_free_bytes += cap;
  } else {
    cudaFreeHost(p);
  }
}

int64_t guac_kernel_launches(guac_ctx* ctx) { return ctx ? ctx->launches : 0; }

int guac_pack_bases(const uint8_t* ascii, int64_t n, uint64_t base_index0, uint8_t* bases2, uint8_t* nmask, uint64_t* exc_pos,
                    uint8_t* exc_base, int64_t exc_cap, int64_t* n_exc) {
  if (!ascii || !bases2 || !nmask || !n_exc || (base_index0 & 7)) return GUAC_E_INVALID_INPUT;
  int64_t ne = 0;
  std::memset(bases2, 0, (size_t)((n + 3) / 4));
  std::memset(nmask, 0, (size_t)((n + 7) / 8));
  for (int64_t i = 0; i < n; ++i) {
    uint8_t b = ascii[i];
    int code;
    switch (b) {
      case 'A': code = 0; break;
      case 'C': code = 1; break;
      case 'G': code = 2; break;
      case 'T': code = 3; break;
      default: code = -1;
    }
    if (code >= 0) {
      bases2[i >> 2] |= (uint8_t)(code << ((i & 3) * 2));
    } else {
      nmask[i >> 3] |= (uint8_t)(1u << (i & 7));
      if (ne < exc_cap) {
        exc_pos[ne] = base_index0 + (uint64_t)i;
        exc_base[ne] = b;
      }
      ++ne;
    }
  }
  *n_exc = ne;
  return ne > exc_cap ? GUAC_E_OOM : GUAC_OK;
}

int guac_ref_reserve(guac_ctx* ctx, int32_t contig_id, int64_t offset, int64_t len) {
  if (!ctx) return GUAC_E_INVALID_INPUT;
  if (len < 0 || offset < 0 || contig_id < 0 || contig_id >= (1 << 22)) return fail(ctx, GUAC_E_INVALID_INPUT, "bad reference window");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  ContigDev& c = ctx->contigs[contig_id];
  if (c.bases) CUDA_TRY(ctx, cudaFree(c.bases));
  c.bases = nullptr;
  c.offset = offset;
  c.len = len;
  c.ref2_dirty = true;
  cudaError_t e = cudaMalloc(&c.bases, (size_t)std::max<int64_t>(len, 16));
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    ctx->contigs.erase(contig_id);
    return fail(ctx, GUAC_E_OOM, "cudaMalloc failed for reference window");
  }
  CUDA_TRY(ctx, cudaMemsetAsync(c.bases, 0, (size_t)std::max<int64_t>(len, 16), ctx->stream));
  ctx->refs_dirty = true;
  return GUAC_OK;
}

int guac_ref_upload(guac_ctx* ctx, int32_t contig_id, const uint8_t* bases, int64_t len, int64_t offset) {
  if (!ctx) return GUAC_E_INVALID_INPUT;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  auto it = ctx->contigs.find(contig_id);
  if (it == ctx->contigs.end()) {
    GUAC_TRY(guac_ref_reserve(ctx, contig_id, offset, len));
    it = ctx->contigs.find(contig_id);
  }
  ContigDev& c = it->second;
  if (offset < c.offset || offset + len > c.offset + c.len)
    return fail(ctx, GUAC_E_INVALID_INPUT, "Illegal reference slice: outside the reserved window");
  if (len > 0) {
    CUDA_TRY(ctx, cudaMemcpyAsync(c.bases + (offset - c.offset), bases, (size_t)len, cudaMemcpyHostToDevice, ctx->stream));
    c.ref2_dirty = true;
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  }
  return GUAC_OK;
}

int guac_reads_upload(guac_ctx* ctx, const guac_read_batch* b) {
  if (!ctx || !b) return GUAC_E_INVALID_INPUT;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  ctx->have_reads = false;
  const int64_t n = b->n_reads;
  if (n < 0 || b->n_samples <= 0 || b->n_samples > 255) return fail(ctx, GUAC_E_INVALID_INPUT, "bad read batch header");
  // Host-side validation of the ordering contract (SlidingWindow.scala:36-37 "Regions must be sorted by start locus";
  // RegionsByContig.scala:22 "Regions are not sorted by contig") and the per-contig spans.
  // It runs on the host while the asynchronous H2D copies below are in flight.
  auto validate = [&]() -> int {
  ctx->contig_spans.clear();
  for (int64_t i = 0; i < n; ++i) {
    if (i == 0 || b->contig_id[i] != b->contig_id[i - 1]) {
      for (auto& sp : ctx->contig_spans)
        if (sp.first == b->contig_id[i]) return fail(ctx, GUAC_E_INVALID_INPUT, "Regions are not sorted by contig");
      if (!ctx->contig_spans.empty()) ctx->contig_spans.back().second.second = i;
      ctx->contig_spans.push_back({b->contig_id[i], {i, n}});
      if (b->contig_id[i] < 0 || b->contig_id[i] >= (1 << 22)) return fail(ctx, GUAC_E_INVALID_INPUT, "contig_id out of range");
    } else if (b->start[i] < b->start[i - 1]) {
      return fail(ctx, GUAC_E_INVALID_INPUT, "Regions must be sorted by start locus");
    }
    if (b->sample_id[i] >= b->n_samples) return fail(ctx, GUAC_E_INVALID_INPUT, "sample_id >= n_samples");
    if (b->start[i] < 0 || b->start[i] >= ((int64_t)1 << 39)) return fail(ctx, GUAC_E_INVALID_INPUT, "read start out of range");
  }
  return GUAC_OK;
  };
  const uint64_t n_bases = n ? b->seq_off[n] : 0;
  const uint32_t n_cigar = n ? b->cigar_off[n] : 0;
  GUAC_TRY(upload(ctx, ctx->b_start, b->start, (size_t)n));
  GUAC_TRY(upload(ctx, ctx->b_contig, b->contig_id, (size_t)n));
  GUAC_TRY(upload(ctx, ctx->b_cigar_off, b->cigar_off, (size_t)n + 1));
  GUAC_TRY(upload(ctx, ctx->b_cigar, b->cigar, (size_t)n_cigar));
  GUAC_TRY(upload(ctx, ctx->b_seq_off, b->seq_off, (size_t)n + 1));
  GUAC_TRY(upload(ctx, ctx->b_bases2, b->bases2, (size_t)((n_bases + 3) / 4), 64));  // padding: word read-ahead / 16-byte bulk copies
  GUAC_TRY(upload(ctx, ctx->b_nmask, b->nmask, (size_t)((n_bases + 7) / 8)));
  GUAC_TRY(upload(ctx, ctx->b_exc_pos, b->exc_pos, (size_t)b->n_exc));
  GUAC_TRY(upload(ctx, ctx->b_exc_base, b->exc_base, (size_t)b->n_exc));
  GUAC_TRY(upload(ctx, ctx->b_qual, b->qual, (size_t)n_bases));
  GUAC_TRY(upload(ctx, ctx->b_mapq, b->mapq, (size_t)n));
  GUAC_TRY(upload(ctx, ctx->b_flags, b->flags, (size_t)n));
  GUAC_TRY(upload(ctx, ctx->b_sample, b->sample_id, (size_t)n));
  int vst = validate();
  if (vst == GUAC_OK && n > 0) {
    // the offsets first (the length check walks the CIGAR words they point at), each with its own answer
    vst = ensure(ctx, ctx->u_bad, sizeof(int));
    int bad = 0;
    for (int pass = 0; vst == GUAC_OK && pass < 2; ++pass) {
      cudaMemsetAsync(ctx->u_bad.p, 0, sizeof(int), ctx->stream);
      if (pass == 0)
        validate_offsets_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(n, ctx->b_cigar_off.as<uint32_t>(), ctx->b_seq_off.as<uint64_t>(),
                                                                           n_cigar, n_bases, ctx->u_bad.as<int>());
      else
        validate_lengths_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(n, ctx->b_cigar_off.as<uint32_t>(), ctx->b_cigar.as<uint32_t>(),
                                                                           ctx->b_seq_off.as<uint64_t>(), ctx->u_bad.as<int>());
      ctx->launches += 1;
      if (cudaMemcpyAsync(&bad, ctx->u_bad.p, sizeof bad, cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess ||
          cudaStreamSynchronize(ctx->stream) != cudaSuccess)
        vst = fail(ctx, GUAC_E_CUDA, "validation of the read batch failed to run");
      else if (bad & 1)
        vst = fail(ctx, GUAC_E_INVALID_INPUT, "cigar_off / seq_off must be non-decreasing and end at the uploaded sizes");
      else if (bad & 2)
        vst = fail(ctx, GUAC_E_INVALID_INPUT, "a read's CIGAR does not consume as many bases as the read has");
    }
  }
  if (vst != GUAC_OK) {
    cudaStreamSynchronize(ctx->stream);
    return vst;
  }
  return finish_reads(ctx, n, b->n_samples, b->n_exc);
}

int guac_pileup_counts(guac_ctx* ctx, const guac_range* loci, int32_t n_ranges, int32_t skip_empty, int32_t** out_contig,
                       int64_t** out_locus, guac_locus_counts** out_counts, int64_t* n_out) {
  if (!ctx || !out_contig || !out_locus || !out_counts || !n_out) return GUAC_E_INVALID_INPUT;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  Plan plan;
  GUAC_TRY(make_plan(ctx, loci, n_ranges, plan));
  const int S = ctx->n_samples;
  *n_out = 0;
  if (plan.n_tiles == 0) {
    GUAC_TRY(host_out(ctx, out_contig, nullptr, 0));
    GUAC_TRY(host_out(ctx, out_locus, nullptr, 0));
    GUAC_TRY(host_out(ctx, out_counts, nullptr, 0));
    return GUAC_OK;
  }
  GUAC_TRY(reset_error(ctx));
  GUAC_TRY(ensure(ctx, ctx->b_visited, (size_t)plan.n_slots));
  GUAC_TRY(ensure(ctx, ctx->b_counts, (size_t)plan.n_slots * S * sizeof(guac_locus_counts)));
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->b_counts.p, 0, (size_t)plan.n_slots * S * sizeof(guac_locus_counts), ctx->stream));
  pileup_counts_kernel<<<(unsigned)plan.n_tiles, CTA_THREADS, 0, ctx->stream>>>(
      ctx->R, ctx->d_refs.as<RefView>(), ctx->b_tiles.as<Tile>(), S, ctx->b_visited.as<uint8_t>(),
      ctx->b_counts.as<guac_locus_counts>(), ctx->b_err.as<DeviceError>());
  ctx->launches += 1;
  CUDA_TRY(ctx, cudaGetLastError());
  GUAC_TRY(check_device_error(ctx));
  int64_t nv = 0;
  GUAC_TRY(compact_visited(ctx, plan, skip_empty, nv));
  TmpDev oc, ol, ocnt;
  GUAC_TRY(ensure(ctx, oc, (size_t)nv * sizeof(int32_t)));
  GUAC_TRY(ensure(ctx, ol, (size_t)nv * sizeof(int64_t)));
  GUAC_TRY(ensure(ctx, ocnt, (size_t)nv * S * sizeof(guac_locus_counts)));
  compact_loci_kernel<<<(unsigned)plan.n_tiles, 64, 0, ctx->stream>>>(ctx->b_tiles.as<Tile>(), ctx->b_visited.as<uint8_t>(),
                                                                      ctx->b_vis_idx.as<int64_t>(), oc.as<int32_t>(), ol.as<int64_t>());
  gather_rows_kernel<guac_locus_counts><<<blocks_for(plan.n_slots * S, 256), 256, 0, ctx->stream>>>(
      ctx->b_counts.as<guac_locus_counts>(), ctx->b_visited.as<uint8_t>(), ctx->b_vis_idx.as<int64_t>(), plan.n_slots, S,
      ocnt.as<guac_locus_counts>());
  ctx->launches += 2;
  CUDA_TRY(ctx, cudaGetLastError());
  GUAC_TRY(host_out(ctx, out_contig, oc.p, (size_t)nv));
  GUAC_TRY(host_out(ctx, out_locus, ol.p, (size_t)nv));
  GUAC_TRY(host_out(ctx, out_counts, ocnt.p, (size_t)nv * S));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  *n_out = nv;
  return GUAC_OK;
}

int guac_pileup_elements(guac_ctx* ctx, const guac_range* loci, int32_t n_ranges, int32_t skip_empty, int32_t** out_contig,
                         int64_t** out_locus, int64_t** out_offsets, guac_element** out_elements, int64_t* n_out) {
  if (!ctx || !out_contig || !out_locus || !out_offsets || !out_elements || !n_out) return GUAC_E_INVALID_INPUT;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  Plan plan;
  GUAC_TRY(make_plan(ctx, loci, n_ranges, plan));
  const int S = ctx->n_samples;
  *n_out = 0;
  if (plan.n_tiles == 0) {
    GUAC_TRY(host_out(ctx, out_contig, nullptr, 0));
    GUAC_TRY(host_out(ctx, out_locus, nullptr, 0));
    GUAC_TRY(host_out(ctx, out_elements, nullptr, 0));
    int64_t zero = 0;
    void* h = nullptr;
    GUAC_TRY(pinned_alloc(ctx, &h, 16));
    std::memcpy(h, &zero, sizeof zero);
    *out_offsets = (int64_t*)h;
    return GUAC_OK;
  }
  GUAC_TRY(reset_error(ctx));
  GUAC_TRY(ensure(ctx, ctx->b_visited, (size_t)plan.n_slots));
  GUAC_TRY(ensure(ctx, ctx->b_totals, (size_t)plan.n_slots * sizeof(int32_t)));
  GUAC_TRY(ensure(ctx, ctx->b_elem_off, (size_t)(plan.n_slots + 1) * sizeof(int64_t)));
  depth_total_kernel<<<(unsigned)plan.n_tiles, CTA_THREADS, 0, ctx->stream>>>(ctx->R, ctx->b_tiles.as<Tile>(), S,
                                                                             ctx->b_visited.as<uint8_t>(), ctx->b_totals.as<int32_t>());
  ctx->launches += 1;
  CUDA_TRY(ctx, cudaGetLastError());
  int64_t* total_dev = ctx->b_elem_off.as<int64_t>() + plan.n_slots;
  GUAC_TRY((device_scan<true, int32_t>(ctx, ctx->b_totals.as<int32_t>(), ctx->b_elem_off.as<int64_t>(), plan.n_slots, OpSum(), total_dev)));
  int64_t n_elems = 0;
  CUDA_TRY(ctx, cudaMemcpyAsync(&n_elems, total_dev, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  int64_t nv = 0;
  GUAC_TRY(compact_visited(ctx, plan, skip_empty, nv));
  TmpDev oc, ol, oo, oe;
  GUAC_TRY(ensure(ctx, oc, (size_t)nv * sizeof(int32_t)));
  GUAC_TRY(ensure(ctx, ol, (size_t)nv * sizeof(int64_t)));
  GUAC_TRY(ensure(ctx, oo, (size_t)(nv + 1) * sizeof(int64_t)));
  GUAC_TRY(ensure(ctx, oe, (size_t)n_elems * sizeof(guac_element)));
  write_elements_kernel<<<(unsigned)plan.n_tiles, CTA_THREADS, 0, ctx->stream>>>(
      ctx->R, ctx->d_refs.as<RefView>(), ctx->b_tiles.as<Tile>(), S, skip_empty, ctx->b_elem_off.as<int64_t>(),
      oe.as<guac_element>(), ctx->b_err.as<DeviceError>());
  compact_loci_kernel<<<(unsigned)plan.n_tiles, 64, 0, ctx->stream>>>(ctx->b_tiles.as<Tile>(), ctx->b_visited.as<uint8_t>(),
                                                                      ctx->b_vis_idx.as<int64_t>(), oc.as<int32_t>(), ol.as<int64_t>());
  gather_offsets_kernel<<<blocks_for(plan.n_slots, 256), 256, 0, ctx->stream>>>(
      ctx->b_elem_off.as<int64_t>(), ctx->b_visited.as<uint8_t>(), ctx->b_vis_idx.as<int64_t>(), plan.n_slots, oo.as<int64_t>());
  ctx->launches += 3;
  CUDA_TRY(ctx, cudaGetLastError());
  CUDA_TRY(ctx, cudaMemcpyAsync(oo.as<int64_t>() + nv, &n_elems, sizeof(int64_t), cudaMemcpyHostToDevice, ctx->stream));
  int st = check_device_error(ctx);
  if (st == GUAC_OK) st = host_out(ctx, out_contig, oc.p, (size_t)nv);
  if (st == GUAC_OK) st = host_out(ctx, out_locus, ol.p, (size_t)nv);
  if (st == GUAC_OK) st = host_out(ctx, out_offsets, oo.p, (size_t)nv + 1);
  if (st == GUAC_OK) st = host_out(ctx, out_elements, oe.p, (size_t)n_elems);
  cudaStreamSynchronize(ctx->stream);
  if (st != GUAC_OK) return st;
  *n_out = nv;
  return GUAC_OK;
}

int guac_somatic_standard(guac_ctx* ctx, const guac_ss_params* params, const guac_range* loci, int32_t n_ranges,
                          guac_ss_call** out_calls, int64_t* n_calls, char** out_arena, int64_t* arena_len,
                          guac_locus_summary** out_summary, int64_t* out_n_visited) {
  if (!ctx || !params || !out_calls || !n_calls || !out_arena || !arena_len) return GUAC_E_INVALID_INPUT;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  Trace trace;
  using TCfg = TileCfg<GUAC_TK_TL>;
  // The streaming tile kernel serves the pileupFlatMapTwoSamples shape (sample ids {0, 1});
  // anything else, or opt_ss_impl = 1, takes the generic gather kernel.
  const bool two_samples = (params->normal_sample_id == 0 && params->tumor_sample_id == 1) ||
                           (params->normal_sample_id == 1 && params->tumor_sample_id == 0);
  const bool use_tiles = ctx->opt_ss_impl != 1 && two_samples;
  Plan plan;
  GUAC_TRY(make_plan(ctx, loci, n_ranges, plan, use_tiles ? TCfg::TL : TILE_LOCI));
  trace.mark("plan");
  *n_calls = 0;
  *arena_len = 0;
  if (out_n_visited) *out_n_visited = 0;
  SSParamsDev P;
  P.normal_sample = params->normal_sample_id;
  P.tumor_sample = params->tumor_sample_id;
  P.n_samples = std::max(params->normal_sample_id, params->tumor_sample_id) + 1;
  P.normal_odds = params->normal_odds_threshold;
  P.tumor_odds = params->tumor_odds_threshold;
  P.max_read_depth = params->max_read_depth;
  P.max_normal_alt = params->max_normal_alternate_read_depth;
  P.min_vaf = (double)params->min_tumor_variant_allele_frequency;
  P.min_alt_reads = 0;
  P.sample_mask = 0;
  if (P.normal_sample < 0 || P.tumor_sample < 0 || P.normal_sample == P.tumor_sample || P.n_samples > ctx->n_samples)
    return fail(ctx, GUAC_E_INVALID_INPUT, "bad normal/tumor sample ids");
  if (plan.n_tiles == 0) {
    GUAC_TRY(host_out(ctx, out_calls, nullptr, 0));
    GUAC_TRY(host_out(ctx, out_arena, nullptr, 0));
    if (out_summary) GUAC_TRY(host_out(ctx, out_summary, nullptr, 0));
    return GUAC_OK;
  }
  GUAC_TRY(reset_error(ctx));
  GUAC_TRY(ensure(ctx, ctx->b_visited, (size_t)plan.n_slots));
  // the dense per-(locus, sample) summary is part of the kernel's output contract; it stays in HBM unless asked for
  GUAC_TRY(ensure(ctx, ctx->b_summary, (size_t)plan.n_slots * 2 * sizeof(guac_locus_summary)));
  // the call's counters share one small block: one memset to clear them, one copy to read them
  struct SSCounters {
    int n_calls, n_fallback, n_heavy;
    unsigned int arena_used;
    unsigned long long n_visited;
  };
  GUAC_TRY(ensure(ctx, ctx->b_ncalls, sizeof(SSCounters)));
  SSCounters* d_cnt = ctx->b_ncalls.as<SSCounters>();
  int* d_ncalls = &d_cnt->n_calls;
  GUAC_TRY(pinned_block(ctx));
  SSCounters& h_cnt = *static_cast<SSCounters*>(ctx->h_counters);  // pinned: no staging copy, no extra wait
  h_cnt = SSCounters{};
  int max_calls = (int)std::max<size_t>(ctx->b_calls.cap / sizeof(SSCallDev), 1 << 16);
  int nc = 0;
  unsigned long long visited_host = 0;  // loci visited, counted by the tile kernels
  bool have_counts = false;             // nc / visited_host / the error record came with the list lengths
  for (int attempt = 0; attempt < 2; ++attempt) {
    have_counts = false;
    GUAC_TRY(ensure(ctx, ctx->b_calls, (size_t)max_calls * sizeof(SSCallDev)));
    CUDA_TRY(ctx, cudaMemsetAsync(d_cnt, 0, sizeof(SSCounters), ctx->stream));
    if (use_tiles) {
      GUAC_TRY(ensure(ctx, ctx->b_fallback, (size_t)(plan.n_tiles + 2) * sizeof(int)));
      int* fb_count = &d_cnt->n_fallback;
      int* fb_list = ctx->b_fallback.as<int>();
      // LOD-bound table for f = the VAF floor (see `prune` in tile_kernels.cuh)
      const int max_heavy = (int)std::min<int64_t>(plan.n_slots, plan.n_slots / 16 + 65536);
      GUAC_TRY(ensure(ctx, ctx->b_heavy, (size_t)max_heavy * sizeof(Tile) + 16));
      int* heavy_count = &d_cnt->n_heavy;
      unsigned long long* visited_count = &d_cnt->n_visited;
      TileArgs TA;
      std::memset(&TA, 0, sizeof TA);
      TA.rflags = ctx->b_rflags.as<uint8_t>();
      TA.hdr = ctx->b_hdr.as<uint4>();
      TA.refs = ctx->d_refs.as<RefView>();
      TA.tiles = ctx->b_tiles.as<Tile>();
      TA.visited = ctx->b_visited.as<uint8_t>();
      TA.summary = ctx->b_summary.as<guac_locus_summary>();
      TA.heavy = ctx->b_heavy.as<Tile>();
      TA.n_heavy = heavy_count;
      TA.max_heavy = max_heavy;
      TA.fallback_tiles = fb_list;
      TA.n_fallback = fb_count;
      TA.force_fallback_mod = (int)ctx->opt_force_fallback_mod;
      TA.n_visited = visited_count;
      TA.err = ctx->b_err.as<DeviceError>();
      if (ctx->opt_ss_impl == 2) {
        // stream implementation: the reads are compared with the reference by a kernel of their own, then the locus pass
        cudaEventRecord(ctx->ev0, ctx->stream);
        GUAC_TRY(run_read_pass(ctx));
        TA.ev_by_read = ctx->b_bins.as<uint2>();
        TA.ev_count = ctx->b_bin_count.as<uint8_t>();
        TA.ev_overflow = reinterpret_cast<const uint2*>(static_cast<const unsigned char*>(ctx->b_ev_overflow.p) + 64);
        ss_tile_kernel<TCfg, 0, true><<<(unsigned)plan.n_tiles, TCfg::THREADS, offsetof(TileSmem<TCfg>, cov2), ctx->stream>>>(ctx->R, P, TA);
      } else {
        cudaEventRecord(ctx->ev0, ctx->stream);
        ss_tile_kernel<TCfg, 0><<<(unsigned)plan.n_tiles, TCfg::THREADS, offsetof(TileSmem<TCfg>, cov2), ctx->stream>>>(ctx->R, P, TA);
      }
      ctx->launches += 1;
      CUDA_TRY(ctx, cudaGetLastError());
      // loci that need the full likelihood: one warp each.  Their number stays on the device: the launch is sized for a
      // full machine and strides over the list, so no host round trip separates it from the locus pass.
      somatic_standard_kernel<<<(unsigned)std::min<int64_t>(blocks_for(max_heavy, WARPS_PER_CTA), 148 * 8), CTA_THREADS, 0, ctx->stream>>>(
          ctx->R, ctx->d_refs.as<RefView>(), ctx->b_heavy.as<Tile>(), nullptr, max_heavy, 1, P, ctx->b_visited.as<uint8_t>(),
          ctx->b_summary.as<guac_locus_summary>(), ctx->b_calls.as<SSCallDev>(), d_ncalls, max_calls, nullptr,
          ctx->b_err.as<DeviceError>(), heavy_count);
      ctx->launches += 1;
      CUDA_TRY(ctx, cudaGetLastError());
      cudaEventRecord(ctx->ev1, ctx->stream);
      // one host round trip for everything the host needs to know now: the two list lengths, the number of calls, the
      // number of visited loci and the error record (every wait costs tens of microseconds, more on a busy host)
      CUDA_TRY(ctx, cudaMemcpyAsync(&h_cnt, d_cnt, sizeof h_cnt, cudaMemcpyDeviceToHost, ctx->stream));
      GUAC_TRY(check_device_error(ctx));  // synchronises
      have_counts = true;
      nc = h_cnt.n_calls;
      visited_host = h_cnt.n_visited;
      const int n_fb = h_cnt.n_fallback, n_heavy = h_cnt.n_heavy;
      ctx->last_fallback_tiles = n_fb;
      ctx->last_heavy_loci = n_heavy;
      if (n_heavy > max_heavy) {
        // more heavy loci than the list holds (pathological input): redo the call with the generic kernel
        const int64_t impl = ctx->opt_ss_impl;
        ctx->opt_ss_impl = 1;
        int st = guac_somatic_standard(ctx, params, loci, n_ranges, out_calls, n_calls, out_arena, arena_len, out_summary,
                                       out_n_visited);
        ctx->opt_ss_impl = impl;
        return st;
      }
      if (n_fb > 0) {  // tiles beyond a shared-memory capacity: generic kernel, still on the GPU
        somatic_standard_kernel<<<(unsigned)n_fb, CTA_THREADS, 0, ctx->stream>>>(
            ctx->R, ctx->d_refs.as<RefView>(), ctx->b_tiles.as<Tile>(), fb_list, n_fb, 0, P, ctx->b_visited.as<uint8_t>(),
            ctx->b_summary.as<guac_locus_summary>(), ctx->b_calls.as<SSCallDev>(), d_ncalls, max_calls,
            visited_count, ctx->b_err.as<DeviceError>());
        ctx->launches += 1;
        CUDA_TRY(ctx, cudaGetLastError());
        cudaEventRecord(ctx->ev1, ctx->stream);  // the timed span then ends here
        have_counts = false;                     // calls and visited loci of these tiles are still to come
      }
    } else {
      cudaEventRecord(ctx->ev0, ctx->stream);
      somatic_standard_kernel<<<(unsigned)plan.n_tiles, CTA_THREADS, 0, ctx->stream>>>(
          ctx->R, ctx->d_refs.as<RefView>(), ctx->b_tiles.as<Tile>(), nullptr, (int)plan.n_tiles, 0, P,
          ctx->b_visited.as<uint8_t>(), ctx->b_summary.as<guac_locus_summary>(), ctx->b_calls.as<SSCallDev>(),
          d_ncalls, max_calls, nullptr, ctx->b_err.as<DeviceError>());
      cudaEventRecord(ctx->ev1, ctx->stream);
      ctx->launches += 1;
      CUDA_TRY(ctx, cudaGetLastError());
    }
    trace.mark("per-locus kernels launched");
    ctx->have_timing = true;
    ctx->timed_loci = plan.n_slots;
    if (!have_counts) {
      CUDA_TRY(ctx, cudaMemcpyAsync(&h_cnt, d_cnt, sizeof h_cnt, cudaMemcpyDeviceToHost, ctx->stream));
      GUAC_TRY(check_device_error(ctx));  // synchronises
      nc = h_cnt.n_calls;
      visited_host = h_cnt.n_visited;
    }
    if (nc <= max_calls) break;
    max_calls = nc;  // rerun with room for every call
  }
  trace.mark("per-locus kernels done");
  // sparse part: the calls are ranked by slot (= (range order, locus) order), then one kernel fills the evidence and the
  // allele strings of every call and writes the finished guac_ss_call records in that order; they come back through
  // pinned memory as they are
  *arena_len = 0;
  const bool have_events = use_tiles && ctx->opt_ss_impl == 2;  // the read pass ran in this call: its Mismatch records are usable
  guac_ss_call* oc = nullptr;
  {
    void* hp = nullptr;
    GUAC_TRY(pinned_alloc(ctx, &hp, std::max<size_t>((size_t)nc * sizeof(guac_ss_call), 16)));
    oc = static_cast<guac_ss_call*>(hp);
  }
  if (nc > 0) {
    int st = ensure(ctx, ctx->b_arena, (size_t)nc * 32 + 4096);
    if (st == GUAC_OK) st = ensure(ctx, ctx->u_flag, (size_t)nc * sizeof(int));
    if (st == GUAC_OK) st = ensure(ctx, ctx->u_off, (size_t)nc * sizeof(guac_ss_call));
    if (st == GUAC_OK) {
      if (nc <= ctx->opt_ss_rank_max) {
        cudaMemsetAsync(ctx->u_flag.p, 0, (size_t)nc * sizeof(int), ctx->stream);
        ss_call_rank_kernel<<<dim3(blocks_for(nc, 256), SS_RANK_SPLIT), 256, 0, ctx->stream>>>(ctx->b_calls.as<SSCallDev>(), nc, ctx->u_flag.as<int>());
        ctx->launches += 1;
      } else {  // too many for the quadratic kernel: bitmap over the slots, linear in the loci of the call
        st = rank_by_slot(ctx, ctx->b_calls.p, sizeof(SSCallDev), offsetof(SSCallDev, slot), nc, plan.n_slots, ctx->u_flag.as<int>());
      }
    }
    for (int attempt = 0; st == GUAC_OK && attempt < 2; ++attempt) {
      unsigned int& used = *reinterpret_cast<unsigned int*>(static_cast<unsigned char*>(ctx->h_counters) + 128);
      used = 0;
      if (attempt > 0) cudaMemsetAsync(&d_cnt->arena_used, 0, sizeof(unsigned int), ctx->stream);  // (cleared with the block the first time)
      ss_evidence_kernel<<<(unsigned)nc, 128, 0, ctx->stream>>>(
          ctx->R, ctx->d_refs.as<RefView>(), P, ctx->b_calls.as<SSCallDev>(), ctx->b_arena.as<char>(), &d_cnt->arena_used,
          (unsigned int)std::min<size_t>(ctx->b_arena.cap, 0xFFFFFFFFu), ctx->b_err.as<DeviceError>(),
          have_events ? ctx->b_hdr.as<uint4>() : nullptr, have_events ? ctx->b_bins.as<uint2>() : nullptr,
          have_events ? ctx->b_bin_count.as<uint8_t>() : nullptr, STREAM_EV_CAP, ctx->u_flag.as<int>(), ctx->u_off.as<guac_ss_call>());
      ctx->launches += 1;
      if (cudaGetLastError() != cudaSuccess ||
          cudaMemcpyAsync(oc, ctx->u_off.p, (size_t)nc * sizeof(guac_ss_call), cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess ||
          cudaMemcpyAsync(&used, &d_cnt->arena_used, sizeof used, cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess) {
        st = fail(ctx, GUAC_E_CUDA, "launch of the evidence kernel failed");
        break;
      }
      st = check_device_error(ctx);  // synchronises
      if (st != GUAC_OK) break;
      *arena_len = used;
      if ((size_t)used <= ctx->b_arena.cap) break;
      st = ensure(ctx, ctx->b_arena, (size_t)used + 16);  // long indel alleles: grow and redo
    }
    if (st != GUAC_OK) {
      guac_free(ctx, oc);
      return st;
    }
  }
  trace.mark("evidence");
  *out_calls = oc;
  *n_calls = nc;
  GUAC_TRY(host_out(ctx, out_arena, ctx->b_arena.p, (size_t)*arena_len));
  trace.mark("calls out");
  int64_t nv = 0;
  if (out_summary || (out_n_visited && !use_tiles)) {
    GUAC_TRY(compact_visited(ctx, plan, 1, nv));
    if (out_n_visited) *out_n_visited = nv;
  } else if (out_n_visited) {  // counted by the tile kernels themselves, read back with the call count
    *out_n_visited = (int64_t)visited_host;
  }
  if (out_summary) {
    TmpDev os;
    GUAC_TRY(ensure(ctx, os, (size_t)nv * 2 * sizeof(guac_locus_summary)));
    // one uint64 per (locus, sample) record; rows of 2
    gather_rows_kernel<unsigned long long><<<blocks_for(plan.n_slots * 2, 256), 256, 0, ctx->stream>>>(
        ctx->b_summary.as<unsigned long long>(), ctx->b_visited.as<uint8_t>(), ctx->b_vis_idx.as<int64_t>(), plan.n_slots, 2,
        os.as<unsigned long long>());
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    int st = host_out(ctx, out_summary, os.p, (size_t)nv * 2);
    cudaStreamSynchronize(ctx->stream);
    if (st != GUAC_OK) return st;
  }
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  trace.mark("visited / summary");
  return GUAC_OK;
}

int guac_somatic_joint(guac_ctx* ctx, const guac_sj_params* params, const guac_range* loci, int32_t n_ranges,
                       const guac_range* force_loci, int32_t n_force, guac_sj_allele_evidence** out_alleles, int64_t* n_alleles,
                       guac_sj_sample_evidence** out_samples, int64_t* n_sample_evidences, char** out_arena, int64_t* arena_len,
                       int64_t* out_n_visited) {
  if (!ctx || !params || !out_alleles || !n_alleles || !out_samples || !n_sample_evidences || !out_arena || !arena_len)
    return GUAC_E_INVALID_INPUT;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  if (!ctx->have_reads) return fail(ctx, GUAC_E_STATE, "no reads uploaded (guac_reads_upload)");
  const int S = params->n_samples;
  if (S != ctx->n_samples || S <= 0 || S > GUAC_SJ_MAX_SAMPLES)
    return fail(ctx, GUAC_E_INVALID_INPUT, "guac_sj_params.n_samples must equal the uploaded batch's n_samples (<= 16)");
  if (params->max_alleles_per_site < 1 || params->max_alleles_per_site > SJ_MAX_ALLELES)
    return fail(ctx, GUAC_E_UNSUPPORTED, "max_alleles_per_site must be in 1..7");
  if (n_force < 0 || (n_force > 0 && !force_loci)) return fail(ctx, GUAC_E_INVALID_INPUT, "bad force-call loci");
  SJParamsDev P;
  std::memset(&P, 0, sizeof P);
  P.n_samples = S;
  for (int s = 0; s < S; ++s) {
    P.kind[s] = params->sample_kind[s];
    if (P.kind[s] < GUAC_SJ_NORMAL_DNA || P.kind[s] > GUAC_SJ_TUMOR_RNA) return fail(ctx, GUAC_E_INVALID_INPUT, "bad sample kind");
  }
  for (int s = 0; s < S; ++s)
    if (P.kind[s] == GUAC_SJ_NORMAL_DNA) P.discovery[P.n_discovery++] = s;
  const int n_normal = P.n_discovery;
  for (int s = 0; s < S; ++s)
    if (P.kind[s] == GUAC_SJ_TUMOR_DNA) P.discovery[P.n_discovery++] = s;
  if (n_normal == 0) return fail(ctx, GUAC_E_INVALID_INPUT, "at least one normal DNA sample is required");  // normalPileups.head (:88)
  P.max_alleles = params->max_alleles_per_site;
  P.min_reads = params->any_allele_min_supporting_reads;
  P.min_percent = params->any_allele_min_supporting_percent;
  P.vaf_floor = params->somatic_vaf_floor;
  // lociSetMinusOne (commands/SomaticJointCaller.scala:143-152): pileups are taken one locus before the alleles
  std::vector<guac_range> shifted;
  for (int i = 0; i < n_ranges; ++i) {
    guac_range r = loci[i];
    r.start = std::max<int64_t>(0, r.start - 1);
    r.end = r.end - 1;
    if (r.end > r.start) shifted.push_back(r);
  }
  Trace trace;
  using TCfg = TileCfg<GUAC_TK_TL>;
  Plan plan;
  GUAC_TRY(make_plan(ctx, shifted.data(), (int32_t)shifted.size(), plan, TCfg::TL));
  *n_alleles = 0;
  *n_sample_evidences = 0;
  *arena_len = 0;
  if (out_n_visited) *out_n_visited = 0;
  if (plan.n_tiles == 0) {
    GUAC_TRY(host_out(ctx, out_alleles, nullptr, 0));
    GUAC_TRY(host_out(ctx, out_samples, nullptr, 0));
    GUAC_TRY(host_out(ctx, out_arena, nullptr, 0));
    return GUAC_OK;
  }
  GUAC_TRY(reset_error(ctx));
  GUAC_TRY(ensure(ctx, ctx->b_visited, (size_t)plan.n_slots));
  GUAC_TRY(ensure(ctx, ctx->b_fallback, (size_t)(plan.n_tiles + 2) * sizeof(int)));
  GUAC_TRY(upload(ctx, ctx->j_force, force_loci, (size_t)n_force));
  GUAC_TRY(sj_logfact_table(ctx));
  // the call's counters: one block on the device, one pinned copy on the host
  struct SJCounters {
    int n_fallback, n_complex, n_simple, n_bail;
    int n_sites_fast, n_jobs_fast, n_sites_gen, n_jobs_gen;
    int n_emits, _pad;
    unsigned long long n_visited;
    long long total_alleles, total_string_bytes;
  };
  GUAC_TRY(ensure(ctx, ctx->b_ncalls, sizeof(SJCounters)));
  SJCounters* d_cnt = ctx->b_ncalls.as<SJCounters>();
  GUAC_TRY(pinned_block(ctx));
  SJCounters& h_cnt = *static_cast<SJCounters*>(ctx->h_counters);
  auto read_counters = [&]() -> int {
    CUDA_TRY(ctx, cudaMemcpyAsync(&h_cnt, d_cnt, sizeof(SJCounters), cudaMemcpyDeviceToHost, ctx->stream));
    return check_device_error(ctx);  // synchronises
  };
  // Dense half.  Stream form: the read pass leaves every read's non-Match elements in its slot, the locus pass (tile
  // kernel, MODE 2) marks the visited loci and lists the pileup loci x where an alternate allele can start at x + 1:
  // `simple` ones (single-base substitutions only; at most four samples) for the fast site kernel, the others — and
  // every force-called locus — for the generic per-locus kernels.
  bool every_contig_has_a_reference = true;
  for (auto& sp : ctx->contig_spans) every_contig_has_a_reference = every_contig_has_a_reference && ctx->contigs.count(sp.first) != 0;
  const bool stream = ctx->opt_sj_impl == 2 && every_contig_has_a_reference;
  const bool classify = ctx->opt_sj_impl != 0 && S <= SJF_SMAX;
  SSParamsDev SP;
  std::memset(&SP, 0, sizeof SP);
  SP.normal_sample = SP.tumor_sample = -1;
  SP.n_samples = S;
  SP.min_alt_reads = std::max(1, params->any_allele_min_supporting_reads);
  for (int d = 0; d < P.n_discovery; ++d) SP.sample_mask |= 1u << P.discovery[d];
  int* fb_list = ctx->b_fallback.as<int>();
  int max_heavy = (int)std::min<int64_t>(plan.n_slots, std::max<int64_t>(plan.n_slots / (classify ? 32 : 4) + 65536, (int64_t)(ctx->b_heavy.cap / sizeof(Tile)) - 1));
  int max_simple = classify ? (int)std::min<int64_t>(plan.n_slots, std::max<int64_t>(plan.n_slots / 16 + 65536, (int64_t)(ctx->j_simple.cap / sizeof(Tile)) - 1)) : 0;
  cudaEventRecord(ctx->ev0, ctx->stream);
  if (stream) GUAC_TRY(run_read_pass(ctx));
  for (int attempt = 0; attempt < 2; ++attempt) {
    GUAC_TRY(ensure(ctx, ctx->b_heavy, (size_t)max_heavy * sizeof(Tile) + 16));
    GUAC_TRY(ensure(ctx, ctx->j_simple, (size_t)max_simple * sizeof(Tile) + 16));
    CUDA_TRY(ctx, cudaMemsetAsync(d_cnt, 0, sizeof(SJCounters), ctx->stream));
    TileArgs TA;
    std::memset(&TA, 0, sizeof TA);
    TA.rflags = ctx->b_rflags.as<uint8_t>();
    TA.hdr = ctx->b_hdr.as<uint4>();
    TA.refs = ctx->d_refs.as<RefView>();
    TA.tiles = ctx->b_tiles.as<Tile>();
    TA.visited = ctx->b_visited.as<uint8_t>();
    TA.heavy = ctx->b_heavy.as<Tile>();
    TA.n_heavy = &d_cnt->n_complex;
    TA.max_heavy = max_heavy;
    TA.fallback_tiles = fb_list;
    TA.n_fallback = &d_cnt->n_fallback;
    TA.force_fallback_mod = (int)ctx->opt_force_fallback_mod;
    TA.n_visited = &d_cnt->n_visited;
    TA.err = ctx->b_err.as<DeviceError>();
    TA.force = ctx->j_force.as<guac_range>();
    TA.n_force = n_force;
    TA.sample_bits = classify ? 1 : 0;
    TA.simple = ctx->j_simple.as<Tile>();
    TA.n_simple = &d_cnt->n_simple;
    TA.max_simple = max_simple;
    TA.log_table = ctx->j_logtab.as<double2>();
    TA.het_prior = params->germline_neg_log10_het_prior;
    TA.hom_prior = params->germline_neg_log10_hom_alt_prior;
    TA.somatic_prior = params->somatic_neg_log10_variant_prior;
    TA.vaf_floor = params->somatic_vaf_floor;
    bool has_rna = false;
    for (int s = 0; s < S && s < 4; ++s) {
      TA.kinds |= (unsigned)P.kind[s] << (2 * s);
      has_rna = has_rna || P.kind[s] == GUAC_SJ_TUMOR_RNA;
    }
    // (with RNA samples the DNA prior of a site depends on their posteriors: no in-tile shortcut)
    TA.prune = classify && stream && !has_rna && ctx->opt_sj_prune ? 1 : 0;
    if (stream) {
      TA.ev_by_read = ctx->b_bins.as<uint2>();
      TA.ev_count = ctx->b_bin_count.as<uint8_t>();
      TA.ev_overflow = reinterpret_cast<const uint2*>(static_cast<const unsigned char*>(ctx->b_ev_overflow.p) + 64);
      ss_tile_kernel<TCfg, 2, true><<<(unsigned)plan.n_tiles, TCfg::THREADS, sizeof(TileSmem<TCfg>), ctx->stream>>>(ctx->R, SP, TA);
    } else {
      ss_tile_kernel<TCfg, 2, false><<<(unsigned)plan.n_tiles, TCfg::THREADS, sizeof(TileSmem<TCfg>), ctx->stream>>>(ctx->R, SP, TA);
    }
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    GUAC_TRY(read_counters());
    if (h_cnt.n_complex <= max_heavy && h_cnt.n_simple <= max_simple) break;
    max_heavy = std::max(max_heavy, h_cnt.n_complex);  // rerun with room for every candidate locus
    max_simple = std::max(max_simple, h_cnt.n_simple);
  }
  const int n_fb = h_cnt.n_fallback, n_complex = h_cnt.n_complex, n_simple = h_cnt.n_simple;
  const unsigned long long n_visited_tiles = h_cnt.n_visited;
  ctx->last_fallback_tiles = n_fb;
  ctx->last_heavy_loci = n_complex;
  ctx->stat_sj_simple = n_simple;
  trace.mark("dense prefilter");
  const size_t raw_stride = offsetof(SJEvidenceDev, slots) + (size_t)(S + 2) * sizeof(SJSlotDev);
  // fast site kernel over the simple candidates: sites, alleles and raw evidence in one pass
  int n_sites_fast = 0, n_bail = 0;
  if (n_simple > 0) {
    GUAC_TRY(ensure(ctx, ctx->j_sites_fast, (size_t)n_simple * sizeof(SJSiteDev)));
    GUAC_TRY(ensure(ctx, ctx->j_bail, (size_t)n_simple * sizeof(Tile)));
    int max_jobs = (int)std::max<int64_t>((int64_t)n_simple + n_simple / 4 + 1024, (int64_t)(ctx->j_raw_fast.cap / raw_stride));
    for (int attempt = 0; attempt < 2; ++attempt) {
      GUAC_TRY(ensure(ctx, ctx->j_raw_fast, (size_t)max_jobs * raw_stride));
      if (attempt > 0) CUDA_TRY(ctx, cudaMemsetAsync(&d_cnt->n_bail, 0, 3 * sizeof(int), ctx->stream));  // n_bail, n_sites_fast, n_jobs_fast
      SJFArgs FA;
      std::memset(&FA, 0, sizeof FA);
      FA.hdr = ctx->b_hdr.as<uint4>();
      FA.log_table = ctx->j_logtab.as<double2>();
      FA.refs = ctx->d_refs.as<RefView>();
      FA.cands = ctx->j_simple.as<Tile>();
      FA.n_cands = &d_cnt->n_simple;
      FA.max_cands = n_simple;
      FA.force = ctx->j_force.as<guac_range>();
      FA.n_force = n_force;
      FA.sites = ctx->j_sites_fast.as<SJSiteDev>();
      FA.n_sites = &d_cnt->n_sites_fast;
      FA.max_sites = n_simple;
      FA.raw = ctx->j_raw_fast.as<unsigned char>();
      FA.raw_stride = raw_stride;
      FA.n_jobs = &d_cnt->n_jobs_fast;
      FA.max_jobs = max_jobs;
      FA.bail = ctx->j_bail.as<Tile>();
      FA.n_bail = &d_cnt->n_bail;
      FA.max_bail = n_simple;
      FA.err = ctx->b_err.as<DeviceError>();
      const unsigned grid = (unsigned)std::min<int64_t>(blocks_for(n_simple, SJF_WARPS), 148 * 16 * 8 / SJF_WARPS);
      sj_snv_site_kernel<<<grid, 32 * SJF_WARPS, SJF_WARPS * sizeof(SJFWarpSmem), ctx->stream>>>(ctx->R, P, *params, FA);
      ctx->launches += 1;
      CUDA_TRY(ctx, cudaGetLastError());
      GUAC_TRY(read_counters());
      if (h_cnt.n_jobs_fast <= max_jobs) break;
      max_jobs = h_cnt.n_jobs_fast;
    }
    n_sites_fast = h_cnt.n_sites_fast;
    n_bail = h_cnt.n_bail;
  }
  ctx->stat_sj_bail = n_bail;
  trace.mark("fast site kernel");
  // generic per-locus kernels: whole tiles the locus pass declined, the complex candidates, the sites the fast kernel
  // handed back
  int n_sites_gen = 0, n_jobs_gen = 0;
  const int64_t max_sites_gen = (int64_t)n_fb * TCfg::TL + n_complex + n_bail;
  if (max_sites_gen > 0) {
    GUAC_TRY(ensure(ctx, ctx->b_sj_sites, (size_t)max_sites_gen * sizeof(SJSiteDev)));
    const int max_sites = (int)std::min<int64_t>(max_sites_gen, INT32_MAX);
    if (n_fb > 0) {
      sj_discover_kernel<<<(unsigned)n_fb, CTA_THREADS, 0, ctx->stream>>>(
          ctx->R, ctx->d_refs.as<RefView>(), ctx->b_tiles.as<Tile>(), fb_list, n_fb, 0, P, ctx->j_force.as<guac_range>(), n_force,
          ctx->b_visited.as<uint8_t>(), ctx->b_sj_sites.as<SJSiteDev>(), &d_cnt->n_sites_gen, max_sites, &d_cnt->n_jobs_gen,
          ctx->b_err.as<DeviceError>());
      ctx->launches += 1;
    }
    if (n_complex > 0) {
      sj_discover_kernel<<<blocks_for(n_complex, WARPS_PER_CTA), CTA_THREADS, 0, ctx->stream>>>(
          ctx->R, ctx->d_refs.as<RefView>(), ctx->b_heavy.as<Tile>(), nullptr, n_complex, 1, P, ctx->j_force.as<guac_range>(), n_force,
          nullptr, ctx->b_sj_sites.as<SJSiteDev>(), &d_cnt->n_sites_gen, max_sites, &d_cnt->n_jobs_gen, ctx->b_err.as<DeviceError>());
      ctx->launches += 1;
    }
    if (n_bail > 0) {
      sj_discover_kernel<<<blocks_for(n_bail, WARPS_PER_CTA), CTA_THREADS, 0, ctx->stream>>>(
          ctx->R, ctx->d_refs.as<RefView>(), ctx->j_bail.as<Tile>(), nullptr, n_bail, 1, P, ctx->j_force.as<guac_range>(), n_force,
          nullptr, ctx->b_sj_sites.as<SJSiteDev>(), &d_cnt->n_sites_gen, max_sites, &d_cnt->n_jobs_gen, ctx->b_err.as<DeviceError>());
      ctx->launches += 1;
    }
    CUDA_TRY(ctx, cudaGetLastError());
    GUAC_TRY(read_counters());
    n_sites_gen = h_cnt.n_sites_gen;
    n_jobs_gen = h_cnt.n_jobs_gen;
    if (n_sites_gen > 0) {
      GUAC_TRY(ensure(ctx, ctx->j_raw_gen, (size_t)std::max(n_jobs_gen, 1) * raw_stride));
      const int64_t ctas = (int64_t)n_sites_gen * P.max_alleles;
      sj_evidence_kernel<<<(unsigned)ctas, 128, 0, ctx->stream>>>(ctx->R, ctx->d_refs.as<RefView>(), P, ctx->b_sj_sites.as<SJSiteDev>(), 0,
                                                                  ctx->j_raw_gen.as<unsigned char>(), raw_stride, ctx->b_err.as<DeviceError>());
      ctx->launches += 1;
      CUDA_TRY(ctx, cudaGetLastError());
    }
  }
  ctx->stat_sj_sites = (int64_t)n_sites_fast + n_sites_gen;
  trace.mark("generic kernels");
  // posteriors, calls, annotations, emit policy: one thread per site; only emitted sites are listed
  const int n_sites_all = n_sites_fast + n_sites_gen;
  int n_emits = 0;
  if (n_sites_all > 0) {
    GUAC_TRY(ensure(ctx, ctx->j_emits, (size_t)n_sites_all * sizeof(SJEmitDev)));
    if (n_sites_fast > 0) {
      sj_posterior_kernel<<<blocks_for(n_sites_fast, 128), 128, 0, ctx->stream>>>(
          *params, ctx->j_logfact.as<double>(), SJ_LOGFACT_N, ctx->j_sites_fast.as<SJSiteDev>(), n_sites_fast, 0,
          ctx->j_raw_fast.as<unsigned char>(), raw_stride, ctx->j_emits.as<SJEmitDev>(), &d_cnt->n_emits);
      ctx->launches += 1;
    }
    if (n_sites_gen > 0) {
      sj_posterior_kernel<<<blocks_for(n_sites_gen, 128), 128, 0, ctx->stream>>>(
          *params, ctx->j_logfact.as<double>(), SJ_LOGFACT_N, ctx->b_sj_sites.as<SJSiteDev>(), n_sites_gen, SJ_SITE_GENERIC,
          ctx->j_raw_gen.as<unsigned char>(), raw_stride, ctx->j_emits.as<SJEmitDev>(), &d_cnt->n_emits);
      ctx->launches += 1;
    }
    CUDA_TRY(ctx, cudaGetLastError());
    GUAC_TRY(read_counters());
    n_emits = h_cnt.n_emits;
  }
  ctx->stat_sj_emitted_sites = n_emits;
  trace.mark("posteriors");
  // emitted sites in (range order, locus) = slot order; offsets of their records and strings by two scans; the finished
  // records written on the device
  int64_t total_alleles = 0, total_strings = 0;
  if (n_emits > 0) {
    GUAC_TRY(ensure(ctx, ctx->r_rank, (size_t)n_emits * sizeof(int)));
    GUAC_TRY(ensure(ctx, ctx->j_order, (size_t)n_emits * sizeof(int)));
    GUAC_TRY(ensure(ctx, ctx->j_cnt, (size_t)n_emits * 2 * sizeof(int32_t)));
    GUAC_TRY(ensure(ctx, ctx->j_off, (size_t)(n_emits + 1) * 2 * sizeof(int64_t)));
    GUAC_TRY(rank_by_slot(ctx, ctx->j_emits.p, sizeof(SJEmitDev), offsetof(SJEmitDev, slot), n_emits, plan.n_slots, ctx->r_rank.as<int>()));
    int32_t* n_by_rank = ctx->j_cnt.as<int32_t>();
    int32_t* len_by_rank = n_by_rank + n_emits;
    int64_t* allele_off = ctx->j_off.as<int64_t>();
    int64_t* string_off = allele_off + n_emits + 1;
    sj_order_kernel<<<blocks_for(n_emits, 256), 256, 0, ctx->stream>>>(ctx->j_emits.as<SJEmitDev>(), ctx->r_rank.as<int>(), n_emits,
                                                                       ctx->j_order.as<int>(), n_by_rank, len_by_rank);
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    GUAC_TRY((device_scan<true, int32_t>(ctx, n_by_rank, allele_off, n_emits, OpSum(), (int64_t*)&d_cnt->total_alleles)));
    GUAC_TRY((device_scan<true, int32_t>(ctx, len_by_rank, string_off, n_emits, OpSum(), (int64_t*)&d_cnt->total_string_bytes)));
    GUAC_TRY(read_counters());
    total_alleles = h_cnt.total_alleles;
    total_strings = h_cnt.total_string_bytes;
    if (total_strings >= ((int64_t)1 << 32)) return fail(ctx, GUAC_E_UNSUPPORTED, "more than 4 GiB of allele strings in one call: split the loci");
    GUAC_TRY(ensure(ctx, ctx->j_out_alleles, (size_t)total_alleles * sizeof(guac_sj_allele_evidence)));
    GUAC_TRY(ensure(ctx, ctx->j_out_samples, (size_t)total_alleles * (S + 2) * sizeof(guac_sj_sample_evidence)));
    GUAC_TRY(ensure(ctx, ctx->b_arena, (size_t)total_strings + 16));
    sj_final_kernel<<<blocks_for(n_emits, 128), 128, 0, ctx->stream>>>(
        ctx->R, ctx->d_refs.as<RefView>(), *params, ctx->j_logfact.as<double>(), SJ_LOGFACT_N, ctx->j_sites_fast.as<SJSiteDev>(),
        ctx->b_sj_sites.as<SJSiteDev>(), ctx->j_raw_fast.as<unsigned char>(), ctx->j_raw_gen.as<unsigned char>(), raw_stride,
        ctx->j_emits.as<SJEmitDev>(), ctx->j_order.as<int>(), n_emits, allele_off, string_off,
        ctx->j_out_alleles.as<guac_sj_allele_evidence>(), ctx->j_out_samples.as<guac_sj_sample_evidence>(), ctx->b_arena.as<char>());
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
  }
  cudaEventRecord(ctx->ev1, ctx->stream);
  ctx->have_timing = true;
  ctx->timed_loci = plan.n_slots;
  trace.mark("records on the device");
  // only the emitted records cross PCIe
  GUAC_TRY(host_out(ctx, out_alleles, ctx->j_out_alleles.p, (size_t)total_alleles));
  {
    int st = host_out(ctx, out_samples, ctx->j_out_samples.p, (size_t)total_alleles * (S + 2));
    if (st == GUAC_OK) st = host_out(ctx, out_arena, ctx->b_arena.p, (size_t)total_strings);
    if (st == GUAC_OK) st = check_device_error(ctx);  // synchronises
    if (st != GUAC_OK) return st;
  }
  *n_alleles = total_alleles;
  *n_sample_evidences = total_alleles * (S + 2);
  *arena_len = total_strings;
  if (out_n_visited) {
    if (n_fb == 0) {
      *out_n_visited = (int64_t)n_visited_tiles;  // counted by the locus pass itself
    } else {
      int64_t nv = 0;
      GUAC_TRY(compact_visited(ctx, plan, 1, nv));
      *out_n_visited = nv;
    }
  }
  trace.mark("records out");
  return GUAC_OK;
}

int guac_germline_standard(guac_ctx* ctx, const guac_gs_params* params, const guac_range* loci, int32_t n_ranges,
                           guac_called_allele** out_calls, int64_t* n_calls, char** out_arena, int64_t* arena_len,
                           int64_t** out_active_loci, int64_t* n_active, int64_t* out_n_visited) {
  if (!ctx || !params || !out_calls || !n_calls || !out_arena || !arena_len || !out_active_loci || !n_active) return GUAC_E_INVALID_INPUT;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  using TCfg = TileCfg<GUAC_TK_TL>;
  Trace trace;
  if (!ctx->have_reads) return fail(ctx, GUAC_E_STATE, "no reads uploaded (guac_reads_upload)");
  if (params->sample_id < 0 || params->sample_id >= ctx->n_samples) return fail(ctx, GUAC_E_INVALID_INPUT, "bad sample id");
  Plan plan;
  GUAC_TRY(make_plan(ctx, loci, n_ranges, plan, TCfg::TL));
  *n_calls = 0;
  *arena_len = 0;
  *n_active = 0;
  if (out_n_visited) *out_n_visited = 0;
  if (plan.n_tiles == 0) {
    GUAC_TRY(host_out(ctx, out_calls, nullptr, 0));
    GUAC_TRY(host_out(ctx, out_arena, nullptr, 0));
    GUAC_TRY(host_out(ctx, out_active_loci, nullptr, 0));
    return GUAC_OK;
  }
  GLParamsDev G;
  G.sample = params->sample_id;
  G.min_alt_reads = params->min_alt_reads;
  G.respect_active_regions = params->respect_active_regions;
  G.min_area_vaf = params->min_area_vaf;
  SSParamsDev P;
  std::memset(&P, 0, sizeof P);
  P.normal_sample = -1;
  P.tumor_sample = params->sample_id;
  P.n_samples = ctx->n_samples;
  P.min_alt_reads = params->min_alt_reads;
  GUAC_TRY(reset_error(ctx));
  GUAC_TRY(ensure(ctx, ctx->b_visited, (size_t)plan.n_slots));
  GUAC_TRY(ensure(ctx, ctx->b_summary, (size_t)plan.n_slots * 2 * sizeof(guac_locus_summary)));
  GUAC_TRY(ensure(ctx, ctx->b_fallback, (size_t)(plan.n_tiles + 2) * sizeof(int)));
  GUAC_TRY(ensure(ctx, ctx->b_vcount, sizeof(unsigned long long)));
  GUAC_TRY(ensure(ctx, ctx->b_boundtab, 256 * sizeof(double)));
  GUAC_TRY(ensure(ctx, ctx->b_ncalls, 2 * sizeof(int)));
  int* fb_count = ctx->b_fallback.as<int>();
  int* fb_list = fb_count + 1;
  int* heavy_count = fb_count + 1 + plan.n_tiles;
  const int max_heavy = (int)std::min<int64_t>(plan.n_slots, plan.n_slots / 8 + 65536);
  GUAC_TRY(ensure(ctx, ctx->b_heavy, (size_t)max_heavy * sizeof(Tile) + 16));
  CUDA_TRY(ctx, cudaMemsetAsync(fb_count, 0, sizeof(int), ctx->stream));
  CUDA_TRY(ctx, cudaMemsetAsync(heavy_count, 0, sizeof(int), ctx->stream));
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->b_vcount.p, 0, sizeof(unsigned long long), ctx->stream));
  TileArgs TA;
  std::memset(&TA, 0, sizeof TA);
  TA.rflags = ctx->b_rflags.as<uint8_t>();
  TA.hdr = ctx->b_hdr.as<uint4>();
  TA.refs = ctx->d_refs.as<RefView>();
  TA.tiles = ctx->b_tiles.as<Tile>();
  TA.visited = ctx->b_visited.as<uint8_t>();
  TA.summary = ctx->b_summary.as<guac_locus_summary>();
  TA.heavy = ctx->b_heavy.as<Tile>();
  TA.n_heavy = heavy_count;
  TA.max_heavy = max_heavy;
  TA.fallback_tiles = fb_list;
  TA.n_fallback = fb_count;
  TA.force_fallback_mod = (int)ctx->opt_force_fallback_mod;
  TA.n_visited = ctx->b_vcount.as<unsigned long long>();
  TA.err = ctx->b_err.as<DeviceError>();
  cudaEventRecord(ctx->ev0, ctx->stream);
  ss_tile_kernel<TCfg, 1><<<(unsigned)plan.n_tiles, TCfg::THREADS, offsetof(TileSmem<TCfg>, cov2), ctx->stream>>>(ctx->R, P, TA);
  ctx->launches += 1;
  CUDA_TRY(ctx, cudaGetLastError());
  int counts[2] = {0, 0};
  CUDA_TRY(ctx, cudaMemcpyAsync(&counts[0], fb_count, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->last_fallback_tiles = counts[0];
  if (counts[0] > 0) {
    germline_dense_fallback_kernel<<<(unsigned)counts[0], CTA_THREADS, 0, ctx->stream>>>(
        ctx->R, ctx->d_refs.as<RefView>(), ctx->b_tiles.as<Tile>(), fb_list, G, ctx->b_visited.as<uint8_t>(),
        ctx->b_summary.as<guac_locus_summary>(), ctx->b_heavy.as<Tile>(), heavy_count, max_heavy,
        ctx->b_vcount.as<unsigned long long>(), ctx->b_err.as<DeviceError>());
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
  }
  CUDA_TRY(ctx, cudaMemcpyAsync(&counts[1], heavy_count, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  const int n_heavy = counts[1];
  ctx->last_heavy_loci = n_heavy;
  if (n_heavy > max_heavy) return fail(ctx, GUAC_E_UNSUPPORTED, "more candidate loci than the heavy-locus list holds; split the loci ranges");
  int nc = 0, na = 0;
  int* d_nc = ctx->b_ncalls.as<int>();
  const int max_calls = std::max(n_heavy, 1);
  GUAC_TRY(ensure(ctx, ctx->b_calls, (size_t)max_calls * sizeof(GLCallDev)));
  DevBuf& d_active = ctx->g_active;  // scratch of this entry point lives in the context: no cudaMalloc / cudaFree per call
  GUAC_TRY(ensure(ctx, d_active, (size_t)max_calls * 2 * sizeof(long long)));
  CUDA_TRY(ctx, cudaMemsetAsync(d_nc, 0, 2 * sizeof(int), ctx->stream));
  if (n_heavy > 0) {
    germline_locus_kernel<<<blocks_for(n_heavy, WARPS_PER_CTA), CTA_THREADS, 0, ctx->stream>>>(
        ctx->R, ctx->d_refs.as<RefView>(), ctx->b_heavy.as<Tile>(), n_heavy, G, ctx->b_calls.as<GLCallDev>(), d_nc, max_calls,
        d_active.as<long long>(), d_nc + 1, max_calls, ctx->b_err.as<DeviceError>());
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
  }
  cudaEventRecord(ctx->ev1, ctx->stream);
  ctx->have_timing = true;
  ctx->timed_loci = plan.n_slots;
  int both[2] = {0, 0};
  CUDA_TRY(ctx, cudaMemcpyAsync(both, d_nc, 2 * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  {
    int st = check_device_error(ctx);
    if (st != GUAC_OK) return st;
  }
  nc = both[0];
  na = both[1];
  trace.mark("per-locus kernels");
  // the records come over through pinned memory and stay where they land; (range order, locus) order = slot order is
  // an index sort
  struct Pinned {
    guac_ctx* ctx;
    std::vector<void*> p;
    ~Pinned() {
      for (void* q : p) guac_free(ctx, q);
    }
    int get(void** out, size_t bytes) {
      GUAC_TRY(pinned_alloc(ctx, out, std::max<size_t>(bytes, 16)));
      p.push_back(*out);
      return GUAC_OK;
    }
  } pinned{ctx, {}};
  GLCallDev* calls_raw = nullptr;
  GUAC_TRY(pinned.get(reinterpret_cast<void**>(&calls_raw), (size_t)nc * sizeof(GLCallDev)));
  std::vector<long long> active((size_t)na * 2);
  int* rank = nullptr;  // position of each record in slot order, counted on the device (no sort on the host)
  GUAC_TRY(pinned.get(reinterpret_cast<void**>(&rank), (size_t)nc * sizeof(int)));
  GUAC_TRY(ensure(ctx, ctx->r_rank, (size_t)std::max(nc, 1) * sizeof(int)));
  GUAC_TRY(rank_by_slot(ctx, ctx->b_calls.p, sizeof(GLCallDev), offsetof(GLCallDev, slot), nc, plan.n_slots, ctx->r_rank.as<int>()));
  if (nc) CUDA_TRY(ctx, cudaMemcpyAsync(calls_raw, ctx->b_calls.p, (size_t)nc * sizeof(GLCallDev), cudaMemcpyDeviceToHost, ctx->stream));
  if (nc) CUDA_TRY(ctx, cudaMemcpyAsync(rank, ctx->r_rank.p, (size_t)nc * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  if (na) CUDA_TRY(ctx, cudaMemcpyAsync(active.data(), d_active.p, (size_t)na * 2 * sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  std::vector<std::pair<int64_t, int>> order((size_t)nc);
  for (int i = 0; i < nc; ++i) order[(size_t)rank[i]] = {calls_raw[i].slot, i};
  struct CallsInOrder {  // range-for over the records in slot order
    const GLCallDev* raw;
    const std::vector<std::pair<int64_t, int>>* order;
    struct It {
      const CallsInOrder* c;
      size_t i;
      const GLCallDev& operator*() const { return c->raw[(*c->order)[i].second]; }
      It& operator++() { ++i; return *this; }
      bool operator!=(const It& o) const { return i != o.i; }
    };
    It begin() const { return It{this, 0}; }
    It end() const { return It{this, order->size()}; }
  } calls{calls_raw, &order};
  // active loci sorted by (contig, locus) like the visiting order within a contig
  {
    std::vector<std::pair<long long, long long>> v;
    for (int i = 0; i < na; ++i) v.push_back({active[2 * i], active[2 * i + 1]});
    std::sort(v.begin(), v.end());
    for (int i = 0; i < na; ++i) {
      active[2 * i] = v[i].first;
      active[2 * i + 1] = v[i].second;
    }
  }
  trace.mark("records to the host, sorted");
  // AlleleEvidence of every called allele
  size_t n_jobs = 0;
  for (const GLCallDev& c : calls) n_jobs += (size_t)c.n_called;
  EvidenceJob* jobs = nullptr;
  guac_allele_evidence* evs = nullptr;
  GUAC_TRY(pinned.get(reinterpret_cast<void**>(&jobs), n_jobs * sizeof(EvidenceJob)));
  GUAC_TRY(pinned.get(reinterpret_cast<void**>(&evs), n_jobs * sizeof(guac_allele_evidence)));
  {
    size_t k = 0;
    for (const GLCallDev& c : calls)
      for (int t = 0; t < c.n_called; ++t) {
        EvidenceJob& j = jobs[k++];
        j.locus = c.locus;
        j.read_lo = c.read_lo;
        j.read_hi = c.read_hi;
        j.key = c.called[t].key;
        j.contig_id = c.contig_id;
        j.ref_index = c.ref_index;
        j.sample = params->sample_id;
        j._pad = 0;
        j.probability = std::exp(c.log_probability);  // :247
      }
  }
  if (n_jobs) {
    DevBuf &d_jobs = ctx->g_jobs, &d_evs = ctx->g_evs;
    GUAC_TRY(upload(ctx, d_jobs, jobs, n_jobs));
    GUAC_TRY(ensure(ctx, d_evs, n_jobs * sizeof(guac_allele_evidence)));
    allele_evidence_kernel<<<(unsigned)n_jobs, 128, 0, ctx->stream>>>(ctx->R, ctx->d_refs.as<RefView>(), d_jobs.as<EvidenceJob>(),
                                                                     d_evs.as<guac_allele_evidence>(), ctx->b_err.as<DeviceError>());
    ctx->launches += 1;
    CUDA_TRY(ctx, cudaGetLastError());
    CUDA_TRY(ctx, cudaMemcpyAsync(evs, d_evs.p, n_jobs * sizeof(guac_allele_evidence), cudaMemcpyDeviceToHost, ctx->stream));
    GUAC_TRY(check_device_error(ctx));
  }
  trace.mark("evidence");
  // keep calls with phredScaledLikelihood > min (:147); lay out the allele strings
  auto phred = [](double probability) {  // util/PhredUtils.scala:7-14 over ADAM's successProbabilityToPhred
    const double p = std::min(1 - 1e-16, probability);
    return (int)std::floor(-10.0 * std::log10(1.0 - p) + 0.5);
  };
  guac_called_allele* out = nullptr;  // at most one record per job, straight into the pinned buffer the caller receives
  {
    void* hp = nullptr;
    GUAC_TRY(pinned_alloc(ctx, &hp, std::max<size_t>(n_jobs * sizeof(guac_called_allele), 16)));
    out = static_cast<guac_called_allele*>(hp);
  }
  size_t n_out = 0, n_strings = 0;
  SJStringJob* strings = nullptr;
  {
    void* hp = nullptr;
    int st = pinned.get(&hp, 2 * n_jobs * sizeof(SJStringJob));
    if (st != GUAC_OK) {
      guac_free(ctx, out);
      return st;
    }
    strings = static_cast<SJStringJob*>(hp);
  }
  uint32_t arena_off = 0;
  auto add_string = [&](int kind, int64_t a, int32_t b, int32_t len, uint32_t& off, uint32_t& out_len) {
    SJStringJob j;
    j.a = a;
    j.b = b;
    j.len = len;
    j.arena_off = arena_off;
    j.kind = kind;
    strings[n_strings++] = j;
    off = arena_off;
    out_len = (uint32_t)len;
    arena_off += (uint32_t)len;
  };
  size_t ji = 0;
  for (const GLCallDev& c : calls)
    for (int t = 0; t < c.n_called; ++t, ++ji) {
      if (!(phred(evs[ji].probability) > params->min_phred_scaled_likelihood)) continue;
      const GLAlleleDev& a = c.called[t];
      guac_called_allele g;
      std::memset(&g, 0, sizeof g);
      g.contig_id = c.contig_id;
      g.n_genotypes = c.n_genotypes;
      g.start = c.locus;
      g.genotype_log_probability = c.log_probability;
      g.evidence = evs[ji];
      if (a.kind == GUAC_ALN_MISMATCH) {
        add_string(1, c.locus, c.ref_index, 1, g.ref_off, g.ref_len);
        add_string(3, 0, a.base, 1, g.alt_off, g.alt_len);
      } else if (a.kind == GUAC_ALN_INSERTION) {
        add_string(0, a.rep_read, a.rep_rp, 1, g.ref_off, g.ref_len);
        add_string(0, a.rep_read, a.rep_rp, a.indel_len + 1, g.alt_off, g.alt_len);
      } else {  // Deletion
        add_string(1, c.locus, c.ref_index, a.indel_len + 1, g.ref_off, g.ref_len);
        add_string(1, c.locus, c.ref_index, 1, g.alt_off, g.alt_len);
      }
      out[n_out++] = g;
    }
  *arena_len = arena_off;
  int st_tail = GUAC_OK;
  if (n_strings) {
    DevBuf& d_strings = ctx->g_strings;
    st_tail = upload(ctx, d_strings, strings, n_strings);
    if (st_tail == GUAC_OK) st_tail = ensure(ctx, ctx->b_arena, arena_off + 16);
    if (st_tail == GUAC_OK) {
      sj_strings_kernel<<<blocks_for((int64_t)n_strings, 128), 128, 0, ctx->stream>>>(ctx->R, ctx->d_refs.as<RefView>(),
                                                                                      d_strings.as<SJStringJob>(), (int)n_strings,
                                                                                      ctx->b_arena.as<char>());
      ctx->launches += 1;
      if (cudaGetLastError() != cudaSuccess) st_tail = fail(ctx, GUAC_E_CUDA, "launch of the string kernel failed");
    }
  }
  trace.mark("strings");
  void* hact = nullptr;
  if (st_tail == GUAC_OK) st_tail = host_out(ctx, out_arena, ctx->b_arena.p, (size_t)arena_off);
  if (st_tail == GUAC_OK) st_tail = pinned_alloc(ctx, &hact, std::max<size_t>(active.size() * sizeof(int64_t), 16));
  if (st_tail != GUAC_OK) {
    guac_free(ctx, out);
    return st_tail;
  }
  if (!active.empty()) std::memcpy(hact, active.data(), active.size() * sizeof(int64_t));
  *out_calls = out;
  *n_calls = (int64_t)n_out;
  *out_active_loci = (int64_t*)hact;
  *n_active = na;
  if (out_n_visited) {
    unsigned long long v = 0;
    CUDA_TRY(ctx, cudaMemcpyAsync(&v, ctx->b_vcount.p, sizeof v, cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    *out_n_visited = (int64_t)v;
  }
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  trace.mark("out");
  return GUAC_OK;
}

void* guac_stream(guac_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

int guac_set_option(guac_ctx* ctx, const char* name, int64_t value) {
  if (!ctx || !name) return GUAC_E_INVALID_INPUT;
  const std::string n(name);
  if (n == "ss_impl") {
    if (value < 0 || value > 2) return fail(ctx, GUAC_E_INVALID_INPUT, "ss_impl must be 0 (tile), 1 (gather) or 2 (stream)");
    ctx->opt_ss_impl = value;
  } else if (n == "ss_force_fallback_mod") {
    ctx->opt_force_fallback_mod = value < 0 ? 0 : value;
  } else if (n == "sj_host_threads") {
    ctx->opt_sj_host_threads = value < 0 ? 0 : std::min<int64_t>(value, 256);  // accepted for old callers; nothing runs on host threads any more
  } else if (n == "synth_profile") {
    if (value < 0 || value > 1) return fail(ctx, GUAC_E_INVALID_INPUT, "synth_profile must be 0 or 1");
    ctx->opt_synth_profile = value;
  } else if (n == "sj_prune") {
    ctx->opt_sj_prune = value != 0;
  } else if (n == "sj_impl") {
    if (value < 0 || value > 2) return fail(ctx, GUAC_E_INVALID_INPUT, "sj_impl must be 0, 1 or 2");
    ctx->opt_sj_impl = value;
  } else if (n == "compact_min_qual_bits") {
    if (value != 2 && value != 4 && value != 8) return fail(ctx, GUAC_E_INVALID_INPUT, "compact_min_qual_bits must be 2, 4 or 8");
    ctx->opt_compact_min_qual_bits = value;
  } else if (n == "compact_ref_relative") {
    ctx->opt_compact_ref_relative = value != 0;
  } else if (n == "compact_narrow") {
    ctx->opt_compact_narrow = value != 0;
  } else if (n == "ss_rank_max") {
    ctx->opt_ss_rank_max = value < 0 ? SS_RANK_MAX : std::min<int64_t>(value, SS_RANK_MAX);
  } else {
    return fail(ctx, GUAC_E_INVALID_INPUT, "unknown option " + n);
  }
  return GUAC_OK;
}

int64_t guac_last_fallback_tiles(guac_ctx* ctx) { return ctx ? ctx->last_fallback_tiles : 0; }
int64_t guac_last_heavy_loci(guac_ctx* ctx) { return ctx ? ctx->last_heavy_loci : 0; }

int64_t guac_last_stat(guac_ctx* ctx, const char* name) {
  if (!ctx || !name) return -1;
  const std::string n(name);
  if (n == "fallback_tiles") return ctx->last_fallback_tiles;
  if (n == "heavy_loci") return ctx->last_heavy_loci;
  if (n == "sj_simple") return ctx->stat_sj_simple;
  if (n == "sj_bail") return ctx->stat_sj_bail;
  if (n == "sj_sites") return ctx->stat_sj_sites;
  if (n == "sj_emitted_sites") return ctx->stat_sj_emitted_sites;
  return -1;
}

int guac_last_kernel_ms(guac_ctx* ctx, float* ms, int64_t* n_loci) {
  if (!ctx || !ms) return GUAC_E_INVALID_INPUT;
  if (!ctx->have_timing) return fail(ctx, GUAC_E_STATE, "no timed kernel yet");
  CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev1));
  CUDA_TRY(ctx, cudaEventElapsedTime(ms, ctx->ev0, ctx->ev1));
  if (n_loci) *n_loci = ctx->timed_loci;
  return GUAC_OK;
}

int guac_synth_generate(guac_ctx* ctx, int32_t contig_id, int64_t contig_len, int32_t n_samples, const int32_t* depths,
                        const uint8_t* tumor, uint64_t seed) {
  if (!ctx || !depths || !tumor) return GUAC_E_INVALID_INPUT;
  if (n_samples <= 0 || n_samples > 16 || contig_len <= 2 * synth::MARGIN) return fail(ctx, GUAC_E_INVALID_INPUT, "bad synth shape");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  ctx->have_reads = false;
  synth::Params P;
  std::memset(&P, 0, sizeof P);
  P.seed = seed;
  P.contig_len = contig_len;
  P.n_samples = n_samples;
  P.profile = (int32_t)ctx->opt_synth_profile;
  // interleaving pattern proportional to depths: round-robin, largest first (synth.py sample_pattern)
  int64_t total_depth = 0;
  int g = 0;
  for (int s = 0; s < n_samples; ++s) {
    if (depths[s] <= 0) return fail(ctx, GUAC_E_INVALID_INPUT, "depths must be positive");
    total_depth += depths[s];
    int a = depths[s], b = g;
    while (b) { int t = a % b; a = b; b = t; }
    g = a;
    P.is_tumor[s] = tumor[s] ? 1 : 0;
  }
  std::vector<int> counts(n_samples), order(n_samples);
  int plen = 0;
  for (int s = 0; s < n_samples; ++s) { counts[s] = depths[s] / g; plen += counts[s]; order[s] = s; }
  if (plen > 64) return fail(ctx, GUAC_E_UNSUPPORTED, "depth ratio pattern longer than 64");
  std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return counts[x] > counts[y]; });
  std::vector<int> remaining = counts;
  int k = 0;
  while (k < plen)
    for (int s : order)
      if (remaining[s]) { P.pattern[k++] = (uint8_t)s; --remaining[s]; }
  P.pattern_len = plen;
  const int64_t n = total_depth * contig_len / synth::READ_LEN;
  P.n_reads = n;
  // reference
  GUAC_TRY(guac_ref_reserve(ctx, contig_id, 0, contig_len));
  synth::reference_kernel<<<blocks_for(contig_len, 256), 256, 0, ctx->stream>>>(seed, contig_len, ctx->contigs[contig_id].bases);
  ctx->contigs[contig_id].ref2_dirty = true;
  // reads
  const int64_t n_bases = n * synth::READ_LEN;
  GUAC_TRY(ensure(ctx, ctx->b_start, (size_t)n * 8 + 16));
  GUAC_TRY(ensure(ctx, ctx->b_contig, (size_t)n * 4 + 16));
  GUAC_TRY(ensure(ctx, ctx->b_mapq, (size_t)n + 16));
  GUAC_TRY(ensure(ctx, ctx->b_flags, (size_t)n + 16));
  GUAC_TRY(ensure(ctx, ctx->b_sample, (size_t)n + 16));
  GUAC_TRY(ensure(ctx, ctx->b_seq_off, (size_t)(n + 1) * 8 + 16));
  GUAC_TRY(ensure(ctx, ctx->b_nops, (size_t)n + 16));
  GUAC_TRY(ensure(ctx, ctx->b_cigar_off, (size_t)(n + 1) * 4 + 16));
  GUAC_TRY(ensure(ctx, ctx->b_cigar, (size_t)n * 3 * 4 + 16));
  GUAC_TRY(ensure(ctx, ctx->b_bases2, (size_t)(n_bases / 4) + 64));
  GUAC_TRY(ensure(ctx, ctx->b_nmask, (size_t)(n_bases / 8) + 16));
  GUAC_TRY(ensure(ctx, ctx->b_qual, (size_t)n_bases + 16));
  GUAC_TRY(ensure(ctx, ctx->b_exc_pos, 16));
  GUAC_TRY(ensure(ctx, ctx->b_exc_base, 16));
  GUAC_TRY(ensure(ctx, ctx->b_elem_off, (size_t)(n + 2) * sizeof(int64_t)));
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->b_nmask.p, 0, (size_t)(n_bases / 8) + 16, ctx->stream));
  synth::read_headers_kernel<<<blocks_for(n + 1, 256), 256, 0, ctx->stream>>>(
      P, ctx->b_start.as<int64_t>(), ctx->b_contig.as<int32_t>(), contig_id, ctx->b_mapq.as<uint8_t>(), ctx->b_flags.as<uint8_t>(),
      ctx->b_sample.as<uint8_t>(), ctx->b_seq_off.as<uint64_t>(), ctx->b_nops.as<uint8_t>());
  int64_t* off64 = ctx->b_elem_off.as<int64_t>();
  GUAC_TRY((device_scan<true, uint8_t>(ctx, ctx->b_nops.as<uint8_t>(), off64, n, OpSum(), off64 + n)));
  synth::read_cigars_kernel<<<blocks_for(n + 1, 256), 256, 0, ctx->stream>>>(P, off64, ctx->b_cigar_off.as<uint32_t>(),
                                                                             ctx->b_cigar.as<uint32_t>());
  synth::read_bases_kernel<<<blocks_for(n_bases / 4, 256), 256, 0, ctx->stream>>>(P, ctx->b_bases2.as<uint8_t>(),
                                                                                  ctx->b_qual.as<uint8_t>());
  ctx->launches += 4;
  CUDA_TRY(ctx, cudaGetLastError());
  ctx->contig_spans.clear();
  ctx->contig_spans.push_back({contig_id, {0, n}});
  return finish_reads(ctx, n, n_samples, 0);
}

int guac_reads_export(guac_ctx* ctx, guac_read_batch* out) {
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
  GUAC_TRY(host_out(ctx, const_cast<int32_t**>(&out->contig_id), ctx->b_contig.p, (size_t)n));
  GUAC_TRY(host_out(ctx, const_cast<int64_t**>(&out->start), ctx->b_start.p, (size_t)n));
  GUAC_TRY(host_out(ctx, const_cast<uint32_t**>(&out->cigar_off), ctx->b_cigar_off.p, (size_t)n + 1));
  GUAC_TRY(host_out(ctx, const_cast<uint32_t**>(&out->cigar), ctx->b_cigar.p, (size_t)n_cigar));
  GUAC_TRY(host_out(ctx, const_cast<uint64_t**>(&out->seq_off), ctx->b_seq_off.p, (size_t)n + 1));
  GUAC_TRY(host_out(ctx, const_cast<uint8_t**>(&out->bases2), ctx->b_bases2.p, (size_t)((n_bases + 3) / 4)));
  GUAC_TRY(host_out(ctx, const_cast<uint8_t**>(&out->nmask), ctx->b_nmask.p, (size_t)((n_bases + 7) / 8)));
  out->n_exc = ctx->R.n_exc;
  GUAC_TRY(host_out(ctx, const_cast<uint64_t**>(&out->exc_pos), ctx->b_exc_pos.p, (size_t)ctx->R.n_exc));
  GUAC_TRY(host_out(ctx, const_cast<uint8_t**>(&out->exc_base), ctx->b_exc_base.p, (size_t)ctx->R.n_exc));
  GUAC_TRY(host_out(ctx, const_cast<uint8_t**>(&out->qual), ctx->b_qual.p, (size_t)n_bases));
  GUAC_TRY(host_out(ctx, const_cast<uint8_t**>(&out->mapq), ctx->b_mapq.p, (size_t)n));
  GUAC_TRY(host_out(ctx, const_cast<uint8_t**>(&out->flags), ctx->b_flags.p, (size_t)n));
  GUAC_TRY(host_out(ctx, const_cast<uint8_t**>(&out->sample_id), ctx->b_sample.p, (size_t)n));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return GUAC_OK;
}

int guac_ref_export(guac_ctx* ctx, int32_t contig_id, int64_t offset, int64_t len, uint8_t* dst) {
  if (!ctx || !dst) return GUAC_E_INVALID_INPUT;
  auto it = ctx->contigs.find(contig_id);
  if (it == ctx->contigs.end()) return fail(ctx, GUAC_E_INVALID_INPUT, "contig not found");
  const ContigDev& c = it->second;
  if (offset < c.offset || offset + len > c.offset + c.len) return fail(ctx, GUAC_E_INVALID_INPUT, "Illegal reference slice");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaMemcpyAsync(dst, c.bases + (offset - c.offset), (size_t)len, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return GUAC_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// guac_pileup_counts_noalloc — the counting closure for callers that must not allocate.
//
// Some hosts of this library may not call cudaMalloc at the moment they need a pileup: a plugin
// running inside an executor whose whole device memory is owned by the host framework's pool
// allocator (an allocation behind its back fails or fragments the pool).  For them this entry
// point does the work of guac_pileup_counts (dense, skip_empty = 0, one range of one contig) with
// every byte of working memory taken from the fixed array below; inputs and the output are device
// buffers the caller already owns.  The call waits on `stream` twice (its two small records travel
// through the caller's stack) and returns with the caller's current device restored.
//
// The working memory holds the per-read end loci, the running maximum used to find a tile's first
// read, the scan's partial totals, the tile descriptors, the visited flags, the one-entry reference
// table, the range record and the device error record.  It is ONE array per device, shared by
// every call: it keeps the previous call's contents (everything that must start from zero is
// cleared by the call itself), and two calls whose GPU work overlaps — different streams or
// different host threads — would race on it.  Every GPU that loads the library holds the array
// (NOALLOC_SCRATCH_BYTES, about 3.9 MiB) until the process exits.
// ------------------------------------------------------------------------------------------------
namespace {

constexpr int64_t NOALLOC_MAX_READS = 1 << 17;  // 131 072 reads per call
constexpr int64_t NOALLOC_MAX_LOCI = 1 << 19;   // 524 288 loci per call
constexpr int64_t NOALLOC_MAX_TILES = NOALLOC_MAX_LOCI / TILE_LOCI;
constexpr int64_t NOALLOC_SCAN_TILES = NOALLOC_MAX_READS / SCAN_TILE + 2;

struct NoAllocScratch {
  int64_t end[NOALLOC_MAX_READS];
  int64_t key[NOALLOC_MAX_READS];
  int64_t runmax[NOALLOC_MAX_READS];
  int64_t scan_totals[NOALLOC_SCAN_TILES];
  Tile tiles[NOALLOC_MAX_TILES];
  uint8_t visited[NOALLOC_MAX_LOCI];
  RefView ref;
  RangeDev range;
  DeviceError err;
};
constexpr size_t NOALLOC_SCRATCH_BYTES = sizeof(NoAllocScratch);

// File-scope device array.  In host code this name is only a stand-in for the symbol: its device
// address is obtained once per device with cudaGetSymbolAddress (below) and passed to the kernels.
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;
NoAllocScratch* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};
thread_local std::string g_noalloc_error;

int noalloc_fail(int status, const std::string& msg) {
  g_noalloc_error = msg;
  return status;
}

#define NA_CUDA(expr)                                                                                     \
  do {                                                                                                    \
    cudaError_t _e = (expr);                                                                              \
    if (_e != cudaSuccess) return noalloc_fail(GUAC_E_CUDA, std::string("CUDA error: ") + cudaGetErrorString(_e) + " at " #expr); \
  } while (0)

}  // namespace

extern "C" {

const char* guac_noalloc_last_error(void) { return g_noalloc_error.c_str(); }

int64_t guac_noalloc_max_reads(void) { return NOALLOC_MAX_READS; }
int64_t guac_noalloc_max_loci(void) { return NOALLOC_MAX_LOCI; }

int guac_pileup_counts_noalloc(int device, const guac_read_batch* d_batch, const uint8_t* d_ref_bases, int64_t ref_offset,
                               int64_t ref_len, int32_t contig_id, int64_t range_start, int64_t range_end,
                               guac_locus_counts* d_out_counts, void* stream_handle) {
  if (!d_batch || !d_ref_bases || !d_out_counts) return noalloc_fail(GUAC_E_INVALID_INPUT, "null argument");
  if (device < 0 || device >= NOALLOC_MAX_DEVICES) return noalloc_fail(GUAC_E_INVALID_INPUT, "device index out of range");
  const int64_t n = d_batch->n_reads, n_loci = range_end - range_start;
  const int S = d_batch->n_samples;
  if (n < 0 || S <= 0 || S > 255 || range_start < 0 || n_loci < 0 || contig_id < 0 || contig_id >= (1 << 22))
    return noalloc_fail(GUAC_E_INVALID_INPUT, "bad read batch header or loci range");
  if (ref_len < 0 || ref_offset < 0) return noalloc_fail(GUAC_E_INVALID_INPUT, "negative reference offset or length");
  if (range_end >= ((int64_t)1 << 40))  // build_tiles_kernel packs contig_id << 40 | locus
    return noalloc_fail(GUAC_E_INVALID_INPUT, "loci range beyond 2^40");
  if (n > NOALLOC_MAX_READS)
    return noalloc_fail(GUAC_E_UNSUPPORTED, "guac_pileup_counts_noalloc: " + std::to_string(n) + " reads exceed the fixed working memory (" +
                                                std::to_string(NOALLOC_MAX_READS) + " reads per call); split the batch or use guac_pileup_counts");
  if (n_loci > NOALLOC_MAX_LOCI)
    return noalloc_fail(GUAC_E_UNSUPPORTED, "guac_pileup_counts_noalloc: " + std::to_string(n_loci) + " loci exceed the fixed working memory (" +
                                                std::to_string(NOALLOC_MAX_LOCI) + " loci per call); split the range or use guac_pileup_counts");
  // the caller's current device comes back on every return path (a plugin must not move its host's device)
  struct DeviceRestore {
    int prev = -1;
    ~DeviceRestore() {
      if (prev >= 0) cudaSetDevice(prev);
    }
  } restore;
  NA_CUDA(cudaGetDevice(&restore.prev));
  if (restore.prev == device) restore.prev = -1;
  else NA_CUDA(cudaSetDevice(device));
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_handle);
  // Set-up, once per device and with that device current: resolve the symbol to its device address.
  // (The first call on a device therefore does a little extra work; make it before any stream capture.)
  if (g_noalloc_ptr[device] == nullptr) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || p == nullptr)
      return noalloc_fail(GUAC_E_CUDA, std::string("cudaGetSymbolAddress(g_noalloc_scratch) failed: ") + cudaGetErrorString(e));
    g_noalloc_ptr[device] = reinterpret_cast<NoAllocScratch*>(p);
  }
  NoAllocScratch* W = g_noalloc_ptr[device];
  if (n_loci == 0) return GUAC_OK;

  // everything that must start from zero: the caller's output (the kernel accumulates with atomics)
  // and the error record; stale ends / tiles / flags of earlier calls are overwritten before use.
  NA_CUDA(cudaMemsetAsync(d_out_counts, 0, (size_t)n_loci * S * sizeof(guac_locus_counts), stream));
  NA_CUDA(cudaMemsetAsync(&W->err, 0, sizeof(DeviceError), stream));

  ReadsView R;
  R.n_reads = n;
  R.start = d_batch->start;
  R.end = W->end;
  R.runmax_key = W->runmax;
  R.cigar_off = d_batch->cigar_off;
  R.cigar = d_batch->cigar;
  R.seq_off = d_batch->seq_off;
  R.bases2 = d_batch->bases2;
  R.nmask = d_batch->nmask;
  R.n_exc = d_batch->n_exc;
  R.exc_pos = d_batch->exc_pos;
  R.exc_base = d_batch->exc_base;
  R.qual = d_batch->qual;
  R.hdr = nullptr;
  R.mapq = d_batch->mapq;
  R.flags = d_batch->flags;
  R.sample_id = d_batch->sample_id;
  if (n > 0) {
    read_ends_kernel<<<blocks_for(n, 256), 256, 0, stream>>>(n, d_batch->start, d_batch->contig_id, d_batch->cigar_off, d_batch->cigar,
                                                            W->end, W->key);
    const int64_t scan_tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    scan_reduce_kernel<int64_t, int64_t, OpMax><<<(unsigned)scan_tiles, SCAN_THREADS, 0, stream>>>(W->key, W->scan_totals, n, OpMax());
    scan_totals_kernel<int64_t, OpMax><<<1, SCAN_THREADS, 0, stream>>>(W->scan_totals, scan_tiles, OpMax());
    scan_apply_kernel<false, int64_t, int64_t, OpMax><<<(unsigned)scan_tiles, SCAN_THREADS, 0, stream>>>(W->key, W->runmax,
                                                                                                       W->scan_totals, n, OpMax());
  }
  RefView ref;
  ref.bases = d_ref_bases;
  ref.offset = ref_offset;
  ref.len = ref_len;
  ref.bad = nullptr;
  ref.ref2 = nullptr;
  RangeDev rg;
  rg.start = range_start;
  rg.end = range_end;
  rg.slot0 = 0;
  rg.tile0 = 0;
  rg.read_clo = 0;
  rg.read_chi = n;  // the batch holds one contig
  rg.contig_id = contig_id;
  rg.ref_index = 0;
  NA_CUDA(cudaMemcpyAsync(&W->ref, &ref, sizeof ref, cudaMemcpyHostToDevice, stream));
  NA_CUDA(cudaMemcpyAsync(&W->range, &rg, sizeof rg, cudaMemcpyHostToDevice, stream));
  NA_CUDA(cudaStreamSynchronize(stream));  // `ref` and `rg` live on this stack frame
  const int64_t n_tiles = (n_loci + TILE_LOCI - 1) / TILE_LOCI;
  build_tiles_kernel<<<blocks_for(n_tiles, 256), 256, 0, stream>>>(&W->range, 1, n_tiles, TILE_LOCI, R, W->tiles);
  pileup_counts_kernel<<<(unsigned)n_tiles, CTA_THREADS, 0, stream>>>(R, &W->ref, W->tiles, S, W->visited, d_out_counts, &W->err);
  NA_CUDA(cudaGetLastError());
  DeviceError err;
  NA_CUDA(cudaMemcpyAsync(&err, &W->err, sizeof err, cudaMemcpyDeviceToHost, stream));
  NA_CUDA(cudaStreamSynchronize(stream));
  if (err.code != DERR_NONE) {
    char buf[192];
    if (err.code == DERR_REF_MISSING)
      snprintf(buf, sizeof buf, "Position %d:%lld missing from reference", err.contig_id, err.locus);
    else
      snprintf(buf, sizeof buf, "invalid CIGAR / element (device error %d) for read %lld at locus %lld", err.code, err.read_index, err.locus);
    return noalloc_fail(GUAC_E_INVALID_INPUT, buf);
  }
  return GUAC_OK;
}

}  // extern "C"

#ifdef GUAC_TK_PHASE_CLOCKS
// development builds only: reads and clears the phase clocks of ss_tile_kernel (tile_kernels.cuh)
extern "C" int guac_debug_phase_clocks(unsigned long long* out32) {
  cudaDeviceSynchronize();
  if (cudaMemcpyFromSymbol(out32, guac::g_tk_phase_clocks, 32 * sizeof(unsigned long long)) != cudaSuccess) return 1;
  unsigned long long zero[32] = {0};
  return cudaMemcpyToSymbol(guac::g_tk_phase_clocks, zero, sizeof zero) != cudaSuccess;
}
#endif
#include "assembly_entry.inl"
#include "compact_entry.inl"
#include "support_entry.inl"
#include "partition_entry.inl"
#include "host_stage.inl"

namespace {
cudaError_t configure_device_functions() {
  using TCfg = TileCfg<GUAC_TK_TL>;
  const int tile_smem = (int)sizeof(TileSmem<TCfg>);
  cudaError_t e = cudaFuncSetAttribute(ss_tile_kernel<TCfg, 0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, tile_smem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(ss_tile_kernel<TCfg, 0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, tile_smem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(ss_tile_kernel<TCfg, 1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, tile_smem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(ss_tile_kernel<TCfg, 2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, tile_smem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(ss_tile_kernel<TCfg, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, tile_smem);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(sj_snv_site_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(SJF_WARPS * sizeof(SJFWarpSmem)));
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(asm_align_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)asm_align_smem(ASM_ALIGN_MAX_S));
  return e;
}
}  // namespace
