// assembly_kernels.cuh — local assembly of the reads around an active locus (SURVEY §8 A17–A20).
//
// Restates, window-parallel, what the reference does one window at a time on one core:
//   DeBruijnGraph.apply / mergeNodes / depthFirstSearch   M/assembly/DeBruijnGraph.scala:10-464
//   AssemblyUtils.discoverHaplotypes                      M/assembly/AssemblyUtils.scala:50-114
//   ReadAlignment.scoreAlignmentPaths                     M/alignment/ReadAlignment.scala:92-195
// Every active window of a call is assembled at once: one CTA counts the window's k-mers into an
// open-addressing table in HBM (k-mers are 3-bit packed, <= 84 bases), one CTA ranks the surviving
// k-mers lexicographically and resolves their edges by table look-ups, one warp lane walks the small
// pruned graph (unipath merge + DFS are inherently sequential, but only over the few hundred nodes
// that survive minOccurrence), and one CTA per (sequence, reference) pair fills the affine-gap
// alignment matrix anti-diagonal by anti-diagonal in shared memory.
#pragma once
#include "device_model.cuh"

namespace guac {

constexpr int ASM_MAX_K = 84;       // 4 words x 21 bases x 3 bits
constexpr int ASM_MAX_PATHS = 17;   // maxPathsToScore + 1 (AssemblyUtils.scala:55,79)
constexpr int ASM_MAX_CIGAR = 64;   // run-length elements kept per path-vs-reference alignment
constexpr int ASM_ALIGN_THREADS = 128;
constexpr int ASM_ALIGN_MAX_S = 2048;

enum : int {
  DERR_ASM_BASE = 20,       // a base other than A C G T N inside an assembly window
  DERR_ASM_COLLISION = 21,  // two distinct k-mers with the same 64-bit hash (never observed; checked exactly)
  DERR_ASM_CAPACITY = 22,   // a per-window scratch bound was exceeded (tangled graph)
  DERR_ASM_CIGAR = 23       // more than ASM_MAX_CIGAR elements in a path alignment
};

struct AsmWindow {
  int64_t locus;
  int64_t read_clo, read_chi;  // reads of the contig
  int64_t read_lo, read_hi;    // span that can overlap [locus - W, locus + W]
  int32_t contig_id, ref_index;
  int32_t n_reads;             // reads of the sample overlapping the window (SlidingWindow.scala:58-85)
  int32_t n_instances;         // k-mer instances = sum over those reads of max(len - K + 1, 0)
  int64_t n_bases;             // bases of those reads
  int64_t table_off;           // slot offset of this window's hash table
  int32_t table_cap;           // power of two
  int32_t n_nodes;             // k-mers with count >= minOccurrence
  int64_t node_off;            // node offset of this window's graph arrays
  int64_t path_off;            // byte offset of the window's path strings
  int32_t path_cap;            // bytes per path
  int32_t n_paths;
  int32_t path_len[ASM_MAX_PATHS];
  int32_t max_read_len;        // longest read of the window
};

struct AsmParams {
  int sample, K, W, min_occurrence, min_mean_kmer_quality, max_paths;
  int ref_len;    // bases of a window's reference slice (2 W for the caller's windows)
  int all_reads;  // 1: the window is every read of the sample (DeBruijnGraph.discoverPathsFromReads over a given read set)
};

// ---- 3-bit packed k-mers, first base in the most significant bits so that word order is string order ----
__host__ __device__ __forceinline__ int asm_base_code(uint8_t b) {  // ASCII order A < C < G < N < T
  switch (b) {
    case 'A': return 0;
    case 'C': return 1;
    case 'G': return 2;
    case 'N': return 3;
    case 'T': return 4;
    default: return -1;
  }
}

struct Kmer {
  uint64_t w[4];
  __host__ __device__ __forceinline__ void clear() { w[0] = w[1] = w[2] = w[3] = 0; }
  __host__ __device__ __forceinline__ int get(int p) const { return (int)((w[p / 21] >> (3 * (20 - p % 21))) & 7); }
  __host__ __device__ __forceinline__ void set(int p, int c) {
    const int sh = 3 * (20 - p % 21);
    w[p / 21] = (w[p / 21] & ~((uint64_t)7 << sh)) | ((uint64_t)c << sh);
  }
  // drop the first base, append c at position K-1
  __host__ __device__ __forceinline__ Kmer shifted_left(int K, int c) const {
    Kmer o;
    const uint64_t M = ((uint64_t)1 << 63) - 1;
    o.w[0] = ((w[0] << 3) | (w[1] >> 60)) & M;
    o.w[1] = ((w[1] << 3) | (w[2] >> 60)) & M;
    o.w[2] = ((w[2] << 3) | (w[3] >> 60)) & M;
    o.w[3] = (w[3] << 3) & M;
    o.set(K - 1, c);
    return o;
  }
  // drop the last base, put c at position 0
  __host__ __device__ __forceinline__ Kmer shifted_right(int K, int c) const {
    Kmer o;
    o.w[3] = (w[3] >> 3) | ((w[2] & 7) << 60);
    o.w[2] = (w[2] >> 3) | ((w[1] & 7) << 60);
    o.w[1] = (w[1] >> 3) | ((w[0] & 7) << 60);
    o.w[0] = w[0] >> 3;
    if (K < ASM_MAX_K) o.set(K, 0);
    o.set(0, c);
    return o;
  }
  __host__ __device__ __forceinline__ bool operator==(const Kmer& b) const {
    return w[0] == b.w[0] && w[1] == b.w[1] && w[2] == b.w[2] && w[3] == b.w[3];
  }
  __host__ __device__ __forceinline__ bool operator<(const Kmer& b) const {
    if (w[0] != b.w[0]) return w[0] < b.w[0];
    if (w[1] != b.w[1]) return w[1] < b.w[1];
    if (w[2] != b.w[2]) return w[2] < b.w[2];
    return w[3] < b.w[3];
  }
  __host__ __device__ __forceinline__ uint64_t hash() const {
    uint64_t h = 0x9E3779B97F4A7C15ull;
    for (int i = 0; i < 4; ++i) {
      h ^= w[i] + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
      h ^= h >> 30;
      h *= 0xBF58476D1CE4E5B9ull;
      h ^= h >> 27;
      h *= 0x94D049BB133111EBull;
      h ^= h >> 31;
    }
    return h ? h : 1;
  }
};

__device__ __forceinline__ bool asm_in_window(const ReadsView& R, int64_t r, const AsmParams& P, int64_t locus) {
  if (__ldg(R.sample_id + r) != P.sample) return false;
  return P.all_reads || (__ldg(R.start + r) <= locus + P.W && __ldg(R.end + r) > locus - P.W);
}
__device__ __forceinline__ int asm_read_len(const ReadsView& R, int64_t r) { return (int)(__ldg(R.seq_off + r + 1) - __ldg(R.seq_off + r)); }

// One warp per window: the read span [read_lo, read_hi) and the sizes the later stages allocate from.
__global__ void asm_window_kernel(ReadsView R, AsmWindow* __restrict__ wins, int n_windows, AsmParams P) {
  const int wi = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  const int lane = threadIdx.x & 31;
  if (wi >= n_windows) return;
  AsmWindow& w = wins[wi];
  const int64_t locus = w.locus;
  int64_t a = w.read_clo, b = w.read_chi;
  while (a < b) {  // first read with start > locus + W
    int64_t m = (a + b) >> 1;
    if (__ldg(R.start + m) <= locus + P.W) a = m + 1; else b = m;
  }
  int64_t hi = a;
  const int64_t key = ((int64_t)w.contig_id << 40) | (locus - P.W);
  a = w.read_clo;
  b = hi;
  while (a < b) {  // first read whose running-max end exceeds locus - W
    int64_t m = (a + b) >> 1;
    if (__ldg(R.runmax_key + m) > key) b = m; else a = m + 1;
  }
  const int64_t lo = P.all_reads ? w.read_clo : a;
  if (P.all_reads) hi = w.read_chi;
  int n_reads = 0, n_inst = 0, max_len = 0;
  long long n_bases = 0;
  for (int64_t r = lo + lane; r < hi; r += 32) {
    if (!asm_in_window(R, r, P, locus)) continue;
    const int len = asm_read_len(R, r);
    n_reads += 1;
    n_inst += max(len - P.K + 1, 0);
    n_bases += len;
    max_len = max(max_len, len);
  }
  for (int o = 16; o > 0; o >>= 1) {
    max_len = max(max_len, __shfl_xor_sync(0xffffffffu, max_len, o));
    n_reads += __shfl_xor_sync(0xffffffffu, n_reads, o);
    n_inst += __shfl_xor_sync(0xffffffffu, n_inst, o);
    n_bases += __shfl_xor_sync(0xffffffffu, n_bases, o);
  }
  if (lane == 0) {
    w.read_lo = lo;
    w.read_hi = hi;
    w.n_reads = n_reads;
    w.n_instances = n_inst;
    w.n_bases = n_bases;
    w.max_read_len = max_len;
  }
}

// The window's reference slice [locus - W, locus + W) as ASCII (GermlineAssemblyCaller.scala:108-116), preceded by
// the base before it (an indel at the first aligned position reports its anchor base; 0 when not in the reference).
__global__ void asm_refwin_kernel(const RefView* __restrict__ refs, const AsmWindow* __restrict__ wins, int n_windows, int W,
                                  uint8_t* __restrict__ out, DeviceError* err) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int stride = 2 * W + 1;
  if (i >= (int64_t)n_windows * stride) return;
  const int wi = (int)(i / stride), j = (int)(i % stride);
  const int64_t locus = wins[wi].locus - W - 1 + j;
  const uint8_t b = refs[wins[wi].ref_index].at(locus);
  if (b == 0 && j > 0) report_error(err, DERR_REF_MISSING, wins[wi].contig_id, locus, -1);
  out[i] = b;
}

struct AsmTable {
  unsigned long long* tag;  // 64-bit hash of the k-mer, 0 = empty
  Kmer* key;                // written once by the thread that claimed the slot
  uint32_t* count;
  uint32_t* node;           // node id after ranking, 0xFFFFFFFF = pruned
};

constexpr int ASM_STAGE = 512;  // bases of a read staged in shared memory per warp (longer reads take the direct path)

// k-mer of read r at offset `at` straight from the packed batch
__device__ __forceinline__ bool asm_pack(const ReadsView& R, int64_t r, int at, int K, Kmer& k) {
  const uint64_t base0 = __ldg(R.seq_off + r) + at;
  k.clear();
  for (int p = 0; p < K; ++p) {
    const int c = asm_base_code(read_base(R, base0 + p));
    if (c < 0) return false;
    k.set(p, c);
  }
  return true;
}
// breeze's mean over the k-mer's qualities as Floats (DeBruijnGraph.scala:357-358): mu += (1 / n) * (q - mu)
__device__ __forceinline__ float asm_mean_quality(const uint8_t* q, int K, const float* inv) {
  float mu = 0.0f;
  for (int p = 0; p < K; ++p) mu = __fadd_rn(mu, __fmul_rn(inv[p], __fsub_rn((float)q[p], mu)));
  return mu;
}

__device__ __forceinline__ int asm_find(const AsmTable& T, int64_t off, int cap, const Kmer& k) {
  const unsigned long long tag = k.hash();
  int slot = (int)(tag & (unsigned long long)(cap - 1));
  for (int probe = 0; probe < cap; ++probe) {
    const unsigned long long t = T.tag[off + slot];
    if (t == tag) return slot;
    if (t == 0) return -1;
    slot = (slot + 1) & (cap - 1);
  }
  return -1;
}

// One CTA per window, one warp per read, each lane a contiguous run of the read's k-mers (rolled base by base out of
// a shared-memory copy of the read).  Pass 0 counts the k-mers that pass the quality gate; pass 1 compares every
// instance with the key stored in its slot, so a 64-bit hash collision is reported, never a silent merge.
__global__ void __launch_bounds__(256) asm_count_kernel(ReadsView R, AsmWindow* __restrict__ wins, AsmTable T, AsmParams P,
                                                        DeviceError* err) {
  AsmWindow& w = wins[blockIdx.x];
  if (w.n_instances == 0) {
    if (threadIdx.x == 0) w.n_nodes = 0;
    return;
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, n_warps = blockDim.x >> 5;
  const int64_t off = w.table_off;
  const int cap = w.table_cap, K = P.K;
  __shared__ int s_nodes;
  __shared__ float s_inv[ASM_MAX_K];
  // the read's base codes as one bit stream, 3 bits per base, first base in the most significant bits of word 0: a k-mer
  // word (21 bases) is then 63 consecutive bits of the stream, cut out with two loads and a funnel shift
  __shared__ unsigned long long s_bits[8][ASM_STAGE * 3 / 64 + 2];
  __shared__ uint8_t s_qual[8][ASM_STAGE];
  if (threadIdx.x == 0) s_nodes = 0;
  for (int p = threadIdx.x; p < K; p += blockDim.x) s_inv[p] = __fdiv_rn(1.0f, (float)(p + 1));
  __syncthreads();
  auto visit = [&](int pass, const Kmer& k, int64_t r) {
    if (pass == 0) {
      const unsigned long long tag = k.hash();
      int slot = (int)(tag & (unsigned long long)(cap - 1));
      while (true) {
        const unsigned long long old = atomicCAS(T.tag + off + slot, 0ull, tag);
        if (old == 0ull) {
          T.key[off + slot] = k;
          break;
        }
        if (old == tag) break;
        slot = (slot + 1) & (cap - 1);
      }
      // the instance that lifts a k-mer to minOccurrence counts it as a node (no scan of the table afterwards)
      if (atomicAdd(T.count + off + slot, 1u) + 1u == (uint32_t)max(P.min_occurrence, 1)) atomicAdd(&s_nodes, 1);
    } else {
      const int slot = asm_find(T, off, cap, k);
      if (!(T.key[off + slot] == k)) report_error(err, DERR_ASM_COLLISION, w.contig_id, w.locus, r);
    }
  };
  for (int pass = 0; pass < 2; ++pass) {
    for (int64_t r = w.read_lo + warp; r < w.read_hi; r += n_warps) {
      if (!asm_in_window(R, r, P, w.locus)) continue;
      const int len = asm_read_len(R, r), n_at = len - K + 1;
      if (n_at <= 0) continue;
      const uint64_t base0 = __ldg(R.seq_off + r);
      if (len <= ASM_STAGE) {
        bool bad = false, zero_q = false;
        const int n_words = (3 * len + 63) / 64 + 1;
        for (int i = lane; i < n_words; i += 32) s_bits[warp][i] = 0ull;
        __syncwarp();
        for (int i = lane; i < len; i += 32) {
          const int c = asm_base_code(read_base(R, base0 + i));
          bad |= c < 0;
          const uint8_t q = __ldg(R.qual + base0 + i);
          zero_q |= q == 0;
          s_qual[warp][i] = q;
          const int t = 3 * i, sh = t & 63;  // stream bits t .. t+2
          const unsigned long long cc = (unsigned long long)(c & 7);
          atomicOr(&s_bits[warp][t >> 6], sh <= 61 ? cc << (61 - sh) : cc >> (sh - 61));
          if (sh > 61) atomicOr(&s_bits[warp][(t >> 6) + 1], cc << (125 - sh));
        }
        bad = __any_sync(0xffffffffu, bad);
        zero_q = __any_sync(0xffffffffu, zero_q);
        __syncwarp();
        if (bad) {
          if (lane == 0) report_error(err, DERR_ASM_BASE, w.contig_id, w.locus, r);
        } else {
          // breeze's mean of the k-mer's qualities is > 0 exactly when one of them is: the running mean starts at the
          // first quality and never returns to 0 from a positive value (every step keeps at least (1 - 1/n) of it).  So a
          // threshold below 0 passes every k-mer, and the default threshold 0 every k-mer of a read without a zero quality.
          const bool all_pass = P.min_mean_kmer_quality < 0 || (P.min_mean_kmer_quality == 0 && !zero_q);
          for (int at = lane; at < n_at; at += 32) {
            if (!all_pass && !(asm_mean_quality(s_qual[warp] + at, K, s_inv) > (float)P.min_mean_kmer_quality)) continue;
            Kmer k;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int nb = min(21, K - 21 * j);  // bases of the k-mer in word j
              unsigned long long v = 0ull;
              if (nb > 0) {
                const int t0 = 3 * (at + 21 * j), sh = t0 & 63;
                v = s_bits[warp][t0 >> 6] << sh;
                if (sh) v |= s_bits[warp][(t0 >> 6) + 1] >> (64 - sh);
                v >>= 1;
                if (nb < 21) v &= ~0ull << (63 - 3 * nb);
              }
              k.w[j] = v;
            }
            visit(pass, k, r);
          }
        }
        __syncwarp();
      } else {
        for (int at = lane; at < n_at; at += 32) {
          float mu = 0.0f;
          for (int p = 0; p < K; ++p) mu = __fadd_rn(mu, __fmul_rn(s_inv[p], __fsub_rn((float)__ldg(R.qual + base0 + at + p), mu)));
          if (!(mu > (float)P.min_mean_kmer_quality)) continue;
          Kmer k;
          if (!asm_pack(R, r, at, K, k)) {
            report_error(err, DERR_ASM_BASE, w.contig_id, w.locus, r);
            continue;
          }
          visit(pass, k, r);
        }
      }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) w.n_nodes = s_nodes;  // (the pass loop ends with a barrier)
}

// Per-node graph arrays (index = node_off + node id; node ids are lexicographic ranks).
struct AsmGraph {
  Kmer* key;          // [n]
  Kmer* tmp_key;      // [n] unsorted survivors
  int32_t* tmp_slot;  // [n]
  int32_t* cgroup;    // [n] prefix group holding the children of the node (-1: none)
  int32_t* pargroup;  // [n] suffix group holding the parents of the node
  int32_t* pgroup;    // [n] prefix group the node belongs to (= smallest member)
  int32_t* sgroup;    // [n] suffix group the node belongs to
  int32_t* plist;     // [n][6] members of prefix group g, in the reference's list order
  int32_t* slist;     // [n][6]
  int32_t* pcount;    // [n]
  int32_t* scount;    // [n]
};

// One CTA per window: compact the surviving k-mers, rank them (string order = allNodes order and the order of
// the children / parents lists), and resolve edges and sibling groups by table look-ups.
__global__ void __launch_bounds__(256) asm_nodes_kernel(ReadsView R, AsmWindow* __restrict__ wins, AsmTable T, AsmGraph G, AsmParams P) {
  const AsmWindow& w = wins[blockIdx.x];
  const int n = w.n_nodes;
  if (n == 0) return;
  const int64_t off = w.table_off, no = w.node_off;
  const int cap = w.table_cap, K = P.K;
  __shared__ int s_next;
  if (threadIdx.x == 0) s_next = 0;
  __syncthreads();
  for (int s = threadIdx.x; s < cap; s += blockDim.x) {
    if (T.tag[off + s] == 0 || T.count[off + s] < (uint32_t)P.min_occurrence) continue;
    const int i = atomicAdd(&s_next, 1);
    G.tmp_key[no + i] = T.key[off + s];
    G.tmp_slot[no + i] = s;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const Kmer k = G.tmp_key[no + i];
    int rank = 0;
    for (int j = 0; j < n; ++j) rank += G.tmp_key[no + j] < k;
    G.key[no + rank] = k;
    T.node[off + G.tmp_slot[no + i]] = (uint32_t)rank;
  }
  __syncthreads();
  auto node_of = [&](const Kmer& k) -> int {
    const int s = asm_find(T, off, cap, k);
    if (s < 0) return -1;
    const uint32_t id = T.node[off + s];
    return id == 0xFFFFFFFFu ? -1 : (int)id;
  };
  for (int x = threadIdx.x; x < n; x += blockDim.x) {
    const Kmer k = G.key[no + x];
    // siblings: same first K-1 bases (vary the last) / same last K-1 bases (vary the first)
    int pg = x, sg = x, ppos = 0, spos = 0, pn = 0, sn = 0;
    for (int c = 0; c < 5; ++c) {
      Kmer a = k;
      a.set(K - 1, c);
      const int ia = (c == k.get(K - 1)) ? x : node_of(a);
      if (ia >= 0) {
        pg = min(pg, ia);
        pn += 1;
        ppos += c < k.get(K - 1);
      }
      Kmer b = k;
      b.set(0, c);
      const int ib = (c == k.get(0)) ? x : node_of(b);
      if (ib >= 0) {
        sg = min(sg, ib);
        sn += 1;
        spos += c < k.get(0);
      }
    }
    G.pgroup[no + x] = pg;
    G.sgroup[no + x] = sg;
    G.plist[(no + pg) * 6 + ppos] = x;
    G.slist[(no + sg) * 6 + spos] = x;
    if (pg == x) G.pcount[no + x] = pn;
    if (sg == x) G.scount[no + x] = sn;
  }
  __syncthreads();
  for (int x = threadIdx.x; x < n; x += blockDim.x) {
    const Kmer k = G.key[no + x];
    int cg = -1, parg = -1;
    for (int c = 0; c < 5 && cg < 0; ++c) {
      const int ch = node_of(k.shifted_left(K, c));
      if (ch >= 0) cg = G.pgroup[no + ch];
    }
    for (int c = 0; c < 5 && parg < 0; ++c) {
      const int pa = node_of(k.shifted_right(K, c));
      if (pa >= 0) parg = G.sgroup[no + pa];
    }
    G.cgroup[no + x] = cg;
    G.pargroup[no + x] = parg;
  }
}

// Sequential scratch of the graph walk; node ids >= n are merged nodes (capacity 2n + 2 per window, at 2 * node_off).
struct AsmWalk {
  int32_t* first;      // [2n+2] first / last constituent k-mer
  int32_t* last;
  int32_t* coff;       // constituents of a merged node: pool[coff .. coff + ccnt)
  int32_t* ccnt;
  uint8_t* alive;
  uint8_t* visited;    // DFS visited set
  int32_t* stamp;      // findMergeable visited set (walk id)
  int32_t* mi_node;    // [n] mergeIndex of an original k-mer: merged node, index in the merged path
  int32_t* mi_idx;
  uint8_t* in_all;     // [n] allNodes
  int32_t* pool;       // [4n+64]
  int32_t* chain;      // [4n+64] scratch for one merge
  int32_t* frontier;   // [10n+64]
  int32_t* path;       // [2n+8]
  int32_t* branch;     // [2][2n+8]
  uint8_t* paths;      // path strings, path_cap bytes each
};

struct AsmWalkCtx {
  const AsmGraph& G;
  const AsmWalk& X;
  int64_t no;   // node offset (per-node arrays of G and [n] arrays of X)
  int64_t mo;   // 2 * no + 2 * window index: offset of the [2n+2] arrays
  int64_t po;   // pool offset
  int n, K;
  __device__ int first(int v) const { return v < n ? v : X.first[mo + v]; }
  __device__ int last(int v) const { return v < n ? v : X.last[mo + v]; }
  __device__ int ccnt(int v) const { return v < n ? 1 : X.ccnt[mo + v]; }
  __device__ int cons(int v, int i) const { return v < n ? v : X.pool[po + X.coff[mo + v] + i]; }
  __device__ int slen(int v) const { return K + ccnt(v) - 1; }
  __device__ uint8_t char_at(int v, int p) const {  // p-th base of the node's string
    const int c = p < K ? G.key[no + cons(v, 0)].get(p) : G.key[no + cons(v, p - K + 1)].get(K - 1);
    return (uint8_t)("ACGNT"[c]);
  }
  __device__ const int32_t* children(int v, int& cnt) const {
    const int g = G.cgroup[no + last(v)];
    cnt = g < 0 ? 0 : G.pcount[no + g];
    return G.plist + (no + max(g, 0)) * 6;
  }
  __device__ const int32_t* parents(int v, int& cnt) const {
    const int g = G.pargroup[no + first(v)];
    cnt = g < 0 ? 0 : G.scount[no + g];
    return G.slist + (no + max(g, 0)) * 6;
  }
  __device__ void list_remove(int32_t* list, int32_t& cnt, int v) const {
    int o = 0;
    for (int i = 0; i < cnt; ++i)
      if (list[i] != v) list[o++] = list[i];
    cnt = o;
  }
  __device__ bool list_push_front(int32_t* list, int32_t& cnt, int v) const {
    if (cnt >= 6) return false;
    for (int i = cnt; i > 0; --i) list[i] = list[i - 1];
    list[0] = v;
    cnt += 1;
    return true;
  }
  __device__ void remove_node(int v) const {  // removeKmer (:58-77)
    X.alive[mo + v] = 0;
    const int pg = G.pgroup[no + first(v)], sg = G.sgroup[no + last(v)];
    list_remove(G.plist + (no + pg) * 6, G.pcount[no + pg], v);
    list_remove(G.slist + (no + sg) * 6, G.scount[no + sg], v);
  }
};

// One lane per window: mergeNodes (:91-126) and depthFirstSearch (:183-300) over node ids instead of strings.
// `win_ref` holds each window's reference slice; its first / last K bases are the source / sink (:446-447).
__global__ void asm_walk_kernel(AsmWindow* __restrict__ wins, int n_windows, AsmTable T, AsmGraph G, AsmWalk X, AsmParams P,
                                const uint8_t* __restrict__ win_ref, DeviceError* err) {
  const int wi = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  if (wi >= n_windows || (threadIdx.x & 31) != 0) return;
  AsmWindow& w = wins[wi];
  w.n_paths = 0;
  const int n = w.n_nodes, K = P.K;
  const int64_t no = w.node_off;
  const int64_t mo = 2 * no + 2 * (int64_t)wi;
  const int64_t po = 4 * no + 64 * (int64_t)wi;
  const int64_t fo = 10 * no + 64 * (int64_t)wi;
  const int64_t so = 2 * no + 8 * (int64_t)wi;  // path / branch stacks
  const int pool_cap = 4 * n + 64, front_cap = 10 * n + 64, stack_cap = 2 * n + 8, node_cap = 2 * n + 2;
  AsmWalkCtx C{G, X, no, mo, po, n, K};
  auto capacity = [&]() { report_error(err, DERR_ASM_CAPACITY, w.contig_id, w.locus, -1); };
  const uint8_t* ref = win_ref + (int64_t)wi * (P.ref_len + 1) + 1;
  Kmer source, sink;
  source.clear();
  sink.clear();
  bool ends_ok = true;
  for (int p = 0; p < K; ++p) {
    const int a = asm_base_code(ref[p]), b = asm_base_code(ref[P.ref_len - K + p]);
    if (a < 0 || b < 0) {
      ends_ok = false;
      break;
    }
    source.set(p, a);
    sink.set(p, b);
  }
  if (!ends_ok) return;  // a reference k-mer with a base no read k-mer can hold: no path can reach it
  auto node_of = [&](const Kmer& k) -> int {
    if (w.n_instances == 0) return -1;
    const int s = asm_find(T, w.table_off, w.table_cap, k);
    if (s < 0) return -1;
    const uint32_t id = T.node[w.table_off + s];
    return id == 0xFFFFFFFFu ? -1 : (int)id;
  };
  // ---- mergeNodes ----
  for (int v = 0; v < node_cap; ++v) {
    X.alive[mo + v] = v < n;
    X.visited[mo + v] = 0;
    X.stamp[mo + v] = 0;
  }
  for (int x = 0; x < n; ++x) {
    X.mi_node[no + x] = -1;
    X.in_all[no + x] = 1;
  }
  int n_total = n, pool_used = 0, walk_id = 0;
  int32_t* chain = X.chain + po;
  for (int node = 0; node < n; ++node) {
    if (!X.in_all[no + node]) continue;
    // backward walk (findMergeable searchForward = false): farthest ancestor first, `node` last
    int n_back = 0;
    {
      ++walk_id;
      int current = node;
      X.stamp[mo + current] = walk_id;
      while (true) {
        int cnt;
        const int32_t* nx = C.parents(current, cnt);
        int cand = -1, n_cand = 0;
        for (int i = 0; i < cnt; ++i)
          if (X.stamp[mo + nx[i]] != walk_id) {
            cand = nx[i];
            ++n_cand;
          }
        if (n_cand != 1) break;
        int prev_cnt;
        C.children(cand, prev_cnt);
        if (prev_cnt != 1) break;
        current = cand;
        X.stamp[mo + current] = walk_id;
        if (n_back + 2 >= pool_cap) {
          capacity();
          return;
        }
        chain[n_back++] = current;
      }
    }
    // chain so far holds the ancestors nearest-first; reverse into path order and append `node`
    for (int i = 0; i < n_back / 2; ++i) {
      const int t = chain[i];
      chain[i] = chain[n_back - 1 - i];
      chain[n_back - 1 - i] = t;
    }
    int n_full = n_back;
    chain[n_full++] = node;
    {
      ++walk_id;
      int current = node;
      X.stamp[mo + current] = walk_id;
      while (true) {
        int cnt;
        const int32_t* nx = C.children(current, cnt);
        int cand = -1, n_cand = 0;
        for (int i = 0; i < cnt; ++i)
          if (X.stamp[mo + nx[i]] != walk_id) {
            cand = nx[i];
            ++n_cand;
          }
        if (n_cand != 1) break;
        int prev_cnt;
        C.parents(cand, prev_cnt);
        if (prev_cnt != 1) break;
        current = cand;
        X.stamp[mo + current] = walk_id;
        if (n_full + 1 >= pool_cap) {
          capacity();
          return;
        }
        chain[n_full++] = current;
      }
    }
    if (n_full > 1) {
      if (n_total >= node_cap) {
        capacity();
        return;
      }
      const int M = n_total++;
      int total = 0;
      for (int i = 0; i < n_full; ++i) total += C.ccnt(chain[i]);
      if (pool_used + total > pool_cap) {
        capacity();
        return;
      }
      const int coff = pool_used;
      for (int i = 0; i < n_full; ++i) {
        const int v = chain[i];
        for (int j = 0; j < C.ccnt(v); ++j) X.pool[po + pool_used++] = C.cons(v, j);
      }
      const int first = C.first(chain[0]), last = C.last(chain[n_full - 1]);
      for (int i = 0; i < n_full; ++i) {
        const int v = chain[i];
        if (v < n) {
          X.in_all[no + v] = 0;
          X.mi_node[no + v] = M;  // mergeIndex(full(i)) = (merged, i)
          X.mi_idx[no + v] = i;
        }
        if (X.alive[mo + v]) C.remove_node(v);
      }
      X.first[mo + M] = first;
      X.last[mo + M] = last;
      X.coff[mo + M] = coff;
      X.ccnt[mo + M] = total;
      X.alive[mo + M] = 1;
      const int pg = G.pgroup[no + first], sg = G.sgroup[no + last];
      if (!C.list_push_front(G.plist + (no + pg) * 6, G.pcount[no + pg], M) ||
          !C.list_push_front(G.slist + (no + sg) * 6, G.scount[no + sg], M)) {
        capacity();
        return;
      }
    } else {
      X.in_all[no + node] = 0;
    }
  }
  // ---- depthFirstSearch(source, sink, maxPaths) ----
  const int src = node_of(source), snk = node_of(sink);
  const bool src_eq_sink = source == sink;
  const bool src_merged = src >= 0 && X.mi_node[no + src] >= 0;
  const bool sink_merged = snk >= 0 && X.mi_node[no + snk] >= 0;
  const int sink_node = sink_merged ? X.mi_node[no + snk] : -1, sink_idx = sink_merged ? X.mi_idx[no + snk] : 0;
  uint8_t* out = X.paths + w.path_off;
  auto emit_char = [&](int pi, int& len, uint8_t c) {
    if (len < w.path_cap) out[(int64_t)pi * w.path_cap + len] = c;
    ++len;
  };
  constexpr int PSEUDO = -2;  // the bottom element when it is not a node of the graph: the raw source k-mer, or
                              // the merged source node cut at the source's index (mergedNode.drop(pathIndex), :207-221)
  int pseudo_node = -1, pseudo_off = 0, pseudo_cgroup = -1;
  int32_t* frontier = X.frontier + fo;
  int32_t* path = X.path + so;
  int32_t* br_index = X.branch + 2 * so;
  int32_t* br_children = br_index + stack_cap;
  int n_front = 0, n_path = 0, n_branch = 0, n_paths = 0, num_backtracks = 0;
  if (src_merged) {
    const int M1 = X.mi_node[no + src], pidx = X.mi_idx[no + src];
    if (sink_merged && sink_node == M1 && sink_idx > pidx) {  // source and sink inside one merged node (:211-214)
      int len = 0;
      for (int p = pidx; p < sink_idx + K; ++p) emit_char(0, len, C.char_at(M1, p));
      if (len > w.path_cap) {
        capacity();
        return;
      }
      w.path_len[0] = len;
      w.n_paths = 1;
      return;
    }
    X.visited[mo + M1] = 1;
    if (pidx == 0) {
      frontier[n_front++] = M1;
    } else {
      pseudo_node = M1;
      pseudo_off = pidx;
      pseudo_cgroup = G.cgroup[no + C.last(M1)];
      frontier[n_front++] = PSEUDO;
    }
  } else if (src >= 0) {
    frontier[n_front++] = src;
  } else {
    for (int c = 0; c < 5 && pseudo_cgroup < 0; ++c) {
      const int ch = node_of(source.shifted_left(K, c));
      if (ch >= 0) pseudo_cgroup = G.pgroup[no + ch];
    }
    frontier[n_front++] = PSEUDO;
  }
  auto elem_len = [&](int v) { return v == PSEUDO ? (pseudo_node < 0 ? K : C.slen(pseudo_node) - pseudo_off) : C.slen(v); };
  auto elem_char = [&](int v, int p) -> uint8_t {
    if (v != PSEUDO) return C.char_at(v, p);
    return pseudo_node < 0 ? ref[p] : C.char_at(pseudo_node, pseudo_off + p);
  };
  while (n_front > 0 && n_paths < P.max_paths) {
    const int next = frontier[--n_front];
    if (n_path >= stack_cap) {
      capacity();
      return;
    }
    path[n_path++] = next;
    const bool found_merged_sink = sink_merged && next >= 0 && next == sink_node;
    const bool is_sink_string = next >= 0 ? (next < n && next == snk) : (elem_len(PSEUDO) == K && src_eq_sink);
    const bool found_sink = is_sink_string || found_merged_sink;
    int cnt = 0;
    const int32_t* kids;
    if (next == PSEUDO) {
      cnt = pseudo_cgroup < 0 ? 0 : G.pcount[no + pseudo_cgroup];
      kids = G.plist + (no + max(pseudo_cgroup, 0)) * 6;
    } else {
      kids = C.children(next, cnt);
    }
    int n_next = 0;
    for (int i = 0; i < cnt; ++i) n_next += !X.visited[mo + kids[i]];
    if (!found_sink && n_next > 0) {
      if (n_next > 1) {
        if (n_branch >= stack_cap) {
          capacity();
          return;
        }
        br_index[n_branch] = n_path;
        br_children[n_branch] = n_next;
        ++n_branch;
      }
      if (n_front + n_next > front_cap) {
        capacity();
        return;
      }
      for (int i = 0; i < cnt; ++i)
        if (!X.visited[mo + kids[i]]) frontier[n_front++] = kids[i];
      if (next >= 0) X.visited[mo + next] = 1;
    } else {
      if (found_sink) {  // minPathLength = 1 always holds
        int len = 0;
        for (int e = 0; e < n_path; ++e) {
          const int v = path[e];
          int el = elem_len(v);
          if (e == n_path - 1 && found_merged_sink) el = max(0, el - (C.slen(sink_node) - (sink_idx + K)));
          if (e == 0) {
            for (int p = 0; p < el; ++p) emit_char(n_paths, len, elem_char(v, p));
          } else {
            const int take = el - K + 1;  // mergeOverlappingSequences(path, kmerSize) (:384-388)
            for (int p = el - take; p < el && take > 0; ++p) emit_char(n_paths, len, elem_char(v, p));
          }
        }
        if (len > w.path_cap) {
          capacity();
          return;
        }
        w.path_len[n_paths++] = len;
      }
      if (n_branch > 0) {
        const int back_to = br_index[n_branch - 1];
        while (n_path > back_to) {
          const int v = path[--n_path];
          if (v >= 0) X.visited[mo + v] = 0;
        }
        num_backtracks += 1;
        if (num_backtracks == br_children[n_branch - 1] - 1) {
          --n_branch;
          num_backtracks = 0;
        }
      }
    }
  }
  w.n_paths = n_paths;
}

// ---- affine-gap alignment (ReadAlignment.scala) ------------------------------------------------
struct AsmAlignJob {
  int64_t seq;       // seq_kind 0: read index; 1: byte offset into `strings`
  int64_t ref;       // byte offset into `strings`
  int32_t seq_len, ref_len;
  int32_t seq_kind;
  int32_t sum_slot;  // >= 0: add -alignmentScore to sums[sum_slot] (AssemblyUtils.scala:97-101)
  int32_t trace;     // >= 0: write the run-length CIGAR of this alignment to cigars[trace]
  int32_t _pad;
  int64_t head_off;  // byte offset of the (seq_len x (ref_len + 1)) back-pointer matrix when trace >= 0
};

struct AsmPenalties {  // -log of the probabilities (ReadAlignment.scala:98-104), computed on the host
  double mismatch, open_gap, no_gap, close_gap, continue_gap;
};

struct AsmCigar {
  int32_t ref_start, ref_end, score, n_elements;
  double raw_score;
  uint8_t op[ASM_MAX_CIGAR];   // 0 '=', 1 'X', 2 'I', 3 'D'
  int32_t len[ASM_MAX_CIGAR];
};

__device__ __forceinline__ double asm_transition_penalty(const AsmPenalties& Q, int next_state, int prev, bool is_end) {  // :118-130
  const bool has_prev = prev != 255;
  const bool next_gap = next_state >= 2, prev_gap = has_prev && prev >= 2;
  const bool open_gap = !(has_prev && prev == next_state) && next_gap;
  const bool close_gap = has_prev && next_state != prev && prev_gap;
  const bool continue_gap = has_prev && prev == next_state && next_gap;
  const double third = continue_gap ? Q.continue_gap : (next_state == 1 ? __dadd_rn(Q.no_gap, Q.mismatch) : Q.no_gap);
  double t = __dadd_rn(open_gap ? Q.open_gap : 0.0, close_gap ? Q.close_gap : 0.0);
  t = __dadd_rn(t, third);
  t = __dadd_rn(t, (is_end && next_gap) ? Q.close_gap : 0.0);
  return t;
}

// One CTA per job; cell (s, r) = best alignment of sequence[s..] starting at reference index r. Anti-diagonal d = s + r
// only needs diagonals d+1 and d+2, kept in shared memory as three rotating rows indexed by s.
__global__ void __launch_bounds__(ASM_ALIGN_THREADS) asm_align_kernel(ReadsView R, const AsmAlignJob* __restrict__ jobs, AsmPenalties Q,
                                                                      const uint8_t* __restrict__ strings, uint8_t* __restrict__ heads,
                                                                      long long* __restrict__ sums, AsmCigar* __restrict__ cigars,
                                                                      DeviceError* err) {
  extern __shared__ __align__(16) unsigned char smem[];
  const AsmAlignJob job = jobs[blockIdx.x];
  const int S = job.seq_len, Rn = job.ref_len;
  double* score = reinterpret_cast<double*>(smem);              // [3][S + 1]
  int32_t* end_idx = reinterpret_cast<int32_t*>(score + 3 * (S + 1));
  uint8_t* head = reinterpret_cast<uint8_t*>(end_idx + 3 * (S + 1));
  uint8_t* seq = head + 3 * (S + 1);                            // [S]
  for (int s = threadIdx.x; s < S; s += blockDim.x)
    seq[s] = job.seq_kind == 0 ? read_base(R, __ldg(R.seq_off + job.seq) + s) : strings[job.seq + s];
  const uint8_t* refs = strings + job.ref;
  uint8_t* H = job.trace >= 0 ? heads + job.head_off : nullptr;
  __shared__ double s_best;
  __shared__ int s_best_r;
  __syncthreads();
  const int W1 = S + 1;
  for (int d = S - 1 + Rn; d >= 0; --d) {
    double* cur_score = score + (d % 3) * W1;
    const double* d1_score = score + ((d + 1) % 3) * W1;
    const double* d2_score = score + ((d + 2) % 3) * W1;
    int32_t* cur_end = end_idx + (d % 3) * W1;
    const int32_t* d1_end = end_idx + ((d + 1) % 3) * W1;
    const int32_t* d2_end = end_idx + ((d + 2) % 3) * W1;
    uint8_t* cur_head = head + (d % 3) * W1;
    const uint8_t* d1_head = head + ((d + 1) % 3) * W1;
    const uint8_t* d2_head = head + ((d + 2) % 3) * W1;
    const int s_lo = max(0, d - Rn), s_hi = min(S - 1, d);
    for (int s = s_lo + threadIdx.x; s <= s_hi; s += blockDim.x) {
      const int r = d - s;
      const bool is_end = s == 0, last_row = s + 1 == S;
      // (s+1, r) -> Insertion
      double from_score = last_row ? 0.0 : d1_score[s + 1];
      int from_head = last_row ? 255 : d1_head[s + 1];
      int from_end = last_row ? r : d1_end[s + 1];
      double best = __dadd_rn(from_score, asm_transition_penalty(Q, 2, from_head, is_end));
      int best_state = 2, best_end = from_end;
      if (r + 1 <= Rn) {
        // (s, r+1) -> Deletion
        double sc = __dadd_rn(d1_score[s], asm_transition_penalty(Q, 3, d1_head[s], is_end));
        if (sc < best) {
          best = sc;
          best_state = 3;
          best_end = d1_end[s];
        }
        // (s+1, r+1) -> Match / Mismatch
        from_score = last_row ? 0.0 : d2_score[s + 1];
        from_head = last_row ? 255 : d2_head[s + 1];
        from_end = last_row ? r + 1 : d2_end[s + 1];
        const int st = seq[s] != refs[r] ? 1 : 0;
        sc = __dadd_rn(from_score, asm_transition_penalty(Q, st, from_head, is_end));
        if (sc < best) {
          best = sc;
          best_state = st;
          best_end = from_end;
        }
      }
      cur_score[s] = best;
      cur_head[s] = (uint8_t)best_state;
      cur_end[s] = best_end;
      if (H) H[(int64_t)s * (Rn + 1) + r] = (uint8_t)best_state;
      if (s == 0) {  // (0, r) for r descending: `<=` keeps the smallest r among equal minima (:82-84 minBy)
        if (r == Rn || best <= s_best) {
          s_best = best;
          s_best_r = r | (best_end << 16);
        }
      }
    }
    __syncthreads();
  }
  if (threadIdx.x != 0) return;
  const int best_r = s_best_r & 0xFFFF, best_end = s_best_r >> 16;
  const int score_int = (int)s_best;  // score.toInt
  if (job.sum_slot >= 0) atomicAdd(reinterpret_cast<unsigned long long*>(sums + job.sum_slot), (unsigned long long)(-(long long)score_int));
  if (job.trace < 0) return;
  AsmCigar& c = cigars[job.trace];
  c.ref_start = best_r;
  c.ref_end = best_end;
  c.score = score_int;
  c.raw_score = s_best;
  int n_el = 0, s = 0, r = best_r;
  while (s < S) {
    const int st = H[(int64_t)s * (Rn + 1) + r];
    if (n_el > 0 && c.op[n_el - 1] == st) {
      c.len[n_el - 1] += 1;
    } else {
      if (n_el >= ASM_MAX_CIGAR) {
        report_error(err, DERR_ASM_CIGAR, -1, -1, -1);
        break;
      }
      c.op[n_el] = (uint8_t)st;
      c.len[n_el] = 1;
      ++n_el;
    }
    if (st == 2) ++s;
    else if (st == 3) ++r;
    else {
      ++s;
      ++r;
    }
  }
  c.n_elements = n_el;
}

// ---- window-wide evidence (GermlineAssemblyCaller.scala:162-192) --------------------------------
struct AsmEvidence {
  int32_t depth, _pad;
  double mean_mapq, median_mapq, mean_baseq, median_baseq;
};

// One warp per window. Medians come from histograms (exact); the two breeze means are running float means,
// taken in batch order (the reference's order is the window queue's heap order).
__global__ void __launch_bounds__(128) asm_evidence_kernel(ReadsView R, const AsmWindow* __restrict__ wins, const int32_t* __restrict__ which,
                                                           int n_windows, AsmParams P, AsmEvidence* __restrict__ out) {
  __shared__ uint32_t s_hist[4][2][256];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int wi = blockIdx.x * 4 + warp;  // index into `which`: only windows that produced variants need evidence
  if (wi >= n_windows) return;
  const AsmWindow w = wins[which[wi]];
  uint32_t* hm = s_hist[warp][0];
  uint32_t* hb = s_hist[warp][1];
  for (int i = lane; i < 256; i += 32) hm[i] = hb[i] = 0;
  __syncwarp();
  float mean_m = 0.0f, mean_b = 0.0f;
  long long n_m = 0, n_b = 0;
  for (int64_t r = w.read_lo; r < w.read_hi; ++r) {
    if (!asm_in_window(R, r, P, w.locus)) continue;
    const uint64_t b0 = __ldg(R.seq_off + r);
    const int len = asm_read_len(R, r);
    for (int i = lane; i < len; i += 32) atomicAdd(&hb[__ldg(R.qual + b0 + i)], 1u);
    if (lane == 0) {
      atomicAdd(&hm[__ldg(R.mapq + r)], 1u);
      n_m += 1;
      mean_m = __fadd_rn(mean_m, __fmul_rn(__fdiv_rn(1.0f, (float)n_m), __fsub_rn((float)__ldg(R.mapq + r), mean_m)));
      for (int i = 0; i < len; ++i) {
        n_b += 1;
        mean_b = __fadd_rn(mean_b, __fmul_rn(__fdiv_rn(1.0f, (float)n_b), __fsub_rn((float)__ldg(R.qual + b0 + i), mean_b)));
      }
    }
  }
  __syncwarp();
  if (lane != 0) return;
  auto median = [](const uint32_t* h, long long n) -> double {  // breeze median over Float
    if (n == 0) return 0.0;
    auto kth = [&](long long k) {
      long long acc = 0;
      for (int v = 0; v < 256; ++v) {
        acc += h[v];
        if (acc > k) return (float)v;
      }
      return 255.0f;
    };
    if (n % 2) return (double)kth((n - 1) / 2);
    return (double)__fdiv_rn(__fadd_rn(kth(n / 2 - 1), kth(n / 2)), 2.0f);
  };
  AsmEvidence e;
  e.depth = (int32_t)n_m;
  e._pad = 0;
  e.mean_mapq = mean_m;
  e.median_mapq = median(hm, n_m);
  e.mean_baseq = mean_b;
  e.median_baseq = median(hb, n_b);
  out[wi] = e;
}

}  // namespace guac
