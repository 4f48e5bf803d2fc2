// tile_kernels.cuh — streaming tile kernel for somatic-standard (the benchmarked path).
//
// One CTA owns a tile of TL consecutive loci; everything about the tile lives in shared memory.
//   A. read-parallel pass, one thread per read: every read that can overlap the tile is visited
//      ONCE.  Its coverage goes into a packed per-sample difference array (+1 at start, -1 at
//      end); its aligned segments are compared with the reference 16 bases (32 bits) at a time in
//      2-bit space, so only NON-matching elements (mismatches, insertion / deletion anchors,
//      mid-deletion loci) cost anything: each becomes an event (locus, read, kind, base, where to
//      find its quality).  The CIGAR is walked per operator, never per base.  Reads that need the
//      exact per-locus cursor (N skips, padding, non-ACGT bases, a leading insertion at locus 0)
//      are rare and are handled one warp per read by element_at().
//   B. ONE block prefix scan over two packed arrays turns the differences into Pileup.depth per
//      sample and buckets the events by locus; referenceDepth = depth - events.
//   C. one thread per locus applies findPotentialVariantAtLocus's early-outs
//      (SomaticStandardCaller.scala:164-172).  For a surviving locus the most frequent tumor
//      variant allele and an exact upper bound of the tumor log-odds come from the events alone
//      (see `prune` below); only loci whose bound reaches the threshold — or whose alleles cannot
//      be identified from the events — are appended to a list of "heavy" loci, which the generic
//      gather kernel (kernels.cuh) evaluates in full afterwards.  Qualities of matching bases
//      are therefore read only where a call is possible.
// Tiles that exceed a shared-memory capacity go to a fallback list and are re-run whole by the
// generic kernel — still on the GPU.
#pragma once
#include "kernels.cuh"

// A/B knobs (tools/build_variant.sh): resident CTAs per SM the register allocation aims for, events per locus of capacity
#ifndef GUAC_TK_MINB
#define GUAC_TK_MINB 7
#endif
#ifndef GUAC_TK_MINB_JOINT
#define GUAC_TK_MINB_JOINT 5  // MODE 2 carries the call bound's FP64 arithmetic: more registers, fewer resident CTAs
#endif
// event records kept per read by the read pass (stream_kernels.cuh); a read with more is walked by the exact cursor
#ifndef GUAC_STREAM_EV_CAP
#define GUAC_STREAM_EV_CAP 8
#endif
constexpr int STREAM_EV_CAP_C = GUAC_STREAM_EV_CAP;
// A read with more events than its slot holds gets a record of the overflow arena (ss_read_events_kernel): its count byte
// reads STREAM_EV_OVERFLOW, the first record of its slot is (arena record, number of events), and the arena record holds
// all its events, up to STREAM_OVF_CAP.  Count 255: no arena record (more events still, or the arena was full).
constexpr int STREAM_EV_OVERFLOW = 254;
constexpr int STREAM_OVF_CAP = 64;
#ifndef GUAC_TK_TL
#define GUAC_TK_TL 512  // loci per tile
#endif
#ifndef GUAC_TK_EMUL
#define GUAC_TK_EMUL 3
#endif

namespace guac {

template <int TL_>
struct TileCfg {
  static constexpr int TL = TL_;  // loci per tile
#ifndef GUAC_TK_THREADS
#define GUAC_TK_THREADS 256
#endif
  static constexpr int THREADS = GUAC_TK_THREADS;
  static constexpr int WARPS = THREADS / 32;
  static constexpr int R_MAX = 4095;     // reads per tile (12-bit read index in an event; depth < 2^16 follows)
  static constexpr int E_MAX = TL * GUAC_TK_EMUL;   // non-match events per tile
  static constexpr int SLOW_MAX = 256;
  static constexpr int EV_SIMPLE = 24;   // a locus with more events goes to the heavy path
};

// per-read derived flags, computed once per upload
enum : uint8_t {
  RF_FAST = 1,    // single M/=/X operator, ACGT only
  RF_HIQ = 2,     // every base quality >= 4, i.e. P(correct) >= 0.60 > 1/2 (licenses the LOD bound)
  RF_MEDIUM = 4,  // operators within {M,=,X,I,D,S,H}, ACGT only, CIGAR consistent with the sequence
  RF_MAPQ_OK = 8, // mapq is 0 (the joint model ignores the read) or >= 8: with RF_HIQ, P(quality) P(mapq) > 1/2 at every base
  RF_NOZEROQ = 16 // every base quality >= 1: P(correct) >= 0.2057, so (1 - p) / p <= 3.8621 (the LOD bound's term for a read
                  // that is not RF_HIQ)
};
// log(1 - f + f (1 - p) / p) <= log(1 + f * LOWQ_ODDS_M1) for every base of quality >= 1
constexpr float LOWQ_ODDS_M1 = 2.8622f;
// weight of a tumor read with a quality-0 base in the low-quality count of a locus: any such read over a locus sends it to
// the exact evaluation; a tile may hold 15 of them before its 16-bit counts would overflow (then the whole tile falls back)
constexpr int ZEROQ_WEIGHT = 4096;
static_assert(RF_MEDIUM == 4, "ss_evidence_kernel (kernels.cuh) tests this bit of the packed header");

__global__ void read_fast_flags_kernel(ReadsView R, uint8_t* __restrict__ rflags) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R.n_reads) return;
  const uint32_t c0 = R.cigar_off[r], c1 = R.cigar_off[r + 1];
  const uint64_t s0 = R.seq_off[r], s1 = R.seq_off[r + 1];
  bool acgt = s1 > s0;
  for (uint64_t i = s0 >> 3; acgt && i <= ((s1 - 1) >> 3); ++i) {
    uint8_t m = R.nmask[i];
    if (i == (s0 >> 3)) m &= (uint8_t)(0xFFu << (s0 & 7));
    if (i == ((s1 - 1) >> 3)) m &= (uint8_t)(0xFFu >> (7 - ((s1 - 1) & 7)));
    if (m) acgt = false;
  }
  bool ops_ok = c1 > c0;
  uint64_t read_len = 0;
  bool seen_ref_op = false, leading_ins_at_zero = false;
  for (uint32_t c = c0; c < c1; ++c) {
    const uint32_t w = R.cigar[c];
    const int op = w & 15;
    if (!(op == OPC_M || op == OPC_EQ || op == OPC_X || op == OPC_I || op == OPC_D || op == OPC_S || op == OPC_H)) ops_ok = false;
    if ((w >> 4) == 0) ops_ok = false;  // zero-length operators: leave to the exact cursor
    if (consumes_read(op)) read_len += w >> 4;
    if (op == OPC_I && !seen_ref_op && R.start[r] == 0) leading_ins_at_zero = true;  // PE.scala:88-89
    if (consumes_ref(op)) seen_ref_op = true;
  }
  const bool medium = acgt && ops_ok && !leading_ins_at_zero && read_len == s1 - s0 && seen_ref_op;
  const bool fast = medium && (c1 - c0 == 1);
  // every quality >= 4  <=>  no byte with (q & 0xFC) == 0; four qualities per aligned word, the bytes outside
  // [s0, s1) forced high.  Every quality >= 1  <=>  no zero byte.
  bool hiq = true, nozero = true;
  if (s1 > s0) {
    const uint32_t* qw = reinterpret_cast<const uint32_t*>(R.qual);
    for (uint64_t w = s0 >> 2; w <= ((s1 - 1) >> 2); ++w) {
      uint32_t v = __ldg(qw + w);
      if (w == (s0 >> 2)) v |= (1u << (8 * (s0 & 3))) - 1;
      if (w == ((s1 - 1) >> 2) && ((s1 & 3) != 0)) v |= ~((1u << (8 * (s1 & 3))) - 1);
      const uint32_t t = v & 0xFCFCFCFCu;
      if ((t - 0x01010101u) & ~t & 0x80808080u) hiq = false;
      if ((v - 0x01010101u) & ~v & 0x80808080u) nozero = false;
    }
  }
  const int mapq = R.mapq[r];
  rflags[r] = (fast ? RF_FAST : 0) | (hiq ? RF_HIQ : 0) | (medium ? RF_MEDIUM : 0) | ((mapq == 0 || mapq >= 8) ? RF_MAPQ_OK : 0) |
              (nozero ? RF_NOZEROQ : 0);
}


// 32 bits starting at bit `bit` of a little-endian word stream
// One 16-byte header per read for the tile kernels: x = low 32 bits of the start, y = reference span (16 bits, 0xFFFF =
// longer: look it up) | sample << 16 | rflags << 24, (z, w) = sequence offset.  One LDG.128 instead of five loads.
__global__ void read_headers_pack_kernel(ReadsView R, const uint8_t* __restrict__ rflags, uint4* __restrict__ hdr) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R.n_reads) return;
  const int64_t st = R.start[r], span = R.end[r] - st;
  const uint64_t s0 = R.seq_off[r];
  uint4 h;
  h.x = (uint32_t)st;
  h.y = (uint32_t)min((long long)span, 0xFFFFll) | ((uint32_t)R.sample_id[r] << 16) | ((uint32_t)rflags[r] << 24);
  h.z = (uint32_t)s0;
  h.w = (uint32_t)(s0 >> 32);
  hdr[r] = h;
}

__device__ __forceinline__ uint32_t load_bits32_g(const uint8_t* __restrict__ bytes, uint64_t bit) {
  const uint32_t* w = reinterpret_cast<const uint32_t*>(bytes) + (bit >> 5);
  return __funnelshift_r(__ldg(w), __ldg(w + 1), (int)(bit & 31));
}
__device__ __forceinline__ uint32_t load_bits32_s(const uint32_t* words, int bit) {
  return __funnelshift_r(words[bit >> 5], words[(bit >> 5) + 1], bit & 31);
}


// event word: [0,11) locus in tile | [11,23) read in tile | 23 sample slot | [24,27) kind | [27,30) code
//   code: Mismatch -> base code (4 = not ACGT); other kinds -> 4 = allele identified by the detail byte,
//         5 = allele must be resolved by the exact path
// second word: Mismatch of an aligned segment -> index of the base's quality relative to the tile's first
//   read (fetched only if the event matters); otherwise bit 31 | detail << 8 | qualityScore
//   (detail: Deletion = deleted length, Insertion = 8-bit hash of the inserted bases).
__device__ __forceinline__ uint32_t pack_event(int l, int rl, int slot, int kind, int code) {
  return (uint32_t)l | ((uint32_t)rl << 11) | ((uint32_t)slot << 23) | ((uint32_t)kind << 24) | ((uint32_t)code << 27);
}
__device__ __forceinline__ int base_code(uint8_t b) { return b == 'A' ? 0 : b == 'C' ? 1 : b == 'G' ? 2 : b == 'T' ? 3 : 4; }

template <class CFG>
struct TileSmem {
  uint32_t ref2[((CFG::TL + CFG::THREADS - 1) / CFG::THREADS) * (CFG::THREADS / 16) + 2];  // whole passes of the loader + read-ahead
  int cov[CFG::TL + 2];     // packed differences: normal | tumor << 16  -> depths after the scan (MODE 2: samples 0 | 1 << 16)
  int aux[CFG::TL + 2];     // low-quality tumor read differences | event count << 16 -> lowq | event end offset
  int nonref[CFG::TL + 1];  // non-Match elements: normal | tumor << 16 (MODE 2 also keeps the locus behind the tile)
  int cursor[CFG::TL + 1];
  uint2 ev[CFG::E_MAX];
  uint16_t ev_order[CFG::E_MAX];  // event indices bucketed by locus (2 B per event instead of a second copy)
  uint16_t slow[CFG::SLOW_MAX];
  int warp_tot[2][CFG::WARPS];
  int n_events, n_slow, overflow, ref_has_mask, n_cand, n_zeroq;
  int warp_tot2[CFG::WARPS];
  int n_scan;  // MODE 2: loci whose events are scanned in a dense second loop
  // MODE 2 only, and last on purpose: the other modes are launched with offsetof(TileSmem, cov2) bytes of shared memory
  int cov2[CFG::TL + 2];    // with at most four samples: depths of samples 2 | 3 << 16
};

struct TileArgs {
  const uint8_t* rflags;
  const uint4* hdr;               // packed per-read headers (read_headers_pack_kernel)
  const RefView* refs;
  const Tile* tiles;
  uint8_t* visited;
  guac_locus_summary* summary;
  Tile* heavy;              // heavy loci as single-locus tiles
  int* n_heavy;
  int max_heavy;
  int* fallback_tiles;
  int* n_fallback;
  int force_fallback_mod;
  unsigned long long* n_visited;  // loci with at least one read in either sample
  DeviceError* err;
  const guac_range* force;        // MODE 2: force-call ranges (allele loci); a site one locus before is always a candidate
  int n_force;
  // STREAM variant (stream_kernels.cuh): the reads were compared with the reference by a kernel of their own
  const uint2* ev_by_read;        // [n_reads][STREAM_EV_CAP] read-event records
  const uint8_t* ev_count;        // [n_reads] (more than the cap: the read is walked by the exact cursor)
  const uint2* ev_overflow;       // [records][STREAM_OVF_CAP] (see STREAM_EV_OVERFLOW)
  // MODE 2 with at most four samples: events carry their sample id and candidate loci are classified — `simple` lists
  // the sites whose alternate alleles can only be single-base substitutions (sj_snv_site_kernel), `heavy` the others
  int sample_bits;
  Tile* simple;
  int* n_simple;
  int max_simple;
  // ... and simple sites that certainly hold no call are dropped in the tile (see `prune` in MODE 2 below)
  int prune;
  unsigned kinds;                 // GUAC_SJ_* of sample s in bits [2 s, 2 s + 2)
  double het_prior, hom_prior, somatic_prior, vaf_floor;
  const double2* log_table;       // (log10 p, log10 (1 - p)) by (mapq << 8 | quality), sj_log_table_kernel
};

// MODE 2, at most four samples: the discovery samples' events of one locus in one pass.  hist[s] holds, 16 bits per
// standard base, how many reads of sample s have that substitution there.  Returns the largest number of equal
// allele-starting elements in one sample (indel anchors of a sample counted together: an upper bound);
// `one_base_alleles`: no sample has min_reads indel anchors there — every allele that can start there is a substitution
// (the call bound applies); `simple`: moreover no read of any sample has an insertion or a non-standard base there —
// every subsequence is one standard base (the fast site kernel applies).  An allele needs min_reads equal elements in ONE discovery
// sample (jointcaller/AlleleAtLocus.scala:99-124); inside a deletion or an N skip the read's element one locus before
// is a variant or Clipped, so those events start nothing.
template <class SM>
__device__ inline int sj_scan_events(const SM& S, int l1, int min_reads, unsigned long long (&hist)[4], bool& simple, bool& one_base_alleles) {
  const int nr1 = S.nonref[l1], e1 = S.aux[l1] >> 16, e0 = e1 - ((nr1 & 0xFFFF) + (nr1 >> 16));
  hist[0] = hist[1] = hist[2] = hist[3] = 0ull;
  unsigned indel = 0;  // 8 bits per sample (saturating)
  simple = one_base_alleles = true;
  for (int i = e0; i < e1; ++i) {
    const uint2 ev = S.ev[S.ev_order[i]];
    const int kind = (ev.x >> 24) & 7, code = (ev.x >> 27) & 7, smp = (int)(ev.x >> 30);
    if (kind == GUAC_ALN_INSERTION || (kind == GUAC_ALN_MISMATCH && code >= 4)) simple = false;
    if (kind == GUAC_ALN_MID_DELETION || kind == GUAC_ALN_CLIPPED) continue;
    if (((ev.x >> 23) & 1) != 1) continue;  // not a discovery sample
    if (kind == GUAC_ALN_MISMATCH && code < 4) {
      const unsigned long long f = (hist[smp] >> (16 * code)) & 0xFFFFull;
      if (f < 0xFFFFull) hist[smp] += 1ull << (16 * code);
    } else if (((indel >> (8 * smp)) & 0xFFu) < 0xFFu) {
      indel += 1u << (8 * smp);
    }
  }
  int best = 0;
#pragma unroll
  for (int s = 0; s < 4; ++s) {
#pragma unroll
    for (int c = 0; c < 4; ++c) best = max(best, (int)((hist[s] >> (16 * c)) & 0xFFFFull));
    const int n_indel = (int)((indel >> (8 * s)) & 0xFFu);
    if (n_indel >= min_reads) simple = one_base_alleles = false;
    best = max(best, n_indel);
  }
  return best;
}

// The same answer for min_reads = 2 (the reference's default, AlleleAtLocus.scala:99) without the counters: an allele
// can start at the locus iff two discovery events there agree in (sample, class), the classes being the four standard
// substitutions and "indel anchor or non-standard base".  One bit per (sample, class) seen so far.
constexpr int SJ_SCAN_SERIAL_MAX = 6;       // events a single lane scans by itself; loci with more go to the warp pass
constexpr int SJ_PRUNE_MAX_ALT_READS = 8;  // with this many equal alternate reads in a sample no bound is attempted
#ifndef GUAC_SJ_PAIRSCAN
#define GUAC_SJ_PAIRSCAN 1
#endif
template <class SM>
__device__ inline bool sj_scan_pairs(const SM& S, int l1, bool& simple, bool& one_base_alleles) {
  const int nr1 = S.nonref[l1], e1 = S.aux[l1] >> 16, e0 = e1 - ((nr1 & 0xFFFF) + (nr1 >> 16));
  unsigned seen = 0, twice = 0;
  bool plain = true;
  for (int i = e0; i < e1; ++i) {
    const uint32_t x = S.ev[S.ev_order[i]].x;
    const unsigned kind = (x >> 24) & 7, code = (x >> 27) & 7;
    const bool sub = kind == GUAC_ALN_MISMATCH && code < 4;
    plain = plain && !(kind == GUAC_ALN_INSERTION || (kind == GUAC_ALN_MISMATCH && code >= 4));
    if (kind == GUAC_ALN_MID_DELETION || kind == GUAC_ALN_CLIPPED || !((x >> 23) & 1)) continue;
    const unsigned bit = 1u << (((x >> 30) << 3) + (sub ? code : 4u));
    twice |= seen & bit;
    seen |= bit;
  }
  const bool indel_twice = (twice & 0x10101010u) != 0;
  simple = plain && !indel_twice;
  one_base_alleles = !indel_twice;
  return twice != 0;
}

// sj_scan_events by a whole warp: lane k takes the events k, k + 32, ... of the locus and the counters are added across
// the lanes (a tile holds fewer than 2^16 events, so no 16-bit field overflows).  Every lane returns the same values.
template <class SM>
__device__ inline int sj_scan_events_warp(const SM& S, int l1, int min_reads, unsigned long long (&hist)[4], bool& simple, bool& one_base_alleles) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int nr1 = S.nonref[l1], e1 = S.aux[l1] >> 16, e0 = e1 - ((nr1 & 0xFFFF) + (nr1 >> 16));
  unsigned long long h0 = 0, h1 = 0, h2 = 0, h3 = 0, indel = 0;  // indel: 16 bits per sample
  bool plain = true;
  for (int i = e0 + lane; i < e1; i += 32) {
    const uint32_t x = S.ev[S.ev_order[i]].x;
    const unsigned kind = (x >> 24) & 7, code = (x >> 27) & 7, smp = x >> 30;
    plain = plain && !(kind == GUAC_ALN_INSERTION || (kind == GUAC_ALN_MISMATCH && code >= 4));
    if (kind == GUAC_ALN_MID_DELETION || kind == GUAC_ALN_CLIPPED || !((x >> 23) & 1)) continue;
    if (kind == GUAC_ALN_MISMATCH && code < 4) {
      const unsigned long long one = 1ull << (16 * code);
      h0 += smp == 0 ? one : 0ull;
      h1 += smp == 1 ? one : 0ull;
      h2 += smp == 2 ? one : 0ull;
      h3 += smp == 3 ? one : 0ull;
    } else {
      indel += 1ull << (16 * smp);
    }
  }
  if (e1 - e0 > 1) {  // (uniform across the warp)
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
      h0 += __shfl_xor_sync(FULL, h0, d);
      h1 += __shfl_xor_sync(FULL, h1, d);
      h2 += __shfl_xor_sync(FULL, h2, d);
      h3 += __shfl_xor_sync(FULL, h3, d);
      indel += __shfl_xor_sync(FULL, indel, d);
    }
  } else {  // at most one event, in lane 0
    h0 = __shfl_sync(FULL, h0, 0);
    h1 = __shfl_sync(FULL, h1, 0);
    h2 = __shfl_sync(FULL, h2, 0);
    h3 = __shfl_sync(FULL, h3, 0);
    indel = __shfl_sync(FULL, indel, 0);
  }
  plain = __all_sync(FULL, plain);
  hist[0] = h0;
  hist[1] = h1;
  hist[2] = h2;
  hist[3] = h3;
  simple = plain;
  one_base_alleles = true;
  int best = 0;
#pragma unroll
  for (int s = 0; s < 4; ++s) {
#pragma unroll
    for (int c = 0; c < 4; ++c) best = max(best, (int)((hist[s] >> (16 * c)) & 0xFFFFull));
    const int n_indel = (int)((indel >> (16 * s)) & 0xFFFFull);
    if (n_indel >= min_reads) simple = one_base_alleles = false;
    best = max(best, n_indel);
  }
  return best;
}

// MODE 0: somatic-standard (two samples).  MODE 1: germline toy caller's dense half — one sample (the "tumor"
// slot; normal_sample = -1), a locus is heavy when depth - referenceDepth >= P.min_alt_reads
// (commands/GermlineAssemblyCaller.scala:135-138).  MODE 2: somatic-joint prefilter — every sample below P.n_samples
// counts for the visit, the samples of P.sample_mask (slot 1) for the events; pileup locus x is heavy when at least
// P.min_alt_reads of their reads have a non-Match element at x + 1 (an alternate allele can only start there,
// jointcaller/pileup_summarization/ReadSubsequence.scala:106-135), when x + 1 is force-called, or when x + 1 lies
// in the next tile.
// Development aid (-DGUAC_TK_PHASE_CLOCKS): SM clocks spent between the phase boundaries of ss_tile_kernel, summed over
// the CTAs by their thread 0; read back with guac_debug_phase_clocks.  Off in the product build.
#ifdef GUAC_TK_PHASE_CLOCKS
__device__ unsigned long long g_tk_phase_clocks[32];
#define TK_PHASE(k)                                                                  \
  do {                                                                               \
    if (threadIdx.x == 0) {                                                          \
      const long long now_ = clock64();                                              \
      atomicAdd(&g_tk_phase_clocks[k], (unsigned long long)(now_ - phase_t_));       \
      phase_t_ = now_;                                                               \
    }                                                                                \
  } while (0)
#else
#define TK_PHASE(k) do { } while (0)
#endif

template <class CFG, int MODE, bool STREAM = false>
__global__ void __launch_bounds__(CFG::THREADS, (MODE == 2 ? GUAC_TK_MINB_JOINT : GUAC_TK_MINB))
ss_tile_kernel(ReadsView R, SSParamsDev P, TileArgs A) {
#ifdef GUAC_TK_PHASE_CLOCKS
  long long phase_t_ = clock64();
#endif
  extern __shared__ __align__(16) unsigned char smem_raw[];
  TileSmem<CFG>& S = *reinterpret_cast<TileSmem<CFG>*>(smem_raw);
  constexpr int TL = CFG::TL;
  const unsigned FULL = 0xffffffffu;
  const Tile tile = A.tiles[blockIdx.x];
  const RefView ref = A.refs[tile.ref_index];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n_loci = tile.n_loci;
  const int64_t t0 = tile.locus0;
  const int64_t n_reads_tile = tile.read_hi - tile.read_lo;
  // loci whose events are kept: the stream form of MODE 2 also keeps the first locus behind the tile, where the alleles
  // of the tile's last site start (every read that has an element there and at the last site is among the tile's reads)
  const int n_ev_loci = (MODE == 2 && STREAM) ? n_loci + 1 : n_loci;

  bool too_big = n_reads_tile > CFG::R_MAX;
  if (A.force_fallback_mod > 0 && blockIdx.x % A.force_fallback_mod == 0) too_big = true;  // test hook
  const uint64_t seq_base = tile.seq_base;  // filled in by build_tiles_kernel
  if (!too_big && n_reads_tile > 0) too_big = (tile.seq_end - seq_base) >= (1ull << 31);
  if (too_big) {  // uniform across the CTA
    if (tid == 0) A.fallback_tiles[atomicAdd(A.n_fallback, 1)] = blockIdx.x;
    return;
  }

  // The first header of every thread (and, STREAM, its event count) depends on the tile descriptor alone: the loads go
  // out here, ahead of the set-up and its barriers, instead of starting a new round trip behind them.
#ifndef GUAC_TK_EARLY
#define GUAC_TK_EARLY 1
#endif
  const uint4* hdrs = A.hdr + tile.read_lo;
  uint4 next_hdr = make_uint4(0, 0, 0, 0);
  int next_cnt = 0;  // STREAM: the read's event count travels with its header
#if GUAC_TK_EARLY
  if (tid < (int)n_reads_tile) {
    next_hdr = __ldg(hdrs + tid);
    if (STREAM) next_cnt = (int)__ldg(A.ev_count + tile.read_lo + tid);
  }
#endif

  // ---- init ------------------------------------------------------------------------------
  for (int i = tid; i < TL + 2; i += CFG::THREADS) {
    S.cov[i] = S.aux[i] = 0;
    if (MODE == 2) S.cov2[i] = 0;
  }
  for (int i = tid; i < TL + 1; i += CFG::THREADS) S.nonref[i] = S.cursor[i] = 0;
  if (tid == 0) S.n_events = S.n_slow = S.overflow = S.ref_has_mask = S.n_cand = S.n_zeroq = S.n_scan = 0;
  if (tid < 4) {  // padding words behind the tile
    S.ref2[TL / 16 + (tid & 1)] = 0;
  }
  __syncthreads();
  // reference tile -> 2-bit words (16 loci per word) and a not-ACGT mask (32 loci per word): one
  // coalesced byte per thread, packed with warp-wide OR reductions
  // the part of the tile the reference window holds, in tile coordinates (the rest reads as missing)
  const long long rel0 = (long long)t0 - (long long)ref.offset;
  const int l_min = (int)min(max(-rel0, 0ll), (long long)n_loci), l_max = (int)max(min((long long)ref.len - rel0, (long long)n_loci), 0ll);
  const uint8_t* ref_tile = ref.bases + rel0;
  if (STREAM) {
    // the read pass already compared against the reference; here only "is every base of the tile A C G T" matters, and
    // build_tiles_kernel answered that from the window's one-bit-per-16-loci map (Tile::_pad)
    if (tile._pad) {  // rare, uniform across the CTA: generic kernel
      if (tid == 0) A.fallback_tiles[atomicAdd(A.n_fallback, 1)] = blockIdx.x;
      return;
    }
  }
  for (int base = 0; !STREAM && base < TL; base += CFG::THREADS) {
    const int l = base + tid;
    const uint8_t b = (l >= l_min && l < l_max) ? __ldg(ref_tile + l) : (l < n_loci ? (uint8_t)0 : (uint8_t)'A');
    const uint32_t d = (uint32_t)b - 65u;  // A C G T are 65 + {0, 2, 6, 19}; their 2-bit code is ((b >> 1) ^ (b >> 2)) & 3
    const bool acgt = d < 20u && ((0x80045u >> d) & 1u);
    const int c = acgt ? (int)(((b >> 1) ^ (b >> 2)) & 3u) : 4;
    const uint32_t bits = c < 4 ? (uint32_t)c << (2 * (lane & 15)) : 0u;
    const uint32_t lo = __reduce_or_sync(FULL, lane < 16 ? bits : 0u);
    const uint32_t hi = __reduce_or_sync(FULL, lane >= 16 ? bits : 0u);
    const uint32_t mask = __ballot_sync(FULL, c == 4);
    if (lane == 0) {
      S.ref2[(base >> 4) + 2 * warp] = lo;
      S.ref2[(base >> 4) + 2 * warp + 1] = hi;
      if (mask) S.ref_has_mask = 1;
    }
  }
  if (!STREAM) {
    __syncthreads();
    if (S.ref_has_mask) {  // a reference base that is not A C G T (or is missing) in the tile: rare, generic kernel
      if (tid == 0) A.fallback_tiles[atomicAdd(A.n_fallback, 1)] = blockIdx.x;
      return;
    }
  }
  uint32_t ev_sample = 0;  // MODE 2: the current read's sample id, bits 30..31 of its event words
  auto push_event = [&](int l, int rl, int slot, int kind, int code, uint32_t second) {
    atomicAdd(&S.nonref[l], slot ? 65536 : 1);
    const int e = atomicAdd(&S.n_events, 1);
    if (e < CFG::E_MAX) S.ev[e] = make_uint2(pack_event(l, rl, slot, kind, code) | ev_sample, second); else S.overflow = 1;
  };
  auto direct = [](int q, int detail) { return 0x80000000u | ((uint32_t)(detail & 0xFF) << 8) | (uint32_t)(q & 0xFF); };

  // ---- A. read-parallel pass ---------------------------------------------------------------
  TK_PHASE(0);
  // The five header loads of a read are independent: issue them together, one read ahead of the one being processed,
  // so the compare work of read i hides the latency of read i + 256's header.
  const uint32_t t0_lo = (uint32_t)t0;
#if !GUAC_TK_EARLY
  if (tid < (int)n_reads_tile) {
    next_hdr = __ldg(hdrs + tid);
    if (STREAM) next_cnt = (int)__ldg(A.ev_count + tile.read_lo + tid);
  }
#endif
  // one record of the read pass -> the tile's events (the part of it inside the tile)
  auto take_event = [&](const uint2 ev, int st, int rl, int slot, uint32_t s0_rel, int64_t r) {
    const int l = st + (int)(ev.x & 0xFFFF);
    const int kind = (int)((ev.x >> 16) & 7), code = (int)((ev.x >> 19) & 7);
    if (kind == GUAC_ALN_MID_DELETION) {  // a run of deleted loci
      const int run = (int)((ev.y >> 16) & 0x7FFF);
      for (int ll = max(l, 0); ll < min(l + run, n_ev_loci); ++ll) push_event(ll, rl, slot, kind, code, ev.y & 0x8000FFFFu);
      return;
    }
    if (l < 0 || l >= n_ev_loci) return;
    uint32_t second = ev.y;
    if (kind == GUAC_ALN_MISMATCH) {
      second = s0_rel + ev.y;  // quality index relative to the tile's first base
    } else if (kind == GUAC_ALN_DELETION && code == 4) {  // the deleted reference bases must exist (PE.scala:95)
      const int nlen = (int)((ev.y >> 8) & 0xFF);
      // (every locus of this tile has an A C G T reference base, or the tile would not be here: only the part of the
      // deletion that runs past the tile's end is looked up)
      for (int k = max(1, n_loci - l); k <= nlen; ++k)
        if (ref.at(t0 + l + k) == 0) report_error(A.err, DERR_REF_MISSING, tile.contig_id, t0 + l + k, r);
    }
    push_event(l, rl, slot, kind, code, second);
  };
  for (int rl = tid; rl < (int)n_reads_tile; rl += CFG::THREADS) {
    const int64_t r = tile.read_lo + rl;
    const uint4 hdr = next_hdr;
    const int cnt = next_cnt;
    if (rl + CFG::THREADS < (int)n_reads_tile) {
      next_hdr = __ldg(hdrs + rl + CFG::THREADS);
      if (STREAM) next_cnt = (int)__ldg(A.ev_count + r + CFG::THREADS);
    }
    const int smp = (int)((hdr.y >> 16) & 0xFF);
    const uint8_t rf = (uint8_t)(hdr.y >> 24);
    if (MODE == 2 ? smp >= P.n_samples : (smp != P.normal_sample && smp != P.tumor_sample)) continue;
    const int slot = MODE == 2 ? (int)((P.sample_mask >> smp) & 1u) : (smp == P.tumor_sample ? 1 : 0);
    if (MODE == 2) ev_sample = A.sample_bits ? (uint32_t)(smp & 3) << 30 : 0u;
    // the reads of a tile start within 2^31 loci of it, so 32-bit wrap-around arithmetic gives the exact offset
    const int st = (int)(hdr.x - t0_lo);
    int en = st + (int)(hdr.y & 0xFFFF);
    if ((hdr.y & 0xFFFF) == 0xFFFF) en = (int)min((long long)(__ldg(R.end + r) - t0), (1ll << 30));  // a span beyond 16 bits
    const int a = max(st, 0), b = min(en, n_loci);  // overlap [a, b) in tile coordinates
    if (a >= b) continue;
    if (MODE == 2 && A.sample_bits) {
      // per-sample depths, kept one locus behind the tile like the events; reads that can have P(correct) <= 1/2 somewhere
      const int bc = min(en, n_ev_loci);
      int* arr = (smp & 2) ? S.cov2 : S.cov;
      const int one = (smp & 1) ? 65536 : 1;
      atomicAdd(&arr[a], one);
      atomicAdd(&arr[bc], -one);
      if ((rf & (RF_HIQ | RF_MAPQ_OK)) != (RF_HIQ | RF_MAPQ_OK)) {
        atomicAdd(&S.aux[a], 1);
        atomicAdd(&S.aux[bc], -1);
      }
    } else {
      const int one = slot ? 65536 : 1;
      atomicAdd(&S.cov[a], one);
      atomicAdd(&S.cov[b], -one);
      if (slot == 1 && !(rf & RF_HIQ)) {  // tumor reads that may have a base below quality 4 over each locus
        const int w = (rf & RF_NOZEROQ) ? 1 : ZEROQ_WEIGHT;
        atomicAdd(&S.aux[a], w);
        atomicAdd(&S.aux[b], -w);
        if (w != 1 && atomicAdd(&S.n_zeroq, 1) >= 15) S.overflow = 1;
      }
    }
    if (!(rf & RF_MEDIUM)) {
      const int i = atomicAdd(&S.n_slow, 1);
      if (i < CFG::SLOW_MAX) S.slow[i] = (uint16_t)rl; else S.overflow = 1;
      continue;
    }
    const uint64_t s0 = (uint64_t)hdr.z | ((uint64_t)hdr.w << 32);
    const uint32_t s0_rel = (uint32_t)(s0 - seq_base);
    if (STREAM) {  // the read was compared by ss_read_events_kernel: take the events that fall inside the tile
      if (cnt == STREAM_EV_OVERFLOW) {  // its events are in the overflow arena: copied after this loop, a warp per such read
        const int i = atomicAdd(&S.n_slow, 1);
        if (i < CFG::SLOW_MAX) S.slow[i] = (uint16_t)(rl | 0x8000); else S.overflow = 1;
        continue;
      }
      if (cnt > STREAM_EV_CAP_C && (rf & RF_FAST) && ref.ref2 != nullptr) {
        // More mismatches than the read's slot holds (a mis-mapped or very noisy read), one aligned operator: the compare of
        // the read pass again, here, for the part of the read inside the tile — against the 2-bit reference in HBM / L2
        const long long ref_rel = (long long)t0 + st - (long long)ref.offset;  // reference-window index of the read's first locus
        const int ca = (int)max(max((long long)a, (long long)st - ref_rel), (long long)st);
        const int cb = (int)min((long long)min(en, n_ev_loci), (long long)st + ((long long)ref.len - ref_rel));
        if (ca < cb) {
          const uint64_t rbit = 2 * (s0 + (uint64_t)(ca - st));
          const uint32_t* rw = reinterpret_cast<const uint32_t*>(R.bases2) + (rbit >> 5);
          const uint64_t fbit = 2 * (uint64_t)(ref_rel + (ca - st));
          const uint32_t* fw = ref.ref2 + (fbit >> 5);
          const int rsh = (int)(rbit & 31), fsh = (int)(fbit & 31);
          uint32_t rlo = __ldg(rw), flo = __ldg(fw);
          for (int l = ca; l < cb; l += 16) {
            const uint32_t rhi = __ldg(++rw), fhi = __ldg(++fw);
            const uint32_t rd = __funnelshift_r(rlo, rhi, rsh);
            const uint32_t x = rd ^ __funnelshift_r(flo, fhi, fsh);
            rlo = rhi;
            flo = fhi;
            uint32_t m = (x | (x >> 1)) & 0x55555555u;
            if (cb - l < 16) m &= (1u << (2 * (cb - l))) - 1;
            while (m) {
              const int bit = __ffs(m) - 1;
              m &= m - 1;
              const int ll = l + (bit >> 1);
              push_event(ll, rl, slot, GUAC_ALN_MISMATCH, (rd >> bit) & 3, s0_rel + (uint32_t)(ll - st));
            }
          }
        }
        continue;
      }
      if (cnt > STREAM_EV_CAP_C) {  // ... of a read with several operators: the exact cursor walks it (rare)
        const int i = atomicAdd(&S.n_slow, 1);
        if (i < CFG::SLOW_MAX) S.slow[i] = (uint16_t)rl; else S.overflow = 1;
        continue;
      }
      const uint2* evs = A.ev_by_read + r * STREAM_EV_CAP_C;
      for (int j = 0; j < cnt; ++j) take_event(__ldg(evs + j), st, rl, slot, s0_rel, r);
      continue;
    }
    // aligned segment: loci [lb, le) (tile coordinates, unclipped) hold read positions rp0..; 16 bases per step
    auto compare_segment = [&](int lb, int le, int rp0) {
      const int ca = max(lb, 0), cb = min(le, n_loci);
      if (ca >= cb) return;
      // both streams advance by exactly one 32-bit word per step, so the shifts are loop invariant and
      // each step loads one new word per stream
      const uint64_t rbit = 2 * (s0 + (uint64_t)(rp0 + (ca - lb)));
      const uint32_t* rw = reinterpret_cast<const uint32_t*>(R.bases2) + (rbit >> 5);
      const int rsh = (int)(rbit & 31), fsh = (2 * ca) & 31;
      const uint32_t* fw = S.ref2 + ((2 * ca) >> 5);
      uint32_t rlo = __ldg(rw), flo = fw[0];
      const uint32_t q_rel = s0_rel + (uint32_t)(rp0 - lb);  // quality index of locus l = q_rel + l
      auto step = [&](int l, uint32_t keep) {
        const uint32_t rhi = __ldg(++rw), fhi = *++fw;
        const uint32_t rd = __funnelshift_r(rlo, rhi, rsh);
        const uint32_t x = rd ^ __funnelshift_r(flo, fhi, fsh);
        rlo = rhi;
        flo = fhi;
        uint32_t m = (x | (x >> 1)) & keep;
        while (m) {
          const int bit = __ffs(m) - 1;
          m &= m - 1;
          push_event(l + (bit >> 1), rl, slot, GUAC_ALN_MISMATCH, (rd >> bit) & 3, q_rel + (uint32_t)(l + (bit >> 1)));
        }
      };
      int l = ca;
      for (; l + 16 <= cb; l += 16) step(l, 0x55555555u);
      if (l < cb) step(l, ((1u << (2 * (cb - l))) - 1) & 0x55555555u);  // the tail word, masked once
    };
    if (rf & RF_FAST) {
      compare_segment(st, en, 0);
      continue;
    }
    // operator walk (PE.scala:59-117 per operator instead of per base)
    const uint32_t c0 = __ldg(R.cigar_off + r), c1 = __ldg(R.cigar_off + r + 1);
    const int read_len = (int)(__ldg(R.seq_off + r + 1) - s0);
    int cl = st, rp = 0;
    for (uint32_t ci = c0; ci < c1; ++ci) {
      const uint32_t w = __ldg(R.cigar + ci);
      const int op = w & 15, len = (int)(w >> 4);
      if (op == OPC_M || op == OPC_EQ || op == OPC_X) {
        int nop = -1, nlen = 0;
        if (ci + 1 < c1) {
          const uint32_t nw = __ldg(R.cigar + ci + 1);
          nop = nw & 15;
          nlen = (int)(nw >> 4);
        }
        // the final base of the operator anchors a following insertion (M/= only) or deletion
        const bool special_last = nop == OPC_D || (nop == OPC_I && op != OPC_X);
        compare_segment(cl, cl + len - (special_last ? 1 : 0), rp);
        if (special_last) {
          const int ll = cl + len - 1, k = rp + len - 1;
          if (ll >= 0 && ll < n_loci) {
            if (nop == OPC_I) {  // qualityScore = min over anchor + inserted qualities (PE.scala:153)
              const int until = min(k + nlen + 1, read_len);
              int q = 255;
              uint32_t h = 2166136261u ^ (uint32_t)(until - k);  // identity hint of the inserted sequence
              for (int j = k; j < until; ++j) {
                q = min(q, (int)__ldg(R.qual + s0 + j));
                const uint64_t gi = s0 + (uint64_t)j;
                h = (h ^ ((__ldg(R.bases2 + (gi >> 2)) >> ((gi & 3) * 2)) & 3)) * 16777619u;
              }
              push_event(ll, rl, slot, GUAC_ALN_INSERTION, 4, direct(q, (int)((h ^ (h >> 8) ^ (h >> 16) ^ (h >> 24)) & 0xFF)));
            } else {  // Deletion: the deleted reference bases must exist (PE.scala:95)
              for (int j = max(1, n_loci - ll); j <= nlen; ++j)  // inside the tile the reference is known to be there
                if (ref.at(t0 + ll + j) == 0) report_error(A.err, DERR_REF_MISSING, tile.contig_id, t0 + ll + j, r);
              push_event(ll, rl, slot, GUAC_ALN_DELETION, nlen < 255 ? 4 : 5, direct(__ldg(R.qual + s0 + k), nlen));
            }
          }
        }
        cl += len;
        rp += len;
      } else if (op == OPC_D) {
        const int mq = __ldg(R.mapq + r);
        for (int ll = max(cl, 0); ll < min(cl + len, n_loci); ++ll) push_event(ll, rl, slot, GUAC_ALN_MID_DELETION, 4, direct(mq, 0));
        cl += len;
      } else if (op == OPC_I || op == OPC_S) {
        rp += len;
      }
    }
  }
  __syncthreads();
  // reads that need the exact cursor: one warp per read, one lane per locus
  TK_PHASE(1);
  {
    const int n_slow = min(S.n_slow, CFG::SLOW_MAX);
    for (int si = warp; si < n_slow; si += CFG::WARPS) {
      const int entry = S.slow[si], rl = entry & 0x7FFF;
      const int64_t r = tile.read_lo + rl;
      if (STREAM && (entry & 0x8000)) {  // the read's events from the overflow arena, a lane each
        const uint4 hdr = __ldg(hdrs + rl);
        const int smp = (int)((hdr.y >> 16) & 0xFF);
        const int slot = MODE == 2 ? (int)((P.sample_mask >> smp) & 1u) : (smp == P.tumor_sample ? 1 : 0);
        if (MODE == 2) ev_sample = A.sample_bits ? (uint32_t)(smp & 3) << 30 : 0u;
        const uint2 head = __ldg(A.ev_by_read + r * STREAM_EV_CAP_C);  // (arena record, number of events)
        const uint2* evs = A.ev_overflow + (size_t)head.x * STREAM_OVF_CAP;
        const uint32_t s0_rel = (uint32_t)(((uint64_t)hdr.z | ((uint64_t)hdr.w << 32)) - seq_base);
        for (int k = lane; k < (int)head.y; k += 32) take_event(__ldg(evs + k), (int)(hdr.x - t0_lo), rl, slot, s0_rel, r);
        continue;
      }
      const int smp = (int)__ldg(R.sample_id + r);
      const int slot = MODE == 2 ? (int)((P.sample_mask >> smp) & 1u) : (smp == P.tumor_sample ? 1 : 0);
      if (MODE == 2) ev_sample = A.sample_bits ? (uint32_t)(smp & 3) << 30 : 0u;
      const int a = (int)max((long long)(__ldg(R.start + r) - t0), 0ll);
      const int b = (int)min((long long)(__ldg(R.end + r) - t0), (long long)n_ev_loci);
      for (int l = a + lane; l < b; l += 32) {
        Elem e;
        const int ec = element_at(R, r, t0 + l, ref, e);
        if (ec != DERR_NONE) {
          report_error(A.err, ec, tile.contig_id, t0 + l, r);
        } else if (e.kind != GUAC_ALN_MATCH) {
          // indel alleles met on the exact path are resolved by the heavy kernel (code 5)
          const int code = e.kind == GUAC_ALN_MISMATCH ? base_code(e.base)
                           : (e.kind == GUAC_ALN_INSERTION || e.kind == GUAC_ALN_DELETION) ? 5 : 4;
          push_event(l, rl, slot, e.kind, code, direct(e.quality, 0));
        }
      }
    }
  }
  __syncthreads();
  if (S.overflow) {  // a capacity was exceeded: hand the whole tile to the generic kernel
    if (tid == 0) A.fallback_tiles[atomicAdd(A.n_fallback, 1)] = blockIdx.x;
    return;
  }

  // ---- B. one scan over both packed arrays ----------------------------------------------------
  TK_PHASE(2);
  {
    constexpr int PER = (TL + CFG::THREADS) / CFG::THREADS;  // covers TL + 1 entries
    const int b0 = tid * PER, e0 = min(n_ev_loci + 1, b0 + PER);
    int s_cov = 0, s_aux = 0, s_cov2 = 0;
    for (int i = b0; i < e0; ++i) {
      if (i < n_ev_loci) {  // add the locus's event count to the high half of aux
        const int nr = S.nonref[i];
        S.aux[i] += ((nr & 0xFFFF) + (nr >> 16)) << 16;
      }
      s_cov += S.cov[i];
      s_aux += S.aux[i];
      if (MODE == 2) s_cov2 += S.cov2[i];
    }
    int i_cov = s_cov, i_aux = s_aux, i_cov2 = s_cov2;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int oc = __shfl_up_sync(FULL, i_cov, d), oa = __shfl_up_sync(FULL, i_aux, d);
      const int o2 = MODE == 2 ? __shfl_up_sync(FULL, i_cov2, d) : 0;
      if (lane >= d) {
        i_cov += oc;
        i_aux += oa;
        i_cov2 += o2;
      }
    }
    if (lane == 31) {
      S.warp_tot[0][warp] = i_cov;
      S.warp_tot[1][warp] = i_aux;
      if (MODE == 2) S.warp_tot2[warp] = i_cov2;
    }
    __syncthreads();
    // totals of the warps before this one: lane w holds warp w's, one warp-wide add each
    static_assert(CFG::WARPS <= 32, "one lane per warp total");
    const bool before = lane < warp;  // (warp < WARPS)
    const int c_cov = __reduce_add_sync(FULL, before ? S.warp_tot[0][lane] : 0);
    const int c_aux = __reduce_add_sync(FULL, before ? S.warp_tot[1][lane] : 0);
    const int c_cov2 = MODE == 2 ? __reduce_add_sync(FULL, before ? S.warp_tot2[lane] : 0) : 0;
    int r_cov = c_cov + i_cov - s_cov, r_aux = c_aux + i_aux - s_aux, r_cov2 = c_cov2 + i_cov2 - s_cov2;
    for (int i = b0; i < e0; ++i) {
      r_cov += S.cov[i];
      r_aux += S.aux[i];
      S.cov[i] = r_cov;
      S.aux[i] = r_aux;
      if (MODE == 2) {
        r_cov2 += S.cov2[i];
        S.cov2[i] = r_cov2;
      }
    }
    __syncthreads();
  }
  // bucket the events by locus: events of locus l occupy [end(l) - count(l), end(l))
  TK_PHASE(3);
  const int n_events = S.n_events;
  for (int i = tid; i < n_events; i += CFG::THREADS) {
    const uint2 ev = S.ev[i];
    const int l = ev.x & 0x7FF;
    const int nr = S.nonref[l];
    const int first = (S.aux[l] >> 16) - ((nr & 0xFFFF) + (nr >> 16));
    S.ev_order[first + atomicAdd(&S.cursor[l], 1)] = (uint16_t)i;
  }
  __syncthreads();

  // ---- C. one thread per locus ------------------------------------------------------------------
  TK_PHASE(4);
  // MODE 2: two lists of tile loci share S.cursor's words as 16-bit halves (the bucketing above was its last use)
  uint16_t* const scan_list = reinterpret_cast<uint16_t*>(S.cursor);
  uint16_t* const prune_list = scan_list + (TL + 1);
  auto route = [&](int l, bool to_simple) {  // pileup locus t0 + l goes to a per-locus kernel's list
    const int h = atomicAdd(to_simple ? A.n_simple : A.n_heavy, 1);
    if (h < (to_simple ? A.max_simple : A.max_heavy)) {
      Tile t;
      t.locus0 = t0 + l;
      t.n_loci = 1;
      t.contig_id = tile.contig_id;
      t.slot0 = tile.slot0 + l;
      t.read_lo = tile.read_lo;
      t.read_hi = tile.read_hi;
      t.ref_index = tile.ref_index;
      t._pad = 0;
      t.seq_base = t.seq_end = 0;
      (to_simple ? A.simple : A.heavy)[h] = t;
    }
  };
  int my_visited = 0;
  for (int l = tid; l < n_loci; l += CFG::THREADS) {
    const int cv = S.cov[l], nr = S.nonref[l];
    const int dN = cv & 0xFFFF, dT = cv >> 16;
    const int nrN = nr & 0xFFFF, nrT = nr >> 16;
    const int64_t slot = tile.slot0 + l, x = t0 + l;
    const bool any_depth = (MODE == 2 && A.sample_bits) ? (cv | S.cov2[l]) != 0 : dN + dT > 0;
    A.visited[slot] = any_depth ? 1 : 0;
    my_visited += any_depth;
    // (a tile with a missing reference base never gets here: it went to the generic kernel, which reports the error)
    bool candidate, simple_site = false, one_base = false;
    if (MODE == 0) {  // early-outs (SomaticStandardCaller.scala:164-172)
      candidate = dT > 0 && dN > 0 && dT <= P.max_read_depth && dN <= P.max_read_depth && nrT != 0 && nrN <= P.max_normal_alt;
    } else if (MODE == 1) {
      candidate = dT > 0 && nrT >= P.min_alt_reads;
    } else {
      candidate = false;
      if (any_depth) {
        // reference.getReferenceBase(contig, locus + 1) (jointcaller/evidence/MultiSampleMultiAlleleEvidence.scala:93)
        if (ref.at(x + 1) == 0) report_error(A.err, DERR_REF_MISSING, tile.contig_id, x + 1, -1);
        bool forced = false;
        for (int i = 0; i < A.n_force; ++i)
          forced = forced || (A.force[i].contig_id == tile.contig_id && A.force[i].start <= x + 1 && x + 1 < A.force[i].end);
        if (forced) {
          candidate = true;  // force-called: always evaluated, by the generic kernels
        } else if (l + 1 >= n_ev_loci) {
          candidate = true;  // (single-kernel form: the events of x + 1 are in the next tile)
        } else if ((S.nonref[l + 1] >> 16) >= P.min_alt_reads) {
          if (A.sample_bits) {
            // its events are looked at in the loop after this one, where every lane has such a locus (here one lane in
            // six does, and the warp would wait for it)
            scan_list[atomicAdd(&S.n_scan, 1)] = (uint16_t)l;
            continue;
          }
          candidate = true;
        }
      }
    }
    int flags = 0;
    if (MODE != 0 && candidate) {  // every candidate locus goes to the per-locus kernel
      route(l, false);
      flags = 1;
    }
    if (MODE == 0 && candidate) {  // the few candidate loci are examined densely in a second pass
      S.cursor[atomicAdd(&S.n_cand, 1)] = l;
      continue;
    }
    if (MODE == 2) continue;  // no per-locus summary in the joint caller
    // dense summary (normal, tumor): one 16-byte store per locus
    uint4 v;
    v.x = (uint32_t)min(dN, 65535) | ((uint32_t)min(dN - nrN, 65535) << 16);
    v.y = (uint32_t)flags << 16;
    v.z = (uint32_t)min(dT, 65535) | ((uint32_t)min(dT - nrT, 65535) << 16);
    v.w = (uint32_t)flags << 16;
    reinterpret_cast<uint4*>(A.summary)[slot] = v;
  }
  my_visited = __reduce_add_sync(FULL, my_visited);
  if (lane == 0 && my_visited) atomicAdd(A.n_visited, (unsigned long long)my_visited);
  if (MODE == 2 && A.sample_bits) {
    __syncthreads();
    TK_PHASE(5);
    const int n_scan = S.n_scan;
    for (int i = tid; i < n_scan; i += CFG::THREADS) {
      const int l = scan_list[i];
      bool candidate, simple_site, one_base;
      {
        // a locus with many events (a germline site: every other read) would keep this lane busy long after the other
        // lanes are done: the warp pass below scans it, a lane per event
        const int nr1 = S.nonref[l + 1];
        if ((nr1 & 0xFFFF) + (nr1 >> 16) > SJ_SCAN_SERIAL_MAX) {
          prune_list[atomicAdd(&S.n_cand, 1)] = (uint16_t)(l | 0x8000);
          continue;
        }
      }
      if (GUAC_SJ_PAIRSCAN && P.min_alt_reads == 2) {
        candidate = sj_scan_pairs(S, l + 1, simple_site, one_base);
      } else {
        unsigned long long hist[4];
        candidate = sj_scan_events(S, l + 1, P.min_alt_reads, hist, simple_site, one_base) >= P.min_alt_reads;
      }
      if (!candidate) continue;
      if (one_base && A.prune) prune_list[atomicAdd(&S.n_cand, 1)] = (uint16_t)l;  // examined below, a warp each
      else route(l, simple_site);
    }
  }
  if (MODE == 2 && A.sample_bits) {
    // prune: a simple site is listed for the site kernel only if one of its possible alleles can be called.  With every
    // read over x + 1 at P(correct) > 1/2 (no flagged read there), a read on the reference base can only lower a
    // het / hom-alt genotype or a tumor mixture against the reference one (its term is log10(1/2) - log10 p,
    // log10((1 - p) / p) or log10(((1 - v) p + v (1 - p)) / p): all <= 0), reads on other bases add exactly 0, and the
    // alternate reads' terms — exact, from their qualities — bound every posterior gain
    // (MultiSampleSingleAlleleEvidence.scala:66-230).  v is bounded above with the event count over the depth.
    __syncthreads();
    TK_PHASE(6);
#ifdef GUAC_TK_PHASE_CLOCKS
    if (threadIdx.x == 0) {
      atomicAdd(&g_tk_phase_clocks[9], 1ull);
      atomicAdd(&g_tk_phase_clocks[10], (unsigned long long)S.n_scan);
      atomicAdd(&g_tk_phase_clocks[11], (unsigned long long)S.n_cand);
      atomicAdd(&g_tk_phase_clocks[12], (unsigned long long)S.n_events);
      atomicAdd(&g_tk_phase_clocks[13], (unsigned long long)n_reads_tile);
    }
#endif
    const int n_cand = S.n_cand;
    const double LOG10_HALF = -0.3010299956639812;
    // One warp per candidate, one lane per event of x + 1: the loads of the alternate reads' mapq and quality and their
    // logarithms run side by side; the sums are combined by fixed butterflies.
    for (int ci = warp; ci < n_cand; ci += CFG::WARPS) {
      const int entry = prune_list[ci], l = entry & 0x7FFF;
      unsigned long long hist[4];
      bool simple_site, one_base;
#ifdef GUAC_TK_PHASE_CLOCKS
      long long sub_t_ = clock64();
#endif
      const int most = sj_scan_events_warp(S, l + 1, P.min_alt_reads, hist, simple_site, one_base);
#ifdef GUAC_TK_PHASE_CLOCKS
      if (threadIdx.x == 0) { const long long n_ = clock64(); atomicAdd(&g_tk_phase_clocks[14], (unsigned long long)(n_ - sub_t_)); sub_t_ = n_; }
#endif
      if (entry & 0x8000) {  // not looked at yet: is it a candidate at all, and one for the bound?
        if (most < P.min_alt_reads) continue;
        if (!(one_base && A.prune)) {
          if (lane == 0) route(l, simple_site);
          continue;
        }
      }
      // a read that may have P(correct) <= 1/2 covers x + 1; or some allele has so many reads that no bound can hold
      // (keeping a site is always right: the site kernel decides)
      bool keep = (S.aux[l + 1] & 0xFFFF) != 0 || most >= SJ_PRUNE_MAX_ALT_READS;
      {
        // Clipped elements at x leave the depths (MultiSampleMultiAlleleEvidence.scala:85-86): rare, no shortcut
        const int nr0 = S.nonref[l], f1 = S.aux[l] >> 16, f0 = f1 - ((nr0 & 0xFFFF) + (nr0 >> 16));
        for (int i = f0; i < f1; ++i) keep = keep || ((S.ev[S.ev_order[i]].x >> 24) & 7) == GUAC_ALN_CLIPPED;
      }
      const int cv = S.cov[l], cv2 = S.cov2[l];
      const int d0 = cv & 0xFFFF, d1 = cv >> 16, d2 = cv2 & 0xFFFF, d3 = cv2 >> 16;
      const int nr1 = S.nonref[l + 1], e1 = S.aux[l + 1] >> 16, e0 = e1 - ((nr1 & 0xFFFF) + (nr1 >> 16));
      auto tumor = [&](int s) { return ((A.kinds >> (2 * s)) & 3u) == GUAC_SJ_TUMOR_DNA; };
      const int dp = (tumor(0) ? d0 : 0) + (tumor(1) ? d1 : 0) + (tumor(2) ? d2 : 0) + (tumor(3) ? d3 : 0);
      if (!keep) {
        // One possible allele (a base with min_reads equal events in a discovery sample) at a time — there is seldom a
        // second — so a lane carries seven sums, not twenty-eight.  Events of other bases add exactly 0 to this allele's
        // sums whatever their quality (they miss every allele of either hypothesis).
        auto count_of = [&](int s, int c) {
          const unsigned long long h = s == 0 ? hist[0] : s == 1 ? hist[1] : s == 2 ? hist[2] : hist[3];
          return (int)((h >> (16 * c)) & 0xFFFFull);
        };
        auto wsum = [&](double v) {
#pragma unroll
          for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(FULL, v, d);
          return v;
        };
        const double margin = -1e-6;  // dwarfs the rounding of the sums; costs nothing in pruning power
        for (int c = 0; c < 4 && !keep; ++c) {
          const int k0 = count_of(0, c), k1 = count_of(1, c), k2 = count_of(2, c), k3 = count_of(3, c);
          if (max(max(k0, k1), max(k2, k3)) < P.min_alt_reads) continue;  // not a possible allele (uniform across the warp)
          const int kp = (tumor(0) ? k0 : 0) + (tumor(1) ? k1 : 0) + (tumor(2) ? k2 : 0) + (tumor(3) ? k3 : 0);
          const double vp = fmin(1.0, fmax(A.vaf_floor, (double)kp / dp));  // (only used with a tumor read: dp > 0 then)
          double het = 0, hom = 0, tp = 0, tm0 = 0, tm1 = 0, tm2 = 0, tm3 = 0;
          bool low = false;
          for (int i = e0 + lane; i < e1; i += 32) {
            const uint2 ev = S.ev[S.ev_order[i]];
            if (((ev.x >> 24) & 7) != GUAC_ALN_MISMATCH || (int)((ev.x >> 27) & 7) != c) continue;
            const int smp = (int)(ev.x >> 30), rl = (int)((ev.x >> 11) & 0xFFF);
            // (the three loads below depend on the event alone: requested together)
            const uint32_t hy = __ldg(A.hdr + tile.read_lo + rl).y;
            const int mapq = (int)__ldg(R.mapq + tile.read_lo + rl);
            const int q = (ev.y >> 31) ? (int)(ev.y & 0xFF) : (int)__ldg(R.qual + seq_base + ev.y);
            {
              // The bound is about one-base alleles.  A read whose next element is not a Match either starts a longer
              // allele (ReadSubsequence.scala:117-121), and with it come multi-base subsequences: the site kernel's case.
              // So is a read the exact cursor walked (an N skip after x + 1 leaves no event), and the tile's last site,
              // whose locus x + 2 lies outside the kept events.
              bool longer = l + 2 >= n_ev_loci || !((hy >> 24) & RF_MEDIUM);
              if (!longer) {
                const int nr2 = S.nonref[l + 2], g1 = S.aux[l + 2] >> 16, g0 = g1 - ((nr2 & 0xFFFF) + (nr2 >> 16));
                for (int j = g0; j < g1; ++j) longer = longer || (int)((S.ev[S.ev_order[j]].x >> 11) & 0xFFF) == rl;
              }
              if (longer) {
                low = true;
                continue;
              }
            }
            if (mapq == 0) continue;  // contributes exactly 0.0 (PileupStats.scala:79-80)
            const double p = c_success_probability[q] * c_success_probability[mapq];
            if (!(p > 0.5)) {
              low = true;
              continue;
            }
            const double2 ab = __ldg(A.log_table + ((mapq << 8) | q));
            const unsigned sample_kind = (A.kinds >> (2 * smp)) & 3u;
            if (sample_kind == GUAC_SJ_NORMAL_DNA) {
              het += fmax(0.0, LOG10_HALF - ab.y);
              hom += fmax(0.0, ab.x - ab.y);
            } else if (sample_kind == GUAC_SJ_TUMOR_DNA) {
              // alt fraction bounded above by (events of this base in the sample) / depth, at most 1
              const int ks = smp == 0 ? k0 : smp == 1 ? k1 : smp == 2 ? k2 : k3;
              const int ds = smp == 0 ? d0 : smp == 1 ? d1 : smp == 2 ? d2 : d3;
              const double v = fmin(1.0, fmax(A.vaf_floor, (double)ks / ds));
              const double t = fmax(0.0, log10(v * p + (1.0 - v) * (1.0 - p)) - ab.y);
              tp += fmax(0.0, log10(vp * p + (1.0 - vp) * (1.0 - p)) - ab.y);
              tm0 += smp == 0 ? t : 0.0;
              tm1 += smp == 1 ? t : 0.0;
              tm2 += smp == 2 ? t : 0.0;
              tm3 += smp == 3 ? t : 0.0;
            }
          }
          if (__any_sync(FULL, low)) {
            keep = true;
            break;
          }
          const double g_het = wsum(het), g_hom = wsum(hom);
          const double g_t = fmax(fmax(fmax(wsum(tm0), wsum(tm1)), fmax(wsum(tm2), wsum(tm3))), wsum(tp));
          keep = !(g_het - A.het_prior <= margin) || !(g_hom - A.hom_prior <= margin) || !(g_t - A.somatic_prior <= margin);
        }
#ifdef GUAC_TK_PHASE_CLOCKS
        if (threadIdx.x == 0) { const long long n_ = clock64(); atomicAdd(&g_tk_phase_clocks[15], (unsigned long long)(n_ - sub_t_)); sub_t_ = n_; }
#endif
      }
      if (!keep || lane != 0) continue;
      // kept: the fast site kernel where every subsequence is one standard base, the generic kernels otherwise
      const int h = atomicAdd(simple_site ? A.n_simple : A.n_heavy, 1);
      if (h < (simple_site ? A.max_simple : A.max_heavy)) {
        Tile t;
        t.locus0 = t0 + l;
        t.n_loci = 1;
        t.contig_id = tile.contig_id;
        t.slot0 = tile.slot0 + l;
        t.read_lo = tile.read_lo;
        t.read_hi = tile.read_hi;
        t.ref_index = tile.ref_index;
        t._pad = 0;
        t.seq_base = t.seq_end = 0;
        (simple_site ? A.simple : A.heavy)[h] = t;
      }
    }
  }
  TK_PHASE(7);
  if (MODE != 0) return;
  __syncthreads();
  const int n_cand = S.n_cand;
  for (int ci = tid; ci < n_cand; ci += CFG::THREADS) {
    const int l = S.cursor[ci];
    const int cv = S.cov[l], nr = S.nonref[l];
    const int dN = cv & 0xFFFF, dT = cv >> 16;
    const int nrN = nr & 0xFFFF, nrT = nr >> 16;
    const int64_t slot = tile.slot0 + l, x = t0 + l;
    int top_t = 0, top_n = 0, flags = 0;
    {
      const int e1 = S.aux[l] >> 16, e0 = e1 - (nrN + nrT);
      bool simple = e1 - e0 <= CFG::EV_SIMPLE;
      // Distinct tumor variant alleles of the locus, straight from its (few) events.  An allele is
      // identified by (kind, base) for SNVs, (kind, length) for deletions, kind alone for mid-deletions
      // and (kind, 8-bit hash) for insertions — the last only when no other insertion shares the hash.
      auto akey = [&](uint2 ev) {
        const int kind = (ev.x >> 24) & 7, code = (ev.x >> 27) & 7;
        return (kind << 8) | (kind == GUAC_ALN_MISMATCH ? code : (kind == GUAC_ALN_MID_DELETION ? 0 : (int)((ev.y >> 8) & 0xFF)));
      };
      int best = -1, bcnt = 0, bfirst = INT_MAX;
      if (simple)
        for (int i = e0; i < e1; ++i) {
          const uint2 ev = S.ev[S.ev_order[i]];
          const int eslot = (ev.x >> 23) & 1, kind = (ev.x >> 24) & 7, code = (ev.x >> 27) & 7;
          if (code == 5 || (kind == GUAC_ALN_MISMATCH && code == 4)) {  // either sample: allele not identified here
            simple = false;
            break;
          }
          if (eslot != 1 || kind == GUAC_ALN_CLIPPED) continue;  // Clipped is not a variant allele
          const int key = akey(ev);
          int c = 0, first = INT_MAX;
          bool seen_before = false;
          for (int j = e0; j < e1; ++j) {
            const uint2 ej = S.ev[S.ev_order[j]];
            if (((ej.x >> 23) & 1) != 1 || ((ej.x >> 24) & 7) == GUAC_ALN_CLIPPED || akey(ej) != key) continue;
            if (j < i) seen_before = true;
            ++c;
            first = min(first, (int)((ej.x >> 11) & 0xFFF));
          }
          if (seen_before) continue;
          // most frequent allele; ties -> lowest first supporting read (documented canonical order)
          if (c > bcnt || (c == bcnt && first < bfirst)) {
            best = key;
            bcnt = c;
            bfirst = first;
          }
        }
      if (simple && best < 0) simple = false;  // maxBy over an empty map (:188): the heavy path reports it
      if (simple && (best >> 8) == GUAC_ALN_INSERTION) {
        int same = 0;
        for (int j = e0; j < e1; ++j) {
          const uint2 ej = S.ev[S.ev_order[j]];
          same += ((ej.x >> 24) & 7) == GUAC_ALN_INSERTION && akey(ej) == best;
        }
        if (same > 1) simple = false;  // equal hashes: the exact path compares the sequences
      }
      if (simple) {
        const double f = fmax(P.min_vaf, (double)bcnt / (double)dT);  // :189
        // prune: tumorAltLOD = sum over tumor elements of log P(e|mix) - log P(e|ref), where an element
        //   of the reference allele contributes log(1 - f + f (1-p)/p)  <= 0 whenever p >= 1/2,
        //   of the called allele             log(1 - f + f p/(1-p)),
        //   of anything else                 log(1) = 0.
        // If every tumor read over the locus has all base qualities >= 4 (lowq == 0) the called allele's
        // terms bound the LOD from above; a bound below the threshold (minus a 0.5 margin that dwarfs
        // rounding) means no call is possible (:231) and the matching bases' qualities are never read.
        // The bound is summed in single precision (error < 1e-3 over a pileup, far inside the 0.5 margin):
        // log(1 - f + f p/(1-p)) = log(1 + f (odds - 1)), odds - 1 from a constant table indexed by the quality.
        float bound = 0.0f;
        int n_alt_normal = 0;
        const float ff = (float)f;
        for (int i = e0; i < e1; ++i) {
          const uint2 ev = S.ev[S.ev_order[i]];
          const int eslot = (ev.x >> 23) & 1, kind = (ev.x >> 24) & 7;
          if (kind == GUAC_ALN_CLIPPED || akey(ev) != best) continue;
          if (eslot == 0) {
            ++n_alt_normal;
          } else {
            const int q = (ev.y >> 31) ? (int)(ev.y & 0xFF) : (int)__ldg(R.qual + seq_base + ev.y);
            bound += __logf(fmaf(ff, c_odds_minus_one[q], 1.0f));
          }
        }
        // tumor reads over the locus that are not RF_HIQ: each adds at most log(1 + f * LOWQ_ODDS_M1) if its element is
        // of the reference allele (nothing otherwise); one with a quality-0 base anywhere sends the locus to the exact path
        const int lowq = S.aux[l] & 0xFFFF;
        bound += (float)lowq * __logf(fmaf(ff, LOWQ_ODDS_M1, 1.0f));
        // a mid-deletion allele has empty alt bases and is never called (:233)
        const bool never_called = (best >> 8) == GUAC_ALN_MID_DELETION;
        if (never_called || (lowq < ZEROQ_WEIGHT && bound < (float)P.tumor_odds - 0.5f)) {
          top_t = bcnt;
          top_n = n_alt_normal;
          flags = 1;
        } else {
          simple = false;
        }
      }
      if (!simple) {  // full evaluation by the generic kernel
        const int h = atomicAdd(A.n_heavy, 1);
        if (h < A.max_heavy) {
          Tile t;
          t.locus0 = x;
          t.n_loci = 1;
          t.contig_id = tile.contig_id;
          t.slot0 = slot;
          t.read_lo = tile.read_lo;
          t.read_hi = tile.read_hi;
          t.ref_index = tile.ref_index;
          t._pad = 0;
          t.seq_base = t.seq_end = 0;
        t.seq_base = t.seq_end = 0;
          A.heavy[h] = t;
        }
        continue;  // the heavy kernel writes this locus's summary
      }
    }
    // dense summary (normal, tumor): one 16-byte store per locus
    uint4 v;
    v.x = (uint32_t)min(dN, 65535) | ((uint32_t)min(dN - nrN, 65535) << 16);
    v.y = (uint32_t)min(top_n, 65535) | ((uint32_t)flags << 16);
    v.z = (uint32_t)min(dT, 65535) | ((uint32_t)min(dT - nrT, 65535) << 16);
    v.w = (uint32_t)min(top_t, 65535) | ((uint32_t)flags << 16);
    reinterpret_cast<uint4*>(A.summary)[slot] = v;
  }
  TK_PHASE(8);
}

}  // namespace guac
