// tile_kernels.cuh — streaming tile kernel for somatic-standard (the benchmarked path).
//
// One CTA owns a tile of TL consecutive loci and everything about it stays in shared memory:
//   A. read-parallel pass: every read that can overlap the tile is visited ONCE; its coverage
//      goes into per-sample difference arrays (+1 at start, -1 at end), and its bases are
//      compared with the reference 32 bits (16 bases) at a time in 2-bit space — only
//      non-matching elements (mismatches, indel anchors, mid-deletions, N-skips) become events.
//      Reads with a multi-operator CIGAR or non-ACGT bases (~3 %) take the exact per-locus
//      cursor (element_at), one warp per read, one lane per locus.
//   B. block prefix scans turn the difference arrays into Pileup.depth per sample, give
//      referenceDepth = depth - non-match events, bucket the events by locus, and index the
//      reads by start so that a locus finds its overlapping reads without a search.
//   C. loci that survive findPotentialVariantAtLocus's early-outs (SomaticStandardCaller.scala:
//      164-172; ~20 % of loci at 1 % error, depth 60/30) are evaluated one warp per locus: the
//      most frequent tumor variant allele comes from the locus's events, the likelihood sums
//      from a gather of the overlapping reads' qualities.  Log terms that do not depend on the
//      locus's allele fraction come from tables; the f-dependent tumor mixture terms are
//      histogrammed by (class, quality) so that one warp-wide FP64 log pass serves the locus.
// Tiles that exceed a shared-memory capacity are appended to a fallback list and re-run by
// the generic gather kernel (kernels.cuh) — still on the GPU.
#pragma once
#include "kernels.cuh"

namespace guac {

template <int TL_>
struct TileCfg {
  static constexpr int TL = TL_;           // loci per tile
  static constexpr int THREADS = 256;
  static constexpr int WARPS = THREADS / 32;
  static constexpr int R_MAX = TL * 2 + 256;  // reads staged per tile
  static constexpr int E_MAX = TL * 4;        // non-match events per tile
  static constexpr int SLOW_MAX = 512;
  static constexpr int QB = 64;               // quality bins of the tumor histogram
};

// log tables indexed [cls][q], cls 0 = element allele is the reference allele, 1 = is the called
// allele, 2 = anything else; filled by init_log_tables_kernel with the device's own log().
struct LogTables {
  double ref[3][256];  // log(P(e | ref genotype))           (Likelihood.scala:199-204, f = 1)
  double het[3][256];  // log(0.5 P(e|ref) + 0.5 P(e|alt))   (germlineVariantGenotype, SomaticStandardCaller.scala:212-218)
};

__global__ void init_log_tables_kernel(LogTables* t) {
  const int q = threadIdx.x;
  const double p = c_success_probability[q];
  for (int cls = 0; cls < 3; ++cls) {
    const double pr = cls == 0 ? p : 1 - p;
    const double pa = cls == 1 ? p : 1 - p;
    t->ref[cls][q] = log(pr * 1.0);
    t->het[cls][q] = log(pr * 0.5 + pa * 0.5);
  }
}

// per-read derived flags, computed once per upload
enum : uint8_t {
  RF_FAST = 1,  // single M/=/X operator and no non-ACGT base: the bit-parallel path applies
  RF_HIQ = 2    // every base quality >= 4, i.e. P(correct) >= 0.60 > 1/2 (licenses the LOD bound)
};

__global__ void read_fast_flags_kernel(ReadsView R, uint8_t* __restrict__ rflags) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= R.n_reads) return;
  const uint32_t c0 = R.cigar_off[r], c1 = R.cigar_off[r + 1];
  bool fast = false;
  if (c1 - c0 == 1) {
    const uint32_t w = R.cigar[c0];
    const int op = w & 15;
    const uint64_t s0 = R.seq_off[r], s1 = R.seq_off[r + 1];
    fast = (op == OPC_M || op == OPC_EQ || op == OPC_X) && (uint64_t)(w >> 4) == s1 - s0 && s1 > s0;
    for (uint64_t i = s0 >> 3; fast && i <= ((s1 - 1) >> 3); ++i) {
      uint8_t m = R.nmask[i];
      if (i == (s0 >> 3)) m &= (uint8_t)(0xFFu << (s0 & 7));
      if (i == ((s1 - 1) >> 3)) m &= (uint8_t)(0xFFu >> (7 - ((s1 - 1) & 7)));
      if (m) fast = false;
    }
  }
  bool hiq = true;
  for (uint64_t i = R.seq_off[r]; i < R.seq_off[r + 1]; ++i) hiq = hiq && R.qual[i] >= 4;
  rflags[r] = (fast ? RF_FAST : 0) | (hiq ? RF_HIQ : 0);
}

__device__ __forceinline__ uint32_t spread16(uint32_t x) {  // bit i -> bit 2i
  x &= 0xFFFF;
  x = (x | (x << 8)) & 0x00FF00FF;
  x = (x | (x << 4)) & 0x0F0F0F0F;
  x = (x | (x << 2)) & 0x33333333;
  x = (x | (x << 1)) & 0x55555555;
  return x;
}

// 32 bits starting at bit `bit` of a little-endian word stream in global memory
__device__ __forceinline__ uint32_t load_bits32_g(const uint8_t* __restrict__ bytes, uint64_t bit) {
  const uint64_t byte = bit >> 3;
  const uint32_t* w = reinterpret_cast<const uint32_t*>(bytes) + (byte >> 2);
  const int sh = (int)(bit & 31);
  const uint32_t lo = __ldg(w);
  if (sh == 0) return lo;
  return __funnelshift_r(lo, __ldg(w + 1), sh);
}
__device__ __forceinline__ uint32_t load_bits32_s(const uint32_t* words, int bit) {
  const int sh = bit & 31;
  const uint32_t lo = words[bit >> 5];
  if (sh == 0) return lo;
  return __funnelshift_r(lo, words[(bit >> 5) + 1], sh);
}

// in-place inclusive prefix sum of a shared int array (all threads of the CTA participate)
template <int THREADS>
__device__ __forceinline__ void block_scan_inplace(int* a, int n, int* warp_tot /*[THREADS/32]*/) {
  const int per = (n + THREADS - 1) / THREADS;
  const int b = threadIdx.x * per, e = min(n, b + per);
  int s = 0;
  for (int i = b; i < e; ++i) s += a[i];
  int inc = s;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    int o = __shfl_up_sync(0xffffffffu, inc, d);
    if ((threadIdx.x & 31) >= d) inc += o;
  }
  if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = inc;
  __syncthreads();
  int carry = 0;
  for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) carry += warp_tot[w];
  int run = carry + inc - s;
  for (int i = b; i < e; ++i) {
    run += a[i];
    a[i] = run;
  }
  __syncthreads();
}

// event word: [0,11) locus in tile | [11,23) read in tile | 23 sample slot | [24,27) kind | [27,30) base code (4 = non-ACGT)
__device__ __forceinline__ uint32_t pack_event(int l, int rl, int slot, int kind, int code) {
  return (uint32_t)l | ((uint32_t)rl << 11) | ((uint32_t)slot << 23) | ((uint32_t)kind << 24) | ((uint32_t)code << 27);
}

__device__ __forceinline__ int base_code(uint8_t b) { return b == 'A' ? 0 : b == 'C' ? 1 : b == 'G' ? 2 : b == 'T' ? 3 : 4; }

template <class CFG>
struct TileSmem {
  uint32_t ref2[CFG::TL / 16 + 2];
  uint32_t refmask[CFG::TL / 32 + 2];
  int diff[2][CFG::TL + 1];        // -> depth after the scan
  int nonref[2][CFG::TL];          // non-Match elements per sample
  int starts[CFG::TL + 1];         // -> reads with start <= locus after the scan
  int lowq[CFG::TL + 1];           // -> tumor reads over the locus that carry a base quality < 4
  int ev_off[CFG::TL + 1];         // events per locus -> offsets
  int cursor[CFG::TL];
  int r_start[CFG::R_MAX];         // relative to the tile's first locus
  int r_end[CFG::R_MAX];
  uint32_t r_s0[CFG::R_MAX];       // seq_off relative to the tile's first read
  uint8_t r_flags[CFG::R_MAX];     // bit0 sample slot (0 normal, 1 tumor), bit1 fast, bit2 relevant sample
  uint32_t ev[CFG::E_MAX];
  uint32_t ev_sorted[CFG::E_MAX];
  uint16_t slow[CFG::SLOW_MAX];
  uint16_t cand[CFG::TL];
  unsigned hist[CFG::WARPS][3 * CFG::QB];
  unsigned clist[CFG::WARPS][3 * CFG::QB];  // compacted non-empty bins: bin << 24 | count
  int warp_tot[CFG::WARPS];
  int n_events, n_slow, n_cand, max_span, overflow;
};

template <class CFG>
__global__ void __launch_bounds__(CFG::THREADS)
ss_tile_kernel(ReadsView R, const uint8_t* __restrict__ rflags, const RefView* __restrict__ refs, const Tile* __restrict__ tiles,
               SSParamsDev P, const LogTables* __restrict__ LT, uint8_t* __restrict__ visited,
               guac_locus_summary* __restrict__ summary, SSCallDev* __restrict__ calls, int* __restrict__ n_calls, int max_calls,
               int* __restrict__ fallback_tiles, int* __restrict__ n_fallback, int force_fallback_mod, DeviceError* err) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  TileSmem<CFG>& S = *reinterpret_cast<TileSmem<CFG>*>(smem_raw);
  constexpr int TL = CFG::TL;
  const unsigned FULL = 0xffffffffu;
  const Tile tile = tiles[blockIdx.x];
  const RefView ref = refs[tile.ref_index];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n_loci = tile.n_loci;
  const int64_t t0 = tile.locus0;
  const int64_t n_reads_tile = tile.read_hi - tile.read_lo;
  const uint64_t seq_base = n_reads_tile > 0 ? __ldg(R.seq_off + tile.read_lo) : 0;

  bool too_big = n_reads_tile > CFG::R_MAX;
  if (!too_big && n_reads_tile > 0) too_big = (__ldg(R.seq_off + tile.read_hi) - seq_base) >= (1ull << 31);
  if (force_fallback_mod > 0 && blockIdx.x % force_fallback_mod == 0) too_big = true;  // test hook
  if (too_big) {  // uniform across the CTA
    if (tid == 0) fallback_tiles[atomicAdd(n_fallback, 1)] = blockIdx.x;
    return;
  }

  // ---- init ------------------------------------------------------------------------------
  for (int i = tid; i < TL + 1; i += CFG::THREADS) {
    S.diff[0][i] = S.diff[1][i] = 0;
    S.starts[i] = 0;
    S.lowq[i] = 0;
    S.ev_off[i] = 0;
  }
  for (int i = tid; i < TL; i += CFG::THREADS) {
    S.nonref[0][i] = S.nonref[1][i] = 0;
    S.cursor[i] = 0;
  }
  for (int i = tid; i < CFG::WARPS * 3 * CFG::QB; i += CFG::THREADS) (&S.hist[0][0])[i] = 0;
  if (tid == 0) S.n_events = S.n_slow = S.n_cand = S.max_span = S.overflow = 0;
  // reference tile -> 2-bit words (16 loci per word) and a non-ACGT mask (32 loci per word)
  for (int w = tid; w < TL / 16 + 2; w += CFG::THREADS) {
    uint32_t bits = 0, mask = 0;
#pragma unroll 4
    for (int k = 0; k < 16; ++k) {
      const int l = w * 16 + k;
      const uint8_t b = l < n_loci ? ref.at(t0 + l) : (uint8_t)'A';
      const int c = base_code(b);
      if (c < 4) bits |= (uint32_t)c << (2 * k); else mask |= 1u << k;
    }
    S.ref2[w] = bits;
    // two consecutive 16-bit halves share a mask word
    reinterpret_cast<uint16_t*>(S.refmask)[w] = (uint16_t)mask;
  }
  __syncthreads();

  // ---- A. read-parallel pass ---------------------------------------------------------------
  for (int rl = tid; rl < (int)n_reads_tile; rl += CFG::THREADS) {
    const int64_t r = tile.read_lo + rl;
    const int64_t st64 = __ldg(R.start + r), en64 = __ldg(R.end + r);
    const int smp = (int)__ldg(R.sample_id + r);
    const bool relevant = smp == P.normal_sample || smp == P.tumor_sample;
    const int slot = smp == P.tumor_sample ? 1 : 0;
    const uint8_t rf = __ldg(rflags + r);
    const int st = (int)max((long long)(st64 - t0), -(1ll << 30)), en = (int)min((long long)(en64 - t0), (1ll << 30));
    S.r_start[rl] = st;
    S.r_end[rl] = en;
    S.r_s0[rl] = (uint32_t)(__ldg(R.seq_off + r) - seq_base);
    S.r_flags[rl] = (uint8_t)(slot | ((rf & RF_FAST) ? 2 : 0) | (relevant ? 4 : 0));
    atomicAdd(&S.starts[max(st, 0)], 1);  // st < n_loci by construction of read_hi
    if (!relevant) continue;
    const int a = max(st, 0), b = min(en, n_loci);  // overlap [a, b) in tile coordinates
    if (a >= b) continue;
    atomicMax(&S.max_span, en - st);
    atomicAdd(&S.diff[slot][a], 1);
    atomicAdd(&S.diff[slot][b], -1);
    if (slot == 1 && !(rf & RF_HIQ)) {
      atomicAdd(&S.lowq[a], 1);
      atomicAdd(&S.lowq[b], -1);
    }
    if (!(rf & RF_FAST)) {
      const int i = atomicAdd(&S.n_slow, 1);
      if (i < CFG::SLOW_MAX) S.slow[i] = (uint16_t)rl; else S.overflow = 1;
      continue;
    }
    // fast read: single M/=/X operator, ACGT only.  Compare 16 bases at a time.
    const uint64_t s0 = seq_base + S.r_s0[rl];
    for (int l = a; l < b; l += 16) {
      const int k = l - st;  // read position of locus l
      const uint32_t rd = load_bits32_g(R.bases2, 2 * (s0 + (uint64_t)k));
      const uint32_t rfw = load_bits32_s(S.ref2, 2 * l);
      uint32_t x = rd ^ rfw;
      uint32_t m = (x | (x >> 1)) & 0x55555555u;
      const uint32_t rmask = (load_bits32_s(S.refmask, l)) & 0xFFFFu;
      m |= spread16(rmask);  // a non-ACGT reference base never equals an ACGT read base
      const int nb = b - l;
      if (nb < 16) m &= (1u << (2 * nb)) - 1;
      while (m) {
        const int bit = __ffs(m) - 1;
        m &= m - 1;
        const int i16 = bit >> 1;
        const int code = (rd >> bit) & 3;
        const int ll = l + i16;
        atomicAdd(&S.nonref[slot][ll], 1);
        const int e = atomicAdd(&S.n_events, 1);
        if (e < CFG::E_MAX) S.ev[e] = pack_event(ll, rl, slot, GUAC_ALN_MISMATCH, code); else S.overflow = 1;
      }
    }
  }
  __syncthreads();
  // slow reads: one warp per read, one lane per locus, exact cursor
  {
    const int n_slow = min(S.n_slow, CFG::SLOW_MAX);
    for (int si = warp; si < n_slow; si += CFG::WARPS) {
      const int rl = S.slow[si];
      const int64_t r = tile.read_lo + rl;
      const int slot = S.r_flags[rl] & 1;
      const int a = max(S.r_start[rl], 0), b = min(S.r_end[rl], n_loci);
      for (int l = a + lane; l < b; l += 32) {
        Elem e;
        const int ec = element_at(R, r, t0 + l, ref, e);
        if (ec != DERR_NONE) {
          report_error(err, ec, tile.contig_id, t0 + l, r);
        } else if (e.kind != GUAC_ALN_MATCH) {
          atomicAdd(&S.nonref[slot][l], 1);
          const int i = atomicAdd(&S.n_events, 1);
          if (i < CFG::E_MAX) S.ev[i] = pack_event(l, rl, slot, e.kind, e.kind == GUAC_ALN_MISMATCH ? base_code(e.base) : 4); else S.overflow = 1;
        }
      }
    }
  }
  __syncthreads();
  if (S.overflow) {  // a capacity was exceeded: hand the whole tile to the generic kernel
    if (tid == 0) fallback_tiles[atomicAdd(n_fallback, 1)] = blockIdx.x;
    return;
  }

  // ---- B. scans -----------------------------------------------------------------------------
  for (int i = tid; i < n_loci; i += CFG::THREADS) S.ev_off[i + 1] = S.nonref[0][i] + S.nonref[1][i];
  __syncthreads();
  block_scan_inplace<CFG::THREADS>(S.diff[0], n_loci, S.warp_tot);
  block_scan_inplace<CFG::THREADS>(S.diff[1], n_loci, S.warp_tot);
  block_scan_inplace<CFG::THREADS>(S.starts, n_loci, S.warp_tot);
  block_scan_inplace<CFG::THREADS>(S.lowq, n_loci, S.warp_tot);
  block_scan_inplace<CFG::THREADS>(S.ev_off, n_loci + 1, S.warp_tot);
  // bucket the events by locus
  const int n_events = S.n_events;
  for (int i = tid; i < n_events; i += CFG::THREADS) {
    const uint32_t ev = S.ev[i];
    const int l = ev & 0x7FF;
    S.ev_sorted[S.ev_off[l] + atomicAdd(&S.cursor[l], 1)] = ev;
  }
  // per-locus early-outs (SomaticStandardCaller.scala:164-172) and the dense summary of
  // loci that need no likelihood
  for (int l = tid; l < n_loci; l += CFG::THREADS) {
    const int dN = S.diff[0][l], dT = S.diff[1][l];
    const int nrN = S.nonref[0][l], nrT = S.nonref[1][l];
    const int64_t slot = tile.slot0 + l;
    // visited: any sample's window is non-empty.  Reads of other samples (< n_samples) count too.
    bool any = dN + dT > 0;
    if (P.n_samples > 2 && !any) {
      // rare configuration (normal/tumor ids not {0,1}): test the staged reads directly
      const int hi = S.starts[l];
      for (int i = 0; i < hi && !any; ++i) any = S.r_start[i] <= l && l < S.r_end[i];
    }
    visited[slot] = any ? 1 : 0;
    if (dN + dT > 0 && ref.at(t0 + l) == 0) report_error(err, DERR_REF_MISSING, tile.contig_id, t0 + l, -1);
    const bool candidate = dT > 0 && dN > 0 && dT <= P.max_read_depth && dN <= P.max_read_depth && nrT != 0 && nrN <= P.max_normal_alt;
    if (candidate) {
      S.cand[atomicAdd(&S.n_cand, 1)] = (uint16_t)l;
    } else {
      guac_locus_summary sn, st;
      sn.depth = (uint16_t)min(dN, 65535);
      sn.reference_depth = (uint16_t)min(dN - nrN, 65535);
      sn.top_alt_count = 0;
      sn.flags = 0;
      st.depth = (uint16_t)min(dT, 65535);
      st.reference_depth = (uint16_t)min(dT - nrT, 65535);
      st.top_alt_count = 0;
      st.flags = 0;
      summary[slot * 2 + 0] = sn;
      summary[slot * 2 + 1] = st;
    }
  }
  __syncthreads();

  // ---- C. candidate loci, one warp each --------------------------------------------------------
  const int n_cand = S.n_cand;
  const int max_span = S.max_span;
  unsigned* hist = S.hist[warp];
  for (int ci = warp; ci < n_cand; ci += CFG::WARPS) {
    const int l = S.cand[ci];
    const int64_t x = t0 + l, slot = tile.slot0 + l;
    const int dN = S.diff[0][l], dT = S.diff[1][l];
    const int nrN = S.nonref[0][l], nrT = S.nonref[1][l];
    // most frequent tumor variant allele from the locus's events
    uint64_t tkey = 0;
    int tcnt = 0, ntab = 0;
    int64_t tfirst = 0;
    const int e0 = S.ev_off[l], e1 = S.ev_off[l + 1];
    for (int base = e0; base < e1; base += 32) {
      const int i = base + lane;
      const bool has = i < e1;
      const uint32_t ev = has ? S.ev_sorted[i] : 0;
      const int rl = (ev >> 11) & 0xFFF, eslot = (ev >> 23) & 1, kind = (ev >> 24) & 7, code = (ev >> 27) & 7;
      const bool var = has && eslot == 1 && kind != GUAC_ALN_CLIPPED;
      const int64_t r = tile.read_lo + rl;
      uint64_t key = 0;
      if (var) {
        if (kind == GUAC_ALN_MISMATCH && code < 4) {
          key = (1ull << 60) | (uint8_t)("ACGT"[code]);
        } else if (kind == GUAC_ALN_MID_DELETION) {
          key = 4ull << 60;
        } else {
          Elem e;
          element_at(R, r, x, ref, e);
          key = allele_key(R, r, e);
        }
      }
      unsigned vm = __ballot_sync(FULL, var);
      while (vm) {
        const int L = __ffs(vm) - 1;
        const uint64_t k = __shfl_sync(FULL, key, L);
        const unsigned same = __ballot_sync(FULL, var && key == k);
        const int c = __popc(same);
        long long mn = (var && key == k) ? (long long)r : LLONG_MAX;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) mn = min(mn, __shfl_xor_sync(FULL, mn, d));
        const unsigned hit = __ballot_sync(FULL, lane < ntab && tkey == k);
        if (hit) {
          if (lane == __ffs(hit) - 1) {
            tcnt += c;
            tfirst = min((long long)tfirst, mn);
          }
        } else if (ntab < 32) {
          if (lane == ntab) {
            tkey = k;
            tcnt = c;
            tfirst = mn;
          }
          ++ntab;
        } else if (lane == 0) {
          report_error(err, DERR_ALLELE_TABLE_FULL, tile.contig_id, x, mn);
        }
        vm &= ~same;
      }
    }
    int bcnt = lane < ntab ? tcnt : -1;
    int64_t bfirst = lane < ntab ? tfirst : INT64_MAX;
    uint64_t bkey = tkey;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
      const int oc = __shfl_xor_sync(FULL, bcnt, d);
      const int64_t of = __shfl_xor_sync(FULL, bfirst, d);
      const uint64_t ok = __shfl_xor_sync(FULL, bkey, d);
      if (oc > bcnt || (oc == bcnt && of < bfirst)) {
        bcnt = oc;
        bfirst = of;
        bkey = ok;
      }
    }
    if (ntab == 0) {  // every non-match tumor element is Clipped: maxBy on an empty map (:188)
      if (lane == 0) report_error(err, DERR_EMPTY_MAXBY, tile.contig_id, x, -1);
      continue;
    }
    const double f = fmax(P.min_vaf, (double)bcnt / (double)dT);  // :189

    // Exact pruning.  tumorAltLOD = sum over tumor elements of log(P(e|mix)) - log(P(e|ref)):
    //   element of the reference allele:  log(1 - f + f (1-p)/p)  <= 0 whenever p >= 1/2,
    //   element of the called allele:     log(1 - f + f p/(1-p)),
    //   any other element:                log(1) = 0.
    // When every tumor read over the locus has all base qualities >= 4 (p >= 0.60; tracked by
    // lowq[]), the sum over the called allele's elements — known from the events — bounds the
    // LOD from above, and a locus whose bound is below the threshold (minus a 0.5 margin that
    // dwarfs any rounding) cannot be called (:231), so its qualities are never gathered.
    double bsum = 0;
    int normalAltEv = 0;
    for (int base = e0; base < e1; base += 32) {
      const int i = base + lane;
      const bool has = i < e1;
      const uint32_t ev = has ? S.ev_sorted[i] : 0;
      const int rl = (ev >> 11) & 0xFFF, eslot = (ev >> 23) & 1, kind = (ev >> 24) & 7, code = (ev >> 27) & 7;
      const bool var = has && kind != GUAC_ALN_CLIPPED;
      uint64_t key = 0;
      int q = 0;
      if (var) {
        if (S.r_flags[rl] & 2) {  // fast read: a mismatch with an ACGT base
          key = (1ull << 60) | (uint8_t)("ACGT"[code]);
          if (eslot == 1 && key == bkey) q = __ldg(R.qual + seq_base + S.r_s0[rl] + (uint64_t)(l - S.r_start[rl]));
        } else {
          Elem e;
          const int64_t r = tile.read_lo + rl;
          if (element_at(R, r, x, ref, e) == DERR_NONE) {
            key = allele_key(R, r, e);
            q = e.quality;
          }
        }
      }
      const bool is_alt = var && key == bkey;
      normalAltEv += __popc(__ballot_sync(FULL, is_alt && eslot == 0));
      if (is_alt && eslot == 1) {
        const double p = c_success_probability[q];
        bsum += log(1.0 - f + f * (p / (1.0 - p)));
      }
    }
    bsum = warp_sum(bsum);
    if (S.lowq[l] == 0 && bsum < (double)P.tumor_odds - 0.5) {
      if (lane == 0) {
        guac_locus_summary sn, st;
        sn.depth = (uint16_t)min(dN, 65535);
        sn.reference_depth = (uint16_t)min(dN - nrN, 65535);
        sn.top_alt_count = (uint16_t)min(normalAltEv, 65535);
        sn.flags = 1;
        st.depth = (uint16_t)min(dT, 65535);
        st.reference_depth = (uint16_t)min(dT - nrT, 65535);
        st.top_alt_count = (uint16_t)min(bcnt, 65535);
        st.flags = 1;
        summary[slot * 2 + 0] = sn;
        summary[slot * 2 + 1] = st;
      }
      continue;
    }
    const bool best_snv = (bkey >> 60) == 1;
    const int best_code = best_snv ? base_code((uint8_t)(bkey & 0xFF)) : 5;
    const int ref_code = (S.ref2[l >> 4] >> (2 * (l & 15))) & 3;
    const bool ref_is_acgt = !((S.refmask[l >> 5] >> (l & 31)) & 1);

    // likelihood gather over the reads that can overlap the locus
    const int hi = S.starts[l];
    const int lo = l - max_span >= 0 ? S.starts[l - max_span] : 0;
    double aTR = 0, aNR = 0, aNA = 0, aTAx = 0;  // aTAx: tumor mixture terms of elements with q >= QB
    int normalAlt = 0;
    for (int base = lo; base < hi; base += 32) {
      const int i = base + lane;
      bool ov = false;
      int cls = 2, q = 0, eslot = 0;
      if (i < hi) {
        const int st = S.r_start[i], en = S.r_end[i];
        const int fl = S.r_flags[i];
        ov = st <= l && l < en && (fl & 4);
        if (ov) {
          eslot = fl & 1;
          if (fl & 2) {  // fast read: Match or Mismatch
            const uint64_t gi = seq_base + S.r_s0[i] + (uint64_t)(l - st);
            q = __ldg(R.qual + gi);
            const int code = (__ldg(R.bases2 + (gi >> 2)) >> ((gi & 3) * 2)) & 3;
            cls = (ref_is_acgt && code == ref_code) ? 0 : (code == best_code ? 1 : 2);
          } else {
            Elem e;
            const int64_t r = tile.read_lo + i;
            if (element_at(R, r, x, ref, e) == DERR_NONE) {
              q = e.quality;
              if (e.kind == GUAC_ALN_MATCH) cls = 0;
              else if (e.kind != GUAC_ALN_CLIPPED && allele_key(R, r, e) == bkey) cls = 1;
              else cls = 2;
            } else {
              ov = false;
            }
          }
        }
      }
      if (ov) {
        if (eslot == 1) {
          aTR += __ldg(&LT->ref[cls][q]);
          if (q < CFG::QB) {
            atomicAdd(&hist[cls * CFG::QB + q], 1u);
          } else {
            const double p = c_success_probability[q];
            const double pr = cls == 0 ? p : 1 - p, pa = cls == 1 ? p : 1 - p;
            aTAx += log(pr * (1.0 - f) + pa * f);
          }
        } else {
          aNR += __ldg(&LT->ref[cls][q]);
          aNA += __ldg(&LT->het[cls][q]);
        }
      }
      normalAlt += __popc(__ballot_sync(FULL, ov && eslot == 0 && cls == 1));
    }
    __syncwarp();
    // tumor mixture terms: compact the non-empty (class, quality) bins, then one log per bin
    double aTA = aTAx;
    {
      unsigned* clist = S.clist[warp];
      int nb = 0;
      for (int s = 0; s < 3 * CFG::QB; s += 32) {
        const unsigned n = hist[s + lane];
        const unsigned nz = __ballot_sync(FULL, n != 0);
        if (n) {
          hist[s + lane] = 0;
          clist[nb + __popc(nz & ((1u << lane) - 1))] = ((unsigned)(s + lane) << 24) | n;
        }
        nb += __popc(nz);
      }
      __syncwarp();
      for (int base = 0; base < nb; base += 32) {
        if (base + lane < nb) {
          const unsigned w = clist[base + lane];
          const int bin = (int)(w >> 24), n = (int)(w & 0xFFFFFF);
          const int cls = bin / CFG::QB, q = bin - cls * CFG::QB;
          const double p = c_success_probability[q];
          const double pr = cls == 0 ? p : 1 - p, pa = cls == 1 ? p : 1 - p;
          aTA += (double)n * log(pr * (1.0 - f) + pa * f);
        }
      }
      __syncwarp();
    }
    aTR = warp_sum(aTR);
    aTA = warp_sum(aTA);
    aNR = warp_sum(aNR);
    aNA = warp_sum(aNA);
    double tR, tA, nR, nA;
    log_normalize2(aTR, aTA, tR, tA);
    log_normalize2(aNR, aNA, nR, nA);
    const double tumorAltLOD = tA - tR, normalRefLOD = nR - nA;
    const bool called = tumorAltLOD > P.tumor_odds && normalRefLOD > P.normal_odds && (bkey >> 60) != 4;  // :231-233
    if (lane == 0) {
      if (called) {
        const int idx = atomicAdd(n_calls, 1);
        if (idx < max_calls) {
          SSCallDev c;
          c.slot = slot;
          c.locus = x;
          c.read_lo = tile.read_lo;
          c.read_hi = tile.read_hi;
          c.key = bkey;
          c.rep_read = bfirst;
          c.contig_id = tile.contig_id;
          c.ref_index = tile.ref_index;
          c.tumor_ref_lp = tR;
          c.tumor_alt_lp = tA;
          c.normal_ref_lp = nR;
          c.normal_alt_lp = nA;
          c.ref_off = c.ref_len = c.alt_off = c.alt_len = 0;
          calls[idx] = c;
        }
      }
      const uint16_t fl = 1 | (called ? 2 : 0);
      guac_locus_summary sn, st;
      sn.depth = (uint16_t)min(dN, 65535);
      sn.reference_depth = (uint16_t)min(dN - nrN, 65535);
      sn.top_alt_count = (uint16_t)min(normalAlt, 65535);
      sn.flags = fl;
      st.depth = (uint16_t)min(dT, 65535);
      st.reference_depth = (uint16_t)min(dT - nrT, 65535);
      st.top_alt_count = (uint16_t)min(bcnt, 65535);
      st.flags = fl;
      summary[slot * 2 + 0] = sn;
      summary[slot * 2 + 1] = st;
    }
  }
}

}  // namespace guac
