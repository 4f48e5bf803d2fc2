// stream_kernels.cuh — read-streaming front half of the tile path ("stream" implementation of somatic-standard).
//
// The tile kernel (tile_kernels.cuh) compares every read with the reference once per tile it overlaps, inside the
// CTA's chain of barriers.  Here that comparison is a kernel of its own: one thread per read, every read visited
// exactly once, no shared memory, no barrier and no atomic.  A read compares its aligned segments, 16 bases per step,
// with a 2-bit copy of the reference kept in HBM (12.5 MB for a 50 Mbp contig: L2-resident) and leaves its non-Match
// elements in its own slot of a per-read event table.  The locus pass (ss_tile_kernel<..., STREAM = true>) then only
// reads the headers and the event slots of the tile's reads.
#pragma once
#include "tile_kernels.cuh"

namespace guac {

// ASCII reference window -> 2 bits per locus, 16 loci per word; anything but A C G T packs as 0 (tiles holding such a
// locus are handed to the generic kernel by the locus pass, so the value never matters).
__global__ void ref2_pack_kernel(const uint8_t* __restrict__ bases, int64_t len, uint32_t* __restrict__ ref2, uint32_t* __restrict__ bad) {
  const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (16 * w >= len) return;
  uint32_t v = 0;
  bool any_bad = false;
  for (int k = 0; k < 16 && 16 * w + k < len; ++k) {
    const uint8_t b = bases[16 * w + k];
    const uint32_t d = (uint32_t)b - 65u;
    if (d < 20u && ((0x80045u >> d) & 1u)) v |= (((uint32_t)(b >> 1) ^ (uint32_t)(b >> 2)) & 3u) << (2 * k);
    else any_bad = true;
  }
  ref2[w] = v;
  if (any_bad) atomicOr(bad + (w >> 5), 1u << (w & 31));  // one bit per 16-locus group
}

// Reads with several CIGAR operators (RF_MEDIUM but not RF_FAST), listed once per upload so that their longer walk runs
// in full warps of its own instead of one lane in many warps of the single-operator reads.
__global__ void medium_reads_flag_kernel(const uint8_t* __restrict__ rflags, int64_t n, uint8_t* __restrict__ flag) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= n) return;
  const uint8_t rf = rflags[r];
  flag[r] = ((rf & RF_MEDIUM) && !(rf & RF_FAST)) ? 1 : 0;
}
// flag + exclusive scan -> the list in ascending read order (neighbouring list entries touch neighbouring memory)
__global__ void medium_reads_scatter_kernel(const uint8_t* __restrict__ flag, const int64_t* __restrict__ offset, int64_t n,
                                            int64_t* __restrict__ list) {
  const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r < n && flag[r]) list[offset[r]] = r;
}

constexpr int STREAM_EV_CAP = STREAM_EV_CAP_C;  // event records kept per read; a read with more is walked by the exact cursor in the locus pass

struct StreamArgs {
  const uint4* hdr;     // packed read headers
  int64_t read_lo, read_hi;  // the reads of one contig
  RefView ref;          // that contig's reference window
  const uint32_t* ref2; // and its 2-bit copy
  uint2* events;        // [n_reads][STREAM_EV_CAP]
  uint8_t* n_events;    // [n_reads] (255: more than the cap)
  int inline_medium;    // 1: no list is given (several contigs in the batch): walk multi-operator reads in place
  // The multi-operator reads of the batch, listed at upload.  Their longer, latency-bound walk runs in CTAs of its own
  // (full warps of such reads), one of them after every `list_period - 1` CTAs of single-operator reads, so that the
  // two kinds share every SM all through the launch.
  const int64_t* list;
  int64_t n_list;
  unsigned list_blocks, list_period;  // list_blocks = 0: no list
  // Overflow arena (see STREAM_EV_OVERFLOW in tile_kernels.cuh): records of STREAM_OVF_CAP events, handed out by an atomic
  // counter (which record a read gets differs from run to run; what it holds does not).  Null: none.
  uint2* overflow;
  int* n_overflow;
  int max_overflow;
  int64_t* overflow_reads;  // [max_overflow] the read of each arena record (ss_read_overflow_kernel fills the records)
};

// Read-event record (per read, independent of the loci of the call):
//   x = reference offset from the read's start (16 bits) | kind << 16 | code << 19 (codes as in the tile kernel's events)
//   y = read position of the mismatching base, or bit 31 | detail << 8 | qualityScore for the other kinds
__device__ __forceinline__ uint32_t pack_read_event(int off, int kind, int code) {
  return (uint32_t)off | ((uint32_t)kind << 16) | ((uint32_t)code << 19);
}

// The walk of one read: its non-Match elements against the 2-bit reference, the first `cap` of them written to `out`.
// Returns their number (4096: an element the record cannot express).
__device__ __forceinline__ int read_events_walk(const ReadsView& R, const StreamArgs& A, int64_t r, const uint4 hdr, uint2* out, int cap) {
  const uint8_t rf = (uint8_t)(hdr.y >> 24);
  int n_ev = 0;
  auto push = [&](int off, int kind, int code, uint32_t second) {
    if (n_ev < cap && off < 65536) out[n_ev] = make_uint2(pack_read_event(off, kind, code), second);
    else if (off >= 65536) n_ev = 4096;  // not representable: treated like an overflow
    ++n_ev;
  };
  auto direct = [](int q, int detail) { return 0x80000000u | ((uint32_t)(detail & 0xFF) << 8) | (uint32_t)(q & 0xFF); };
  {
    const RefView& ref = A.ref;
    const uint32_t* ref2 = A.ref2;
    const uint64_t s0 = (uint64_t)hdr.z | ((uint64_t)hdr.w << 32);
    // reads and windows lie within 2^31 loci of each other on a contig: 32-bit wrap-around arithmetic is exact
    const int span = (hdr.y & 0xFFFF) == 0xFFFF ? (int)min((long long)(__ldg(R.end + r) - __ldg(R.start + r)), (1ll << 30)) : (int)(hdr.y & 0xFFFF);
    if (rf & RF_MEDIUM) {  // the others need the exact cursor: the locus pass handles them, one warp per read
      const long long ref_rel = (long long)(int)(hdr.x - (uint32_t)ref.offset);  // reference-window index of the read's first locus
      // aligned segment: loci [lb, le) (offsets from the read's start) hold read positions rp0..
      auto compare_segment = [&](int lb, int le, int rp0) {
        const int ca = (int)max((long long)lb, -ref_rel), cb = (int)min((long long)le, (long long)ref.len - ref_rel);  // inside the window
        if (ca >= cb) return;
        const uint64_t rbit = 2 * (s0 + (uint64_t)(rp0 + (ca - lb)));
        const uint32_t* rw = reinterpret_cast<const uint32_t*>(R.bases2) + (rbit >> 5);
        const uint64_t fbit = 2 * (uint64_t)(ref_rel + ca);
        const uint32_t* fw = ref2 + (fbit >> 5);
        const int rsh = (int)(rbit & 31), fsh = (int)(fbit & 31);
        uint32_t rlo = __ldg(rw), flo = __ldg(fw);
        const int q_rel = rp0 - lb;  // read position of offset l = q_rel + l
#ifndef GUAC_K1_PAIR
#define GUAC_K1_PAIR 1
#endif
        // one 16-base word pair: the mismatch mask (bit 2k: base k differs)
        auto diff = [&](uint32_t& rd) {
          const uint32_t rhi = __ldg(++rw), fhi = __ldg(++fw);
          rd = __funnelshift_r(rlo, rhi, rsh);
          const uint32_t x = rd ^ __funnelshift_r(flo, fhi, fsh);
          rlo = rhi;
          flo = fhi;
          return (x | (x >> 1)) & 0x55555555u;
        };
        int l = ca;
#if GUAC_K1_PAIR
        // Two steps at a time: the second step's mask rides in the odd bits, so the record path below (entered by a warp
        // whenever any of its lanes has a mismatch) is entered once per 32 bases instead of once per 16.
        for (; l + 32 <= cb; l += 32) {
          uint32_t rd0, rd1;
          uint32_t m = diff(rd0);
          m |= diff(rd1) << 1;
          while (m) {
            const int bit = __ffs(m) - 1;
            m &= m - 1;
            const int off = l + ((bit & 1) << 4) + (bit >> 1);
            push(off, GUAC_ALN_MISMATCH, (((bit & 1) ? rd1 : rd0) >> (bit & 30)) & 3, (uint32_t)(q_rel + off));
          }
        }
#endif
        for (; l < cb; l += 16) {
          uint32_t rd;
          uint32_t m = diff(rd);
          if (cb - l < 16) m &= (1u << (2 * (cb - l))) - 1;
          while (m) {
            const int bit = __ffs(m) - 1;
            m &= m - 1;
            push(l + (bit >> 1), GUAC_ALN_MISMATCH, (rd >> bit) & 3, (uint32_t)(q_rel + l + (bit >> 1)));
          }
        }
      };
      if (rf & RF_FAST) {
        compare_segment(0, span, 0);
      } else {  // operator walk (PE.scala:59-117 per operator instead of per base)
        const uint32_t c0 = __ldg(R.cigar_off + r), c1 = __ldg(R.cigar_off + r + 1);
        const int read_len = (int)(__ldg(R.seq_off + r + 1) - s0);
        int cl = 0, rp = 0;
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
            const bool special_last = nop == OPC_D || (nop == OPC_I && op != OPC_X);
            compare_segment(cl, cl + len - (special_last ? 1 : 0), rp);
            if (special_last) {
              const int ll = cl + len - 1, k = rp + len - 1;
              if (nop == OPC_I) {  // qualityScore = min over anchor + inserted qualities (PE.scala:153)
                const int until = min(k + nlen + 1, read_len);
                int q = 255;
                uint32_t h = 2166136261u ^ (uint32_t)(until - k);
                for (int j = k; j < until; ++j) {
                  q = min(q, (int)__ldg(R.qual + s0 + j));
                  const uint64_t gi = s0 + (uint64_t)j;
                  h = (h ^ ((__ldg(R.bases2 + (gi >> 2)) >> ((gi & 3) * 2)) & 3)) * 16777619u;
                }
                push(ll, GUAC_ALN_INSERTION, 4, direct(q, (int)((h ^ (h >> 8) ^ (h >> 16) ^ (h >> 24)) & 0xFF)));
              } else {
                // (the locus pass checks that the deleted reference bases exist, PE.scala:95, when the anchor is inside a tile)
                push(ll, GUAC_ALN_DELETION, nlen < 255 ? 4 : 5, direct(__ldg(R.qual + s0 + k), nlen));
              }
            }
            cl += len;
            rp += len;
          } else if (op == OPC_D) {
            // one record per run of deleted loci: bits 16..30 of y carry the run length
            const int mq = __ldg(R.mapq + r);
            for (int done = 0; done < len; done += 0x7FFF)
              push(cl + done, GUAC_ALN_MID_DELETION, 4, direct(mq, 0) | ((uint32_t)min(len - done, 0x7FFF) << 16));
            cl += len;
          } else if (op == OPC_I || op == OPC_S) {
            rp += len;
          }
        }
      }
    }
  }
  return n_ev;
}

// One thread per read: the read's non-Match elements against the 2-bit reference.  No shared memory, no barrier, no atomic.
// 8 CTAs per SM (32 registers, a dozen bytes of spills): the kernel lives on loads in flight, and the two extra CTAs are
// worth more than the spills cost (measured 2.44 -> 2.36 ms for the step; 7 CTAs 2.37 ms).
#ifndef GUAC_K1_MINB
#define GUAC_K1_MINB 8
#endif
__global__ void __launch_bounds__(256, GUAC_K1_MINB) ss_read_events_kernel(const __grid_constant__ ReadsView R, const __grid_constant__ StreamArgs A) {
  // block -> a block of the list or a block of the read range (see StreamArgs)
  unsigned b = blockIdx.x;
  bool from_list = false;
  if (A.list_blocks) {
    const unsigned g = b / A.list_period, k = b % A.list_period;
    if (g < A.list_blocks) {
      from_list = k == A.list_period - 1;
      b = from_list ? g : g * (A.list_period - 1) + k;
    } else {
      b -= A.list_blocks;
    }
  }
  int64_t r;
  if (from_list) {
    const int64_t i = (int64_t)b * blockDim.x + threadIdx.x;
    if (i >= A.n_list) return;
    r = __ldg(A.list + i);
  } else {
    r = A.read_lo + (int64_t)b * blockDim.x + threadIdx.x;
    if (r >= A.read_hi) return;
  }
  const uint4 hdr = __ldg(A.hdr + r);
  const uint8_t rf = (uint8_t)(hdr.y >> 24);
  if (!from_list && !A.inline_medium && (rf & RF_MEDIUM) && !(rf & RF_FAST)) return;  // visited through the list
  uint2* mine = A.events + r * STREAM_EV_CAP;
  int n_ev = read_events_walk(R, A, r, hdr, mine, STREAM_EV_CAP);
#ifndef GUAC_K1_PAD
#define GUAC_K1_PAD 1
#endif
#if GUAC_K1_PAD
  // a 32-byte sector of the slot that holds an event is written whole (zeros behind the last record): L2 then owns
  // every byte of the sector and evicts it without first fetching the rest from HBM
  if (n_ev < STREAM_EV_CAP)
    for (int k = n_ev; k & 3; ++k) mine[k] = make_uint2(0u, 0u);
#endif
  if (n_ev > STREAM_EV_CAP) {
    // more events than the slot holds: the read gets a record of the overflow arena, filled by ss_read_overflow_kernel —
    // a second walk, with every lane of its warps on such a read instead of one lane here and 31 waiting
    int ovf = INT_MAX;
    if (n_ev <= STREAM_OVF_CAP && A.overflow != nullptr) ovf = atomicAdd(A.n_overflow, 1);
    if (ovf < A.max_overflow) {
      A.overflow_reads[ovf] = r;
      mine[0] = make_uint2((uint32_t)ovf, (uint32_t)n_ev);
      n_ev = STREAM_EV_OVERFLOW;
    } else {
      n_ev = 255;
    }
  }
  A.n_events[r] = (uint8_t)n_ev;
}

// The arena records of the reads ss_read_events_kernel listed: one thread per record, the read walked again with room for
// all its events.  The launch is sized before the number of records is known and strides over them.
__global__ void __launch_bounds__(128) ss_read_overflow_kernel(const __grid_constant__ ReadsView R, const __grid_constant__ StreamArgs A) {
  const int n = min(*A.n_overflow, A.max_overflow);
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
    const int64_t r = A.overflow_reads[t];
    // (a batch of several contigs has one launch pair per contig, each with that contig's reference, and one arena: the
    // records of the other contigs' reads belong to their launches)
    if (A.list_blocks == 0 && (r < A.read_lo || r >= A.read_hi)) continue;
    read_events_walk(R, A, r, __ldg(A.hdr + r), A.overflow + (size_t)t * STREAM_OVF_CAP, STREAM_OVF_CAP);
  }
}

}  // namespace guac
