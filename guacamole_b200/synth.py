"""Seeded synthetic tumor/normal reads (SURVEY.md §8(d) generator), host (numpy) flavour.

The generator is *stateless*: every quantity of read j is a pure function of
(seed, j, base index) through a 64-bit mixing hash, so any window of reads can be produced
independently — on the host here, and bit-identically by the CUDA twin in
csrc/synth_kernels.cuh, which fills HBM directly for the full-size benchmark.  The host
flavour feeds the parity tests (and the CPU oracle) with exactly the reads the device has.

Model: reference iid uniform ACGT; read length 100; reads evenly spread with jitter so that
the batch is sorted by start; samples interleaved in proportion to their depths;
mapq 60/20/0 at 97/2/1 %; qualities {40,30,20,10} at 70/20/8/2 %; substitution errors at
10^(-q/10); CIGARs 100M 97 %, aMbIcM / aMbDcM / sS(100-s)M 1 % each; germline SNVs 1 per kbp
(2/3 het), somatic SNVs 1 per 10 kbp in tumor samples at VAF 0.10 / 0.25 / 0.50.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np

from .io import ContigWindow, Reference
from .reads import FLAG_PAIRED, FLAG_POSITIVE_STRAND, ReadBatch

READ_LEN = 100
MARGIN = 120  # keeps read ends (incl. deletions) inside the contig
U64 = np.uint64
_M1, _M2 = U64(0xBF58476D1CE4E5B9), U64(0x94D049BB133111EB)
_G1, _G2 = U64(0x9E3779B97F4A7C15), U64(0xC2B2AE3D27D4EB4F)

# streams
S_REF, S_START, S_READ, S_BASE, S_QUAL, S_ERR, S_SITE, S_INS, S_DENS, S_TAIL = (U64(i) for i in range(1, 11))
# error thresholds on 24 bits: round(10^(-q/10) * 2^24)
ERR_THR = {40: 1678, 30: 16777, 20: 167772, 10: 1677722}


def mix(z):
    with np.errstate(over="ignore"):
        z = np.asarray(z, dtype=U64)
        z = (z ^ (z >> U64(30))) * _M1
        z = (z ^ (z >> U64(27))) * _M2
        return z ^ (z >> U64(31))


def rnd(seed: int, stream, a, b=0):
    """64 random bits as a pure function of (seed, stream, a, b)."""
    with np.errstate(over="ignore"):
        key = mix(U64(seed) ^ (U64(stream) << U64(32)))
        return mix(mix(key ^ (np.asarray(a, dtype=U64) * _G1)) ^ (np.asarray(b, dtype=U64) * _G2))


def reference_bases(seed: int, lo: int, hi: int) -> np.ndarray:
    """ASCII reference for loci [lo, hi)."""
    codes = (rnd(seed, S_REF, np.arange(lo, hi, dtype=U64)) & U64(3)).astype(np.uint8)
    return np.frombuffer(b"ACGT", np.uint8)[codes]


def sample_pattern(depths: Sequence[int]) -> np.ndarray:
    """Interleaving pattern of sample ids proportional to the depths (e.g. [30, 60] -> [1, 1, 0])."""
    g = int(np.gcd.reduce(np.asarray(depths)))
    counts = [d // g for d in depths]
    pat = []
    # round-robin, largest first, so that samples are spread inside the period
    remaining = counts[:]
    while sum(remaining):
        for s in sorted(range(len(depths)), key=lambda s: -counts[s]):
            if remaining[s]:
                pat.append(s)
                remaining[s] -= 1
    return np.asarray(pat, np.uint8)


def n_reads_total(contig_len: int, depths: Sequence[int]) -> int:
    return int(sum(depths)) * contig_len // READ_LEN


def site_info(seed: int, pos: np.ndarray):
    """Planted variant sites.  Returns (germline kind 0/1 het/2 hom, germline alt offset 1..3,
    somatic vaf threshold on 16 bits or 0, somatic alt offset 1..3)."""
    h = rnd(seed, S_SITE, pos.astype(U64))
    germ = (h % U64(1000)) == U64(0)
    gkind = np.where(germ, np.where(((h >> U64(20)) % U64(3)) == U64(2), 2, 1), 0).astype(np.uint8)
    galt = (((h >> U64(24)) % U64(3)) + U64(1)).astype(np.uint8)
    som = ((h >> U64(32)) % U64(10000)) == U64(0)
    vsel = ((h >> U64(48)) % U64(3)).astype(np.int64)
    vthr = np.where(som, np.asarray([6554, 16384, 32768], np.int64)[vsel], 0)  # 0.10 0.25 0.50 of 2^16
    salt = (((h >> U64(52)) % U64(3)) + U64(1)).astype(np.uint8)
    return gkind, galt, vthr, salt


def generate_reads(contig_len: int, depths: Sequence[int], tumor: Sequence[bool], seed: int, read_lo: int = 0,
                   read_hi: int = None, contig_name: str = "synth1", profile: int = 0) -> ReadBatch:
    """Reads [read_lo, read_hi) of the synthetic batch (all samples interleaved, sorted by start).

    profile 0: the workload of SURVEY.md 8(d).  profile 1: the "realistic" variant — read density modulated in 256-locus
    blocks (depth varies like a Poisson pileup's), 15 % of reads end in 5..30 quality-2 bases with 10 % errors, 10 %
    multi-operator reads (4 % insertions, 4 % deletions, 2 % soft clips), 1 % of reads with 15 % errors throughout."""
    N = n_reads_total(contig_len, depths)
    read_hi = N if read_hi is None else min(read_hi, N)
    read_lo = max(0, read_lo)
    n = max(0, read_hi - read_lo)
    j = np.arange(read_lo, read_hi, dtype=np.int64)
    ju = j.astype(U64)
    pat = sample_pattern(depths)
    sample = pat[j % len(pat)]
    span = contig_len - MARGIN
    # start = floor((j + u) * span / N), u in [0,1) on 16 bits -> monotone in j
    u16 = (rnd(seed, S_START, ju) & U64(0xFFFF)).astype(np.int64)
    js = j * span  # floor((j * 65536 + u16) * span / (N * 65536)) without int64 overflow
    start = js // N + ((js % N) * 65536 + u16 * span) // (N * 65536)
    if profile == 1:  # density modulation inside 256-locus blocks, monotone in the start
        blk = start >> 8
        t = start & 255
        cut = 128 + (rnd(seed, S_DENS, blk.astype(U64)) % U64(61)).astype(np.int64) - 30
        remapped = (blk << 8) + np.where(t < cut, t * 128 // cut, 128 + (t - cut) * 128 // (256 - cut))
        start = np.where(((blk + 1) << 8) <= span, remapped, start)
    hr = rnd(seed, S_READ, ju)
    positive = (hr & U64(1)).astype(bool)
    hap = ((hr >> U64(1)) & U64(1)).astype(np.uint8)
    mq = ((hr >> U64(8)) % U64(100)).astype(np.int64)
    mapq = np.where(mq < 97, 60, np.where(mq < 99, 20, 0)).astype(np.uint8)
    ck = ((hr >> U64(16)) % U64(100)).astype(np.int64)  # 0..96 plain, 97 ins, 98 del, 99 softclip
    b_len = ((hr >> U64(24)) % U64(10)).astype(np.int64) + 1  # indel length 1..10
    # a in [5, 100 - b - 5] for insertion (a + b + c = 100), a in [5, 95] for deletion (a + c = 100)
    a_ins = 5 + ((hr >> U64(32)) % (U64(91) - b_len.astype(U64))).astype(np.int64)
    a_del = 5 + ((hr >> U64(32)) % U64(91)).astype(np.int64)
    s_clip = ((hr >> U64(40)) % U64(20)).astype(np.int64) + 1
    is_ins, is_del, is_clip = ck == 97, ck == 98, ck == 99
    tail = np.zeros(n, np.int64)
    noisy = np.zeros(n, bool)
    if profile == 1:
        is_ins, is_del, is_clip = (ck >= 90) & (ck < 94), (ck >= 94) & (ck < 98), ck >= 98
        noisy = ((hr >> U64(48)) % U64(100)) == U64(0)
        ht = rnd(seed, S_TAIL, ju)
        tail = np.where(((ht >> U64(32)) % U64(100)) < U64(15), 5 + (ht % U64(26)).astype(np.int64), 0)

    # CIGAR words
    n_ops = np.where(is_ins | is_del, 3, np.where(is_clip, 2, 1))
    cigar_off = np.zeros(n + 1, np.uint32)
    cigar_off[1:] = np.cumsum(n_ops)
    cigar = np.zeros(int(cigar_off[-1]), np.uint32)
    o = cigar_off[:-1].astype(np.int64)
    plain = ~(is_ins | is_del | is_clip)
    cigar[o[plain]] = (READ_LEN << 4) | 0
    cigar[o[is_ins]] = (a_ins[is_ins] << 4) | 0
    cigar[o[is_ins] + 1] = (b_len[is_ins] << 4) | 1
    cigar[o[is_ins] + 2] = ((READ_LEN - a_ins[is_ins] - b_len[is_ins]) << 4) | 0
    cigar[o[is_del]] = (a_del[is_del] << 4) | 0
    cigar[o[is_del] + 1] = (b_len[is_del] << 4) | 2
    cigar[o[is_del] + 2] = ((READ_LEN - a_del[is_del]) << 4) | 0
    cigar[o[is_clip]] = (s_clip[is_clip] << 4) | 4
    cigar[o[is_clip] + 1] = ((READ_LEN - s_clip[is_clip]) << 4) | 0

    # per-base reference position (or -1 for inserted / clipped bases)
    k = np.arange(READ_LEN, dtype=np.int64)[None, :]
    st = start[:, None]
    pos = st + k
    a_i, b_i = a_ins[:, None], b_len[:, None]
    pos = np.where(is_ins[:, None], np.where(k < a_i, st + k, np.where(k < a_i + b_i, -1, st + k - b_i)), pos)
    pos = np.where(is_del[:, None], np.where(k < a_del[:, None], st + k, st + k + b_i), pos)
    pos = np.where(is_clip[:, None], np.where(k < s_clip[:, None], -1, st + k - s_clip[:, None]), pos)
    aligned = pos >= 0
    safe = np.where(aligned, pos, 0)
    gi = ju[:, None] * U64(READ_LEN) + k.astype(U64)  # global base index
    ref_code = (rnd(seed, S_REF, safe.astype(U64)) & U64(3)).astype(np.uint8)
    gkind, galt, vthr, salt = site_info(seed, safe)
    is_tumor = np.asarray(tumor, bool)[sample][:, None]
    code = ref_code.copy()
    germ_alt = (gkind == 2) | ((gkind == 1) & (hap[:, None] == 1))
    code = np.where(germ_alt, (ref_code + galt) & 3, code)
    som_draw = (rnd(seed, S_SITE, ju[:, None], safe.astype(U64) + U64(1)) & U64(0xFFFF)).astype(np.int64)
    som_alt = is_tumor & (vthr > 0) & (som_draw < vthr)
    code = np.where(som_alt, (ref_code + salt) & 3, code)
    rand_code = (rnd(seed, S_INS, gi) & U64(3)).astype(np.uint8)
    code = np.where(aligned, code, rand_code)
    # qualities and errors
    hq = (rnd(seed, S_QUAL, gi) % U64(100)).astype(np.int64)
    qual = np.where(hq < 70, 40, np.where(hq < 90, 30, np.where(hq < 98, 20, 10))).astype(np.uint8)
    he = rnd(seed, S_ERR, gi)
    thr = np.where(qual == 40, ERR_THR[40], np.where(qual == 30, ERR_THR[30], np.where(qual == 20, ERR_THR[20], ERR_THR[10])))
    in_tail = k >= (READ_LEN - tail)[:, None]
    qual = np.where(in_tail, 2, qual).astype(np.uint8)
    thr = np.where(in_tail, ERR_THR[10], thr)          # the quality-2 tail: 10 % errors
    thr = np.where(noisy[:, None], np.maximum(thr, 2516582), thr)  # 15 % errors
    is_err = (he & U64(0xFFFFFF)).astype(np.int64) < thr
    err_off = (((he >> U64(24)) % U64(3)) + U64(1)).astype(np.uint8)
    code = np.where(is_err, (code + err_off) & 3, code).astype(np.uint8)
    bases = np.frombuffer(b"ACGT", np.uint8)[code]

    flags = (np.where(positive, FLAG_POSITIVE_STRAND, 0) | FLAG_PAIRED).astype(np.uint8)
    seq_off = (np.arange(n + 1, dtype=np.uint64) * np.uint64(READ_LEN))
    return ReadBatch(n_samples=len(depths), contig_names=[contig_name], contig_id=np.zeros(n, np.int32), start=start,
                     cigar_off=cigar_off, cigar=cigar, seq_off=seq_off, bases_ascii=np.ascontiguousarray(bases.reshape(-1)),
                     qual=np.ascontiguousarray(qual.reshape(-1)), mapq=mapq, flags=flags, sample_id=sample.astype(np.uint8))


def read_window_for_loci(contig_len: int, depths: Sequence[int], lo: int, hi: int) -> Tuple[int, int]:
    """Index range of reads that can overlap loci [lo, hi) (halo included)."""
    N = n_reads_total(contig_len, depths)
    span = contig_len - MARGIN
    # (profile 1 moves a start by less than a 256-locus block: the window is widened by one block on either side)
    j_lo = max(0, (max(0, lo - READ_LEN - 16 - 256) * N) // span - 2)
    j_hi = min(N, ((hi + 256) * N) // span + 2)
    return int(j_lo), int(j_hi)


def make_reference(contig_len: int, seed: int, contig_name: str = "synth1") -> Reference:
    return Reference({contig_name: ContigWindow(contig_name, 0, reference_bases(seed, 0, contig_len), contig_len)})


def make_tumor_normal(contig_len: int = 20000, seed: int = 3000, depths=(30, 60)) -> Tuple[Reference, ReadBatch]:
    """BASELINE config #3 shape (somatic-standard, normal 30x = sample 0, tumor 60x = sample 1)."""
    return make_reference(contig_len, seed), generate_reads(contig_len, depths, (False, True), seed)
