"""ctypes wrapper of oracle/libguac_oracle.so — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module.  It is the parity checker (a CPU restatement of the reference,
see guac_oracle.hpp), never the thing measured as the product or shipped.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Optional, Sequence, Tuple

import numpy as np

from guacamole_b200.io import Reference
from guacamole_b200.native import (ELEMENT_DTYPE, LOCUS_COUNTS_DTYPE, LOCUS_SUMMARY_DTYPE, SS_CALL_DTYPE, ElementC,
                                   LocusCountsC, LocusSummaryC, SSCallC, SSParamsC, i32p, i64p, make_ranges, ptr,
                                   ss_params, u8p, u32p, u64p)
from guacamole_b200.reads import ReadBatch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libguac_oracle.so")


class OracleReads(C.Structure):
    _fields_ = [("n_reads", C.c_int64), ("n_samples", C.c_int32), ("contig_id", i32p), ("start", i64p), ("cigar_off", u32p),
                ("cigar", u32p), ("seq_off", u64p), ("bases_ascii", u8p), ("qual", u8p), ("mapq", u8p), ("flags", u8p),
                ("sample_id", u8p)]


class OracleContig(C.Structure):
    _fields_ = [("contig_id", C.c_int32), ("_pad", C.c_int32), ("offset", C.c_int64), ("len", C.c_int64), ("bases", u8p)]


def build(force: bool = False) -> str:
    if force or not os.path.exists(LIB_PATH) or any(
            os.path.getmtime(os.path.join(_HERE, f)) > os.path.getmtime(LIB_PATH)
            for f in ("guac_oracle.cpp", "guac_oracle.hpp", "guac_oracle_joint.hpp")):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(LIB_PATH)
        _lib.oracle_last_error.restype = C.c_char_p
        _lib.oracle_free.argtypes = [C.c_void_p]
        _lib.oracle_phred_to_success_probability.restype = C.c_double
        _lib.oracle_phred_to_success_probability.argtypes = [C.c_int]
        _lib.oracle_success_probability_to_phred.argtypes = [C.c_double]
    return _lib


class OracleError(RuntimeError):
    pass


def _check(st):
    if st != 0:
        raise OracleError(lib().oracle_last_error().decode())


def _take(p, dtype, n):
    if n == 0 or not p:
        out = np.zeros(0, dtype)
    else:
        out = np.frombuffer(C.string_at(C.cast(p, C.c_void_p), int(n) * np.dtype(dtype).itemsize), dtype=dtype).copy()
    if p:
        lib().oracle_free(C.cast(p, C.c_void_p))
    return out


def _reads(batch: ReadBatch) -> OracleReads:
    b = OracleReads()
    b.n_reads, b.n_samples = batch.n_reads, batch.n_samples
    b.contig_id, b.start = ptr(batch.contig_id, i32p), ptr(batch.start, i64p)
    b.cigar_off, b.cigar, b.seq_off = ptr(batch.cigar_off, u32p), ptr(batch.cigar, u32p), ptr(batch.seq_off, u64p)
    b.bases_ascii, b.qual = ptr(batch.bases_ascii, u8p), ptr(batch.qual, u8p)
    b.mapq, b.flags, b.sample_id = ptr(batch.mapq, u8p), ptr(batch.flags, u8p), ptr(batch.sample_id, u8p)
    return b


def _contigs(reference: Reference, contig_names: Sequence[str]):
    present = [(i, reference.contigs[n]) for i, n in enumerate(contig_names) if n in reference.contigs]
    arr = (OracleContig * max(1, len(present)))()
    for k, (cid, w) in enumerate(present):
        arr[k].contig_id, arr[k].offset, arr[k].len, arr[k].bases = cid, w.offset, w.bases.shape[0], ptr(w.bases, u8p)
    return arr, len(present)


def pileup_counts(batch: ReadBatch, reference: Reference, ranges, skip_empty=True):
    cs, nc = _contigs(reference, batch.contig_names)
    b = _reads(batch)
    r = make_ranges(ranges)
    oc, ol, cnt, n = i32p(), i64p(), C.POINTER(LocusCountsC)(), C.c_int64()
    _check(lib().oracle_pileup_counts(C.byref(b), cs, nc, r, len(ranges), int(skip_empty), C.byref(oc), C.byref(ol),
                                      C.byref(cnt), C.byref(n)))
    return (_take(oc, np.int32, n.value), _take(ol, np.int64, n.value),
            _take(cnt, LOCUS_COUNTS_DTYPE, n.value * batch.n_samples).reshape(n.value, batch.n_samples))


def pileup_elements(batch: ReadBatch, reference: Reference, ranges, skip_empty=True):
    cs, nc = _contigs(reference, batch.contig_names)
    b = _reads(batch)
    r = make_ranges(ranges)
    oc, ol, off, els, order, n = i32p(), i64p(), i64p(), C.POINTER(ElementC)(), i64p(), C.c_int64()
    _check(lib().oracle_pileup_elements(C.byref(b), cs, nc, r, len(ranges), int(skip_empty), C.byref(oc), C.byref(ol),
                                        C.byref(off), C.byref(els), C.byref(order), C.byref(n)))
    offsets = _take(off, np.int64, n.value + 1)
    total = int(offsets[-1])
    return (_take(oc, np.int32, n.value), _take(ol, np.int64, n.value), offsets, _take(els, ELEMENT_DTYPE, total),
            _take(order, np.int64, total))


def pileup_alleles(batch: ReadBatch, reference: Reference, contig: str, locus: int, sample: int = 0):
    """[(ref, alt, kind, qualityScore)] of every element of `sample` at `locus`, in batch order."""
    cs, nc = _contigs(reference, batch.contig_names)
    b = _reads(batch)
    text = C.c_void_p()
    _check(lib().oracle_pileup_alleles(C.byref(b), cs, nc, batch.contig_names.index(contig), C.c_int64(locus), sample,
                                       C.byref(text)))
    s = C.string_at(text).decode()
    lib().oracle_free(text)
    out = []
    for line in s.split("\n"):
        if line:
            ref, alt, kind, q = line.split("\t")
            out.append((ref, alt, int(kind), int(q)))
    return out


def likelihoods_of_genotypes(batch, reference, contig, locus, genotypes, sample=0, include_alignment=False,
                             log_space=False, normalize=False):
    """genotypes: list of lists of ((ref, alt), fraction)."""
    cs, nc = _contigs(reference, batch.contig_names)
    b = _reads(batch)
    n_alleles = np.array([len(g) for g in genotypes], np.int32)
    text = b"".join(a[0].encode() + b"\0" + a[1].encode() + b"\0" for g in genotypes for a, _ in g)
    fr = np.array([f for g in genotypes for _, f in g], np.float64)
    out = np.zeros(len(genotypes), np.float64)
    _check(lib().oracle_likelihoods_of_genotypes(C.byref(b), cs, nc, batch.contig_names.index(contig), C.c_int64(locus), sample,
                                                 len(genotypes), ptr(n_alleles, i32p), text, fr.ctypes.data_as(C.POINTER(C.c_double)),
                                                 int(include_alignment), int(log_space), int(normalize),
                                                 out.ctypes.data_as(C.POINTER(C.c_double))))
    return out


def all_genotypes(batch, reference, contig, locus, sample=0, include_alignment=False, log_space=False):
    """[((ref1, alt1), (ref2, alt2), value)] — probabilitiesOfAllPossibleGenotypesFromPileup."""
    cs, nc = _contigs(reference, batch.contig_names)
    b = _reads(batch)
    text = C.c_void_p()
    _check(lib().oracle_all_genotypes(C.byref(b), cs, nc, batch.contig_names.index(contig), C.c_int64(locus), sample,
                                      int(include_alignment), int(log_space), C.byref(text)))
    s = C.string_at(text).decode()
    lib().oracle_free(text)
    out = []
    for line in s.split("\n"):
        if line:
            r1, a1, r2, a2, v = line.split("\t")
            out.append(((r1, a1), (r2, a2), float(v)))
    return out


def somatic_standard(batch: ReadBatch, reference: Reference, ranges, params: Optional[SSParamsC] = None, n_threads=1,
                     want_summary=False):
    params = params or ss_params()
    cs, nc = _contigs(reference, batch.contig_names)
    b = _reads(batch)
    r = make_ranges(ranges)
    calls, n, arena, alen = C.POINTER(SSCallC)(), C.c_int64(), C.c_void_p(), C.c_int64()
    summ, nv = C.POINTER(LocusSummaryC)(), C.c_int64()
    _check(lib().oracle_somatic_standard(C.byref(b), cs, nc, C.byref(params), r, len(ranges), n_threads, C.byref(calls),
                                         C.byref(n), C.byref(arena), C.byref(alen), C.byref(summ) if want_summary else None,
                                         C.byref(nv)))
    arena_b = C.string_at(arena, alen.value)
    lib().oracle_free(arena)
    out = dict(calls=_take(calls, SS_CALL_DTYPE, n.value), arena=arena_b, n_visited=nv.value)
    if want_summary:
        out["summary"] = _take(summ, LOCUS_SUMMARY_DTYPE, nv.value * 2).reshape(nv.value, 2)
    return out


def phred_to_success_probability(q: int) -> float:
    return lib().oracle_phred_to_success_probability(q)


def success_probability_to_phred(p: float) -> int:
    return lib().oracle_success_probability_to_phred(p)


def synth_generate(contig_len: int, depths, tumor, seed: int, read_lo: int = 0, read_hi=None, contig_name="synth1",
                   n_threads: int = 0, profile: int = 0) -> ReadBatch:
    """C twin of guacamole_b200.synth.generate_reads (fast; used for large CPU-side samples)."""
    from guacamole_b200 import synth
    N = synth.n_reads_total(contig_len, depths)
    read_hi = N if read_hi is None else min(read_hi, N)
    read_lo = max(0, read_lo)
    n = max(0, read_hi - read_lo)
    pat = synth.sample_pattern(depths)
    d = np.asarray(depths, np.int32)
    t = np.asarray(tumor, np.uint8)
    start = np.zeros(n, np.int64)
    cigar_off = np.zeros(n + 1, np.uint32)
    cigar = np.zeros(3 * n + 1, np.uint32)
    bases = np.zeros(100 * n, np.uint8)
    qual = np.zeros(100 * n, np.uint8)
    mapq, flags, sample = np.zeros(n, np.uint8), np.zeros(n, np.uint8), np.zeros(n, np.uint8)
    lib().oracle_synth_generate(C.c_int64(contig_len), C.c_int32(len(d)), ptr(d, i32p), ptr(t, u8p), ptr(pat, u8p),
                                C.c_int32(len(pat)), C.c_uint64(seed), C.c_int64(read_lo), C.c_int64(read_hi), ptr(start, i64p),
                                ptr(cigar_off, u32p), ptr(cigar, u32p), ptr(bases, u8p), ptr(qual, u8p), ptr(mapq, u8p),
                                ptr(flags, u8p), ptr(sample, u8p), C.c_int(n_threads or (os.cpu_count() or 1)), C.c_int(profile))
    return ReadBatch(n_samples=len(d), contig_names=[contig_name], contig_id=np.zeros(n, np.int32), start=start,
                     cigar_off=cigar_off, cigar=cigar[:int(cigar_off[-1])].copy(),
                     seq_off=np.arange(n + 1, dtype=np.uint64) * np.uint64(100), bases_ascii=bases, qual=qual, mapq=mapq,
                     flags=flags, sample_id=sample)


def synth_reference(seed: int, lo: int, hi: int) -> np.ndarray:
    out = np.zeros(hi - lo, np.uint8)
    lib().oracle_synth_reference(C.c_uint64(seed), C.c_int64(lo), C.c_int64(hi), ptr(out, u8p))
    return out


def somatic_joint(batch: ReadBatch, reference: Reference, loci, params, force_loci=(), raw_records=False):
    """SomaticJoint.makeCalls over one partition -> list of per-allele evidence dicts
    (guacamole_b200.commands.somatic_joint.decode form) plus the number of loci visited."""
    from guacamole_b200.commands import somatic_joint as sj
    cs, nc = _contigs(reference, batch.contig_names)
    b = _reads(batch)
    r, f = make_ranges(loci), make_ranges(force_loci)
    al, na, sm, ns, arena, alen, nv = C.c_void_p(), C.c_int64(), C.c_void_p(), C.c_int64(), C.c_void_p(), C.c_int64(), C.c_int64()
    _check(lib().oracle_somatic_joint(C.byref(b), cs, nc, C.byref(params), r, len(loci), f, len(force_loci), C.byref(al), C.byref(na),
                                      C.byref(sm), C.byref(ns), C.byref(arena), C.byref(alen), C.byref(nv)))
    arena_b = C.string_at(arena, alen.value)
    lib().oracle_free(arena)
    alleles = _take(al, sj.ALLELE_EVIDENCE_DTYPE, na.value)
    samples = _take(sm, sj.SAMPLE_EVIDENCE_DTYPE, ns.value)
    if raw_records:
        return alleles, samples, arena_b, nv.value
    return sj.decode(alleles, samples, arena_b, batch.n_samples), nv.value


def read_subsequences(batch, reference, contig, locus, length=1, sample=0):
    """[(fixed_length_sequence or None, next_alt_sequence or None)] per element (ReadSubsequence probes)."""
    cs, nc = _contigs(reference, batch.contig_names)
    b = _reads(batch)
    text = C.c_void_p()
    _check(lib().oracle_read_subsequences(C.byref(b), cs, nc, batch.contig_names.index(contig), C.c_int64(locus), sample, length,
                                          C.byref(text)))
    s = C.string_at(text).decode()
    lib().oracle_free(text)
    rows = [l.split("\t") for l in s.split("\n") if l]
    out = []
    for i in range(0, len(rows), 2):
        fx, al = rows[i], rows[i + 1]
        out.append((None if fx[1] == "-" else fx[1], None if al[1] == "-" else al[1]))
    return out


def pileup_stats(batch, reference, contig, locus, ref_seq, mixtures=(), sample=0):
    """(total depth, [(allele, depth)...] most frequent first, [log10 likelihood per mixture])."""
    cs, nc = _contigs(reference, batch.contig_names)
    b = _reads(batch)
    n_keys = np.array([len(m) for m in mixtures] or [0], np.int32)
    keys = b"".join(k.encode() + b"\0" for m in mixtures for k, _ in m)
    fr = np.array([f for m in mixtures for _, f in m] or [0.0], np.float64)
    text = C.c_void_p()
    _check(lib().oracle_pileup_stats(C.byref(b), cs, nc, batch.contig_names.index(contig), C.c_int64(locus), sample, ref_seq.encode(),
                                     len(mixtures), ptr(n_keys, i32p), keys, fr.ctypes.data_as(C.POINTER(C.c_double)), C.byref(text)))
    s = C.string_at(text).decode()
    lib().oracle_free(text)
    total, depths, lls = 0, [], []
    for l in s.split("\n"):
        if l.startswith("total="):
            total = int(l[6:])
        elif l.startswith("LL\t"):
            lls.append(float(l[3:]))
        elif l:
            a, c = l.rsplit("=", 1)
            depths.append((a, int(c)))
    return total, depths, lls


def germline_standard(batch: ReadBatch, reference: Reference, ranges, params=None):
    """The germline toy caller (shortcut branch of discoverGermlineVariants) over one partition."""
    from guacamole_b200.native import CALLED_ALLELE_DTYPE, gs_params
    params = params or gs_params()
    cs, nc = _contigs(reference, batch.contig_names)
    b = _reads(batch)
    r = make_ranges(ranges)
    calls, n, arena, alen, act, na, nv = (C.c_void_p(), C.c_int64(), C.c_void_p(), C.c_int64(), C.c_void_p(), C.c_int64(),
                                          C.c_int64())
    _check(lib().oracle_germline_standard(C.byref(b), cs, nc, C.byref(params), r, len(ranges), C.byref(calls), C.byref(n),
                                          C.byref(arena), C.byref(alen), C.byref(act), C.byref(na), C.byref(nv)))
    arena_b = C.string_at(arena, alen.value)
    lib().oracle_free(arena)
    active = _take(act, np.int64, na.value * 2).reshape(-1, 2)
    return dict(calls=_take(calls, CALLED_ALLELE_DTYPE, n.value), arena=arena_b,
                active_loci=[(int(a), int(b)) for a, b in active], n_visited=nv.value)


class GAParamsC(C.Structure):
    _fields_ = [("sample_id", C.c_int32), ("kmer_size", C.c_int32), ("assembly_window_range", C.c_int32), ("min_occurrence", C.c_int32),
                ("min_mean_kmer_quality", C.c_int32), ("min_alt_reads", C.c_int32), ("min_phred_scaled_likelihood", C.c_int32),
                ("shortcut_assembly", C.c_int32), ("min_area_vaf", C.c_float), ("_pad", C.c_int32)]


def ga_params(sample_id=0, kmer_size=45, assembly_window_range=20, min_occurrence=3, min_mean_kmer_quality=0, min_alt_reads=2,
              min_phred_scaled_likelihood=0, shortcut_assembly=False, min_area_vaf_percent=5):
    p = GAParamsC()
    p.sample_id, p.kmer_size, p.assembly_window_range = sample_id, kmer_size, assembly_window_range
    p.min_occurrence, p.min_mean_kmer_quality, p.min_alt_reads = min_occurrence, min_mean_kmer_quality, min_alt_reads
    p.min_phred_scaled_likelihood, p.shortcut_assembly = min_phred_scaled_likelihood, int(shortcut_assembly)
    p.min_area_vaf = float(np.float32(min_area_vaf_percent) / np.float32(100.0))
    return p


def _text_call(fn, *args):
    text = C.c_void_p()
    _check(fn(*args, C.byref(text)))
    s = C.string_at(text).decode()
    lib().oracle_free(text)
    return [l for l in s.split("\n") if l]


def debruijn(sequences, kmer_size, op, min_occurrence=1, min_mean_kmer_quality=0, merge_nodes=False, qualities=None, source="",
             sink="", max_paths=10):
    """De Bruijn graph probe: op 'counts' | 'dfs' | 'forward' | 'backward' | 'merge_index'."""
    ops = {"counts": 0, "dfs": 1, "forward": 2, "backward": 3, "merge_index": 4}
    seqs = b"".join(s.encode() + b"\0" for s in sequences)
    q = None
    if qualities is not None:
        qa = np.array([x for qs in qualities for x in qs], np.uint8)
        q = ptr(qa, u8p)
    lines = _text_call(lib().oracle_debruijn, seqs, len(sequences), q, kmer_size, min_occurrence, min_mean_kmer_quality, int(merge_nodes),
                       ops[op], source.encode(), sink.encode(), max_paths)
    if op == "counts":
        return {l.split("=")[0]: int(l.split("=")[1]) for l in lines}
    if op == "merge_index":
        return [(l.split("=")[0], l.split("=")[1]) for l in lines]
    return lines


def read_alignment(sequence, reference, mismatch_probability=0.0, open_gap_probability=0.0, close_gap_probability=0.0):
    """ReadAlignment(sequence, reference) -> (cigar, refStartIdx, refEndIdx, alignmentScore, rawScore); zeros = defaults."""
    l = _text_call(lib().oracle_read_alignment, sequence.encode(), reference.encode(), C.c_double(mismatch_probability),
                   C.c_double(open_gap_probability), C.c_double(close_gap_probability))
    f = (l[0] if l else "\t0\t0\t0\t0").split("\t")
    return f[0], int(f[1]), int(f[2]), int(f[3]), float(f[4])


def germline_assembly(batch: ReadBatch, reference: Reference, ranges, params):
    """discoverGermlineVariants -> list of tuples: ('A', contig, locus, ref, alt, depth, ...), ('P', ...), ('W', contig, locus, nPaths)."""
    cs, nc = _contigs(reference, batch.contig_names)
    b = _reads(batch)
    r = make_ranges(ranges)
    out = []
    for l in _text_call(lib().oracle_germline_assembly, C.byref(b), cs, nc, C.byref(params), r, len(ranges)):
        f = l.split("\t")
        out.append(tuple([f[0], int(f[1]), int(f[2])] + f[3:]))
    return out


# ---- the small consumers of the operator, restated over the oracle's own pileups -------------------------
def variant_support(batch: ReadBatch, reference: Reference, loci):
    """VariantSupport.Caller.pileupsToAlleleCounts (M/commands/VariantSupport.scala:100-112) through
    pileupFlatMapMultipleSamples(skipEmpty = true) (:83-89): per visited locus, per sample, pileup.elements.groupBy(_.allele)
    with the group sizes.  -> [(sample_id, contig_id, locus, ref, alt, count)] in (range order, locus), sample, (ref, alt)
    order (the Scala's order inside a pileup is a hash map's; VariantSupportSuite sorts before it compares)."""
    import collections
    contigs, positions, _ = pileup_counts(batch, reference, loci, skip_empty=True)
    out = []
    for c, x in zip(contigs, positions):
        for s in range(batch.n_samples):
            groups = collections.Counter((ref, alt) for ref, alt, _, _ in pileup_alleles(batch, reference, batch.contig_names[int(c)], int(x), s))
            out += [(s, int(c), int(x), ref, alt, n) for (ref, alt), n in sorted(groups.items())]
    return out


def vaf_histogram(batch: ReadBatch, reference: Reference, loci, bins=20, min_read_depth=0, min_vaf_percent=0):
    """VariantLocus.apply (M/commands/VAFHistogram.scala:37-58), the filter of variantLociFromReads (:214-229) and
    generateVAFHistograms (:194-212), Float arithmetic as written there.  -> (hist [n_samples, 101], loci per sample,
    [(sample_id, contig_id, locus, vaf float32)])."""
    assert 1 <= bins <= 100, "Bins should be between 1 and 100"
    contigs, positions, counts = pileup_counts(batch, reference, loci, skip_empty=True)
    S = batch.n_samples
    hist, tot, recs = np.zeros((S, 101), np.int64), np.zeros(S, np.int64), []
    for i, (c, x) in enumerate(zip(contigs, positions)):
        for s in range(S):
            depth, ref_depth = int(counts[i, s]["depth"]), int(counts[i, s]["reference_depth"])
            if depth < min_read_depth or ref_depth == depth:
                continue
            vaf = np.float32(depth - ref_depth) / np.float32(depth)   # (depth - referenceDepth).toFloat / depth
            if float(vaf) < min_vaf_percent / 100.0:
                continue
            percent = int(np.float32(vaf) * np.float32(100))          # (variantAlleleFrequency * 100).toInt
            b = percent - (percent % (100 // bins))
            hist[s, b] += 1
            tot[s] += 1
            recs.append((s, int(c), int(x), np.float32(vaf)))
    return hist, tot, recs


# ---- coverage-aware ("capped") partitioning, restated literally; pure Python: small cases only ------------
def contig_coverage(intervals, half_window=0, eligible=None):
    """ContigCoverageIterator (M/readsets/iterator/ContigCoverageIterator.scala:28-117) restricted to `eligible` loci (its
    `intersect`, M/readsets/rdd/CoverageRDD.scala:134-137): [(locus, depth, starts)] for every evaluated locus with coverage.
    An interval counts towards depth at every locus of [start - halfWindow, end + halfWindow) and towards starts at the first
    EVALUATED one of them.  `intervals` sorted by start; `eligible` a sorted iterable of loci (None: every locus)."""
    last = 0
    for s, _ in intervals:
        if s < last:
            raise ValueError(f"Consecutive regions rewinding from {last} to {s}")  # RegionsNotSortedException (:71-76)
        last = s
    if not intervals:
        return []
    if eligible is None:
        eligible = range(max(0, min(s for s, _ in intervals) - half_window), max(e for _, e in intervals) + half_window)
    counted, out = set(), []
    for x in eligible:
        covering = [i for i, (s, e) in enumerate(intervals) if s <= x + half_window and e > max(0, x - half_window)]
        if covering:
            out.append((x, len(covering), sum(1 for i in covering if i not in counted)))
            counted.update(covering)
    return out


def take_loci(coverages, max_regions_per_partition, trim_ranges=False):
    """TakeLociIterator (M/loci/set/TakeLociIterator.scala:25-196), literally: `coverages` = [(contig, locus, depth, starts)]
    sorted by (contig, locus) -> list of loci sets, each a list of (contig, start, end) intervals in the order they were
    built."""
    by_contig, order = {}, []
    for c, x, d, st in coverages:
        if c not in by_contig:
            by_contig[c] = []
            order.append(c)
        by_contig[c].append((x, d, st))
    heads = {c: 0 for c in order}

    def take(contig, num_regions, drop_above):  # takeLoci (:108-196)
        its, intervals, cur, cur_iv = by_contig[contig], [], 0, None
        while heads[contig] < len(its):
            locus, depth, starts = its[heads[contig]]
            new = depth if cur == 0 else starts
            if cur + new > num_regions:
                if cur_iv:
                    intervals.append(cur_iv)
                    cur_iv = None
                if new > drop_above:
                    heads[contig] += 1
                else:
                    return (cur, intervals) if intervals else None
            else:
                cur += new
                if cur_iv and (not trim_ranges or locus == cur_iv[1]):
                    cur_iv = (cur_iv[0], locus + 1)
                else:
                    if cur_iv:
                        intervals.append(cur_iv)
                    cur_iv = (locus, locus + 1)
                heads[contig] += 1
        if cur_iv:
            intervals.append(cur_iv)
        return (cur, intervals) if intervals else None

    sets, ci = [], 0
    while ci < len(order):  # _advance (:36-100), once per loci set
        cur_regions, contigs, done = 0, [], False
        while ci < len(order) and not done:
            c = order[ci]
            got = take(c, max_regions_per_partition - cur_regions, max_regions_per_partition)
            if got is None:
                if heads[c] < len(by_contig[c]):
                    done = True
                else:
                    ci += 1
            else:
                cur_regions += got[0]
                contigs += [(c, a, b) for a, b in got[1]]
        if contigs:
            sets.append(contigs)
        elif not done:
            break
    return sets


def capped_partitioning(batch: ReadBatch, loci, half_window=0, max_reads_per_partition=100000, trim_ranges=False):
    """CappedRegionsPartitioner.partition (M/loci/partitioning/CappedRegionsPartitioner.scala:66-88): coverage of all the
    batch's reads over the eligible loci, then TakeLociIterator.  -> [(contig_id, start, end, partition)]."""
    ends = batch.ends()
    cov = []
    for c in sorted({r[0] for r in loci}):
        sel = np.flatnonzero(batch.contig_id == c)
        intervals = [(int(batch.start[i]), int(ends[i])) for i in sel]
        eligible = [x for (cc, s, e) in loci if cc == c for x in range(s, e)]
        cov += [(c, x, d, st) for x, d, st in contig_coverage(intervals, half_window, eligible)]
    out = []
    for p, contigs in enumerate(take_loci(cov, max_reads_per_partition, trim_ranges)):
        out += [(c, a, b, p) for c, a, b in contigs]
    return out
