"""ctypes binding of libguac_b200.so (include/guac_b200.h).

This is the tested consumer of the C ABI in this repository (the Scala/JNI shim of
INTEGRATION.md binds the same symbols).  There is no CPU fallback: if the shared library
is missing, or no CUDA device is present, constructing a Context raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence, Tuple

import numpy as np

from .io import Reference
from .reads import ReadBatch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GUAC_B200_LIB") or os.path.join(_HERE, "csrc", "libguac_b200.so")  # override: kernel A/B builds


class GuacError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"guac status {status}: {message}")
        self.status = status
        self.message = message


GUAC_OK, GUAC_E_INVALID_INPUT, GUAC_E_CUDA, GUAC_E_NO_DEVICE, GUAC_E_OOM, GUAC_E_UNSUPPORTED, GUAC_E_STATE = 0, -1, -2, -3, -4, -5, -6

u8p, u32p, u64p, i32p, i64p = (C.POINTER(t) for t in (C.c_uint8, C.c_uint32, C.c_uint64, C.c_int32, C.c_int64))


class ReadBatchC(C.Structure):
    _fields_ = [("n_reads", C.c_int64), ("n_samples", C.c_int32), ("contig_id", i32p), ("start", i64p),
                ("cigar_off", u32p), ("cigar", u32p), ("seq_off", u64p), ("bases2", u8p), ("nmask", u8p),
                ("n_exc", C.c_int64), ("exc_pos", u64p), ("exc_base", u8p), ("qual", u8p), ("mapq", u8p),
                ("flags", u8p), ("sample_id", u8p)]


class RangeC(C.Structure):
    _fields_ = [("contig_id", C.c_int32), ("_pad", C.c_int32), ("start", C.c_int64), ("end", C.c_int64)]


class LocusCountsC(C.Structure):
    _fields_ = [(n, C.c_uint32) for n in ("depth", "reference_depth", "positive_depth", "mismatches", "insertions",
                                          "deletions", "mid_deletions", "clipped")]


class ElementC(C.Structure):
    _fields_ = [("read_index", C.c_int64), ("read_position", C.c_int32), ("cigar_element_index", C.c_int32),
                ("cigar_element_locus", C.c_int64), ("index_within_cigar_element", C.c_int32), ("kind", C.c_uint8),
                ("base", C.c_uint8), ("quality", C.c_uint8), ("sample_id", C.c_uint8), ("indel_length", C.c_int32),
                ("_pad", C.c_int32)]


class AlleleEvidenceC(C.Structure):
    _fields_ = [("probability", C.c_double), ("read_depth", C.c_int32), ("allele_read_depth", C.c_int32),
                ("forward_depth", C.c_int32), ("allele_forward_depth", C.c_int32), ("mean_mapping_quality", C.c_double),
                ("median_mapping_quality", C.c_double), ("mean_base_quality", C.c_double),
                ("median_base_quality", C.c_double), ("median_mismatches_per_read", C.c_double)]


class SSParamsC(C.Structure):
    _fields_ = [("normal_sample_id", C.c_int32), ("tumor_sample_id", C.c_int32), ("normal_odds_threshold", C.c_int32),
                ("tumor_odds_threshold", C.c_int32), ("max_read_depth", C.c_int32),
                ("max_normal_alternate_read_depth", C.c_int32), ("min_tumor_variant_allele_frequency", C.c_float),
                ("_pad", C.c_int32)]


class SSCallC(C.Structure):
    _fields_ = [("contig_id", C.c_int32), ("_pad", C.c_int32), ("start", C.c_int64), ("ref_off", C.c_uint32),
                ("ref_len", C.c_uint32), ("alt_off", C.c_uint32), ("alt_len", C.c_uint32), ("somatic_log_odds", C.c_double),
                ("tumor_ref_log_prob", C.c_double), ("tumor_alt_log_prob", C.c_double), ("normal_ref_log_prob", C.c_double),
                ("normal_alt_log_prob", C.c_double), ("tumor_variant_evidence", AlleleEvidenceC),
                ("normal_reference_evidence", AlleleEvidenceC)]


class LocusSummaryC(C.Structure):
    _fields_ = [("depth", C.c_uint16), ("reference_depth", C.c_uint16), ("top_alt_count", C.c_uint16), ("flags", C.c_uint16)]


class GSParamsC(C.Structure):
    _fields_ = [("sample_id", C.c_int32), ("min_alt_reads", C.c_int32), ("min_phred_scaled_likelihood", C.c_int32),
                ("respect_active_regions", C.c_int32), ("min_area_vaf", C.c_float), ("_pad", C.c_int32)]


def gs_params(sample_id=0, min_alt_reads=2, min_phred_scaled_likelihood=0, respect_active_regions=0, min_area_vaf_percent=5):
    """Defaults of GermlineAssemblyCaller (commands/GermlineAssemblyCaller.scala:40-47,92) and AssemblyArgs
    (assembly/AssemblyArgs.scala:33); minAreaVaf is percent / 100.0f as a float (:70)."""
    p = GSParamsC()
    p.sample_id, p.min_alt_reads, p.min_phred_scaled_likelihood = sample_id, min_alt_reads, min_phred_scaled_likelihood
    p.respect_active_regions = int(respect_active_regions)
    p.min_area_vaf = float(np.float32(min_area_vaf_percent) / np.float32(100.0))
    return p


ALLELE_COUNT_DTYPE = np.dtype([("contig_id", np.int32), ("sample_id", np.int32), ("locus", np.int64), ("ref_off", np.uint32),
                               ("ref_len", np.uint32), ("alt_off", np.uint32), ("alt_len", np.uint32), ("count", np.int32), ("_pad", np.int32)])
VARIANT_LOCUS_DTYPE = np.dtype([("sample_id", np.int32), ("contig_id", np.int32), ("locus", np.int64),
                                ("variant_allele_frequency", np.float32), ("_pad", np.int32)])
LOCUS_COUNTS_DTYPE = np.dtype([(n, np.uint32) for n, _ in LocusCountsC._fields_])
ELEMENT_DTYPE = np.dtype([("read_index", np.int64), ("read_position", np.int32), ("cigar_element_index", np.int32),
                          ("cigar_element_locus", np.int64), ("index_within_cigar_element", np.int32), ("kind", np.uint8),
                          ("base", np.uint8), ("quality", np.uint8), ("sample_id", np.uint8), ("indel_length", np.int32),
                          ("_pad", np.int32)])
_EV = [("probability", np.float64), ("read_depth", np.int32), ("allele_read_depth", np.int32), ("forward_depth", np.int32),
       ("allele_forward_depth", np.int32), ("mean_mapping_quality", np.float64), ("median_mapping_quality", np.float64),
       ("mean_base_quality", np.float64), ("median_base_quality", np.float64), ("median_mismatches_per_read", np.float64)]
SS_CALL_DTYPE = np.dtype([("contig_id", np.int32), ("_pad", np.int32), ("start", np.int64), ("ref_off", np.uint32),
                          ("ref_len", np.uint32), ("alt_off", np.uint32), ("alt_len", np.uint32),
                          ("somatic_log_odds", np.float64), ("tumor_ref_log_prob", np.float64),
                          ("tumor_alt_log_prob", np.float64), ("normal_ref_log_prob", np.float64),
                          ("normal_alt_log_prob", np.float64), ("tumor_variant_evidence", np.dtype(_EV)),
                          ("normal_reference_evidence", np.dtype(_EV))])
class GAParamsC(C.Structure):
    _fields_ = [("sample_id", C.c_int32), ("kmer_size", C.c_int32), ("assembly_window_range", C.c_int32), ("min_occurrence", C.c_int32),
                ("min_mean_kmer_quality", C.c_int32), ("min_alt_reads", C.c_int32), ("min_phred_scaled_likelihood", C.c_int32),
                ("shortcut_assembly", C.c_int32), ("min_area_vaf", C.c_float), ("_pad", C.c_int32)]


def ga_params(sample_id=0, kmer_size=45, assembly_window_range=20, min_occurrence=3, min_mean_kmer_quality=0, min_alt_reads=2,
              min_phred_scaled_likelihood=0, shortcut_assembly=False, min_area_vaf_percent=5):
    """Defaults of AssemblyArgs (assembly/AssemblyArgs.scala:13-45) and GermlineAssemblyCaller.Arguments
    (commands/GermlineAssemblyCaller.scala:33-47,92).  The reference's own defaults (range 20 < k-mer 45) make its window
    jump move backward after the first assembled window; its tests run with range 120, k-mer 47."""
    p = GAParamsC()
    p.sample_id, p.kmer_size, p.assembly_window_range = sample_id, kmer_size, assembly_window_range
    p.min_occurrence, p.min_mean_kmer_quality, p.min_alt_reads = min_occurrence, min_mean_kmer_quality, min_alt_reads
    p.min_phred_scaled_likelihood, p.shortcut_assembly = min_phred_scaled_likelihood, int(shortcut_assembly)
    p.min_area_vaf = float(np.float32(min_area_vaf_percent) / np.float32(100.0))
    return p


ASSEMBLED_WINDOW_DTYPE = np.dtype([("contig_id", np.int32), ("n_paths", np.int32), ("locus", np.int64), ("n_graph_nodes", np.int32),
                                   ("consumed", np.int32)])
LOCUS_SUMMARY_DTYPE = np.dtype([("depth", np.uint16), ("reference_depth", np.uint16), ("top_alt_count", np.uint16),
                                ("flags", np.uint16)])
CALLED_ALLELE_DTYPE = np.dtype([("contig_id", np.int32), ("n_genotypes", np.int32), ("start", np.int64), ("ref_off", np.uint32),
                                ("ref_len", np.uint32), ("alt_off", np.uint32), ("alt_len", np.uint32),
                                ("genotype_log_probability", np.float64), ("evidence", np.dtype(_EV))])
assert LOCUS_COUNTS_DTYPE.itemsize == C.sizeof(LocusCountsC)
assert ELEMENT_DTYPE.itemsize == C.sizeof(ElementC)
assert SS_CALL_DTYPE.itemsize == C.sizeof(SSCallC)


def ptr(a: np.ndarray, t):
    return a.ctypes.data_as(t)


def make_ranges(ranges: Sequence[Tuple[int, int, int]]):
    arr = (RangeC * max(1, len(ranges)))()
    for i, (c, s, e) in enumerate(ranges):
        arr[i].contig_id, arr[i].start, arr[i].end = c, s, e
    return arr


def ss_params(normal_sample_id=0, tumor_sample_id=1, normal_odds_threshold=5, tumor_odds_threshold=10,
              max_read_depth=2**31 - 1, max_normal_alternate_read_depth=4, min_tumor_variant_allele_frequency_percent=3):
    """Defaults of SomaticStandard.Arguments (commands/SomaticStandardCaller.scala:44-70);
    the VAF floor is percent / 100.0f as a float (:97)."""
    p = SSParamsC()
    p.normal_sample_id, p.tumor_sample_id = normal_sample_id, tumor_sample_id
    p.normal_odds_threshold, p.tumor_odds_threshold = normal_odds_threshold, tumor_odds_threshold
    p.max_read_depth = max_read_depth
    p.max_normal_alternate_read_depth = max_normal_alternate_read_depth
    p.min_tumor_variant_allele_frequency = float(np.float32(min_tumor_variant_allele_frequency_percent) / np.float32(100.0))
    return p


_lib = None


def load_library() -> C.CDLL:
    """Loads the CUDA library.  Raises (loudly) if it has not been built: there is no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(there is no CPU fallback)")
        _lib = C.CDLL(LIB_PATH)
        _lib.guac_last_error.restype = C.c_char_p
        _lib.guac_last_error.argtypes = [C.c_void_p]
        _lib.guac_kernel_launches.restype = C.c_int64
        _lib.guac_kernel_launches.argtypes = [C.c_void_p]
        _lib.guac_free.argtypes = [C.c_void_p, C.c_void_p]
        _lib.guac_free.restype = None
        _lib.guac_ctx_destroy.argtypes = [C.c_void_p]
        _lib.guac_ctx_destroy.restype = None
    return _lib


def pack_bases(ascii_bases: np.ndarray):
    """guac_pack_bases over a whole batch -> dict(bases2, nmask, exc_pos, exc_base)."""
    lib = load_library()
    n = int(ascii_bases.shape[0])
    bases2 = np.zeros((n + 3) // 4 + 8, np.uint8)
    nmask = np.zeros((n + 7) // 8 + 8, np.uint8)
    cap = max(16, int(np.count_nonzero(~np.isin(ascii_bases, (65, 67, 71, 84)))))
    exc_pos = np.zeros(cap, np.uint64)
    exc_base = np.zeros(cap, np.uint8)
    n_exc = C.c_int64(0)
    a = np.ascontiguousarray(ascii_bases)
    st = lib.guac_pack_bases(ptr(a, u8p), C.c_int64(n), C.c_uint64(0), ptr(bases2, u8p), ptr(nmask, u8p), ptr(exc_pos, u64p),
                             ptr(exc_base, u8p), C.c_int64(cap), C.byref(n_exc))
    if st != 0:
        raise GuacError(st, "guac_pack_bases failed")
    return dict(bases2=bases2, nmask=nmask, exc_pos=exc_pos[:max(1, n_exc.value)], exc_base=exc_base[:max(1, n_exc.value)],
                n_exc=n_exc.value)


def batch_struct(batch: ReadBatch) -> ReadBatchC:
    if batch.packed is None:
        batch.packed = pack_bases(batch.bases_ascii)
    p = batch.packed
    b = ReadBatchC()
    b.n_reads, b.n_samples = batch.n_reads, batch.n_samples
    b.contig_id, b.start = ptr(batch.contig_id, i32p), ptr(batch.start, i64p)
    b.cigar_off, b.cigar = ptr(batch.cigar_off, u32p), ptr(batch.cigar, u32p)
    b.seq_off = ptr(batch.seq_off, u64p)
    b.bases2, b.nmask = ptr(p["bases2"], u8p), ptr(p["nmask"], u8p)
    b.n_exc, b.exc_pos, b.exc_base = p["n_exc"], ptr(p["exc_pos"], u64p), ptr(p["exc_base"], u8p)
    b.qual, b.mapq, b.flags, b.sample_id = ptr(batch.qual, u8p), ptr(batch.mapq, u8p), ptr(batch.flags, u8p), ptr(batch.sample_id, u8p)
    return b


u16p = C.POINTER(C.c_uint16)


class ReadBatchCompactC(C.Structure):
    """guac_read_batch_compact: the PCIe wire form (lengths instead of offsets, contig runs, 4-bit quality codes)."""
    _fields_ = [("n_reads", C.c_int64), ("n_samples", C.c_int32), ("n_contig_runs", C.c_int32), ("run_contig_id", i32p),
                ("run_first_read", i64p), ("start", i64p), ("read_len", u16p), ("n_cigar_ops", u16p), ("cigar", u32p),
                ("bases2", u8p), ("n_exc", C.c_int64), ("exc_pos", u64p), ("exc_base", u8p), ("qual_bits", C.c_int32),
                ("_pad", C.c_int32), ("qual_table", C.c_uint8 * 16), ("qual", u8p), ("mapq", u8p), ("flags", u8p),
                ("sample_id", u8p),
                # narrow forms of start / n_cigar_ops / read_len (NULL / 0: the wide column is given)
                ("run_start_base", i64p), ("start_rel", u32p), ("n_cigar_ops8", u8p), ("uniform_read_len", C.c_int32),
                ("_pad2", C.c_int32),
                # reference-relative bases (NULL: bases2 holds every read)
                ("mm_count", u8p), ("mm_words", u16p), ("n_mm_words", C.c_int64), ("n_explicit_bases", C.c_int64)]

    def wire_bytes(self, n_bases: int, n_cigar_words: int) -> int:
        """Bytes this batch puts on the wire (what guac_reads_upload_compact copies to the device)."""
        n = int(self.n_reads)
        per_read = (4 if self.start_rel else 8) + (0 if self.uniform_read_len > 0 else 2) + (1 if self.n_cigar_ops8 else 2) + 3
        q = n_bases if self.qual_bits == 8 else (n_bases * self.qual_bits + 7) // 8
        bases = (n_bases + 3) // 4
        if self.mm_count:  # reference-relative: a count per read, the difference words, the explicit reads' bases
            bases = n + 2 * int(self.n_mm_words) + (int(self.n_explicit_bases) + 3) // 4
        return n * per_read + 4 * n_cigar_words + bases + q + 9 * int(self.n_exc) + 20 * int(self.n_contig_runs) + 8


def compact_struct(batch: ReadBatch, narrow: bool = True, ref=None) -> ReadBatchCompactC:
    """Host-side packer of the compact wire form (numpy); keeps the arrays alive on the returned struct.
    narrow: use the narrow start / operator-count / read-length columns where the batch allows them.
    ref (a `Reference`): reads that can travel as their differences from it do (mm_count / mm_words)."""
    if batch.packed is None:
        batch.packed = pack_bases(batch.bases_ascii)
    p = batch.packed
    n = batch.n_reads
    keep = {}
    b = ReadBatchCompactC()
    b.n_reads, b.n_samples = n, batch.n_samples
    change = np.flatnonzero(np.diff(batch.contig_id)) + 1 if n else np.zeros(0, np.int64)
    first = np.concatenate([[0], change, [n]]).astype(np.int64) if n else np.zeros(1, np.int64)
    keep["run_first"] = first
    keep["run_contig"] = np.ascontiguousarray(batch.contig_id[first[:-1]], np.int32) if n else np.zeros(0, np.int32)
    b.n_contig_runs = len(keep["run_contig"])
    keep["read_len"] = np.diff(batch.seq_off.astype(np.int64)).astype(np.uint16)
    keep["n_ops"] = np.diff(batch.cigar_off.astype(np.int64)).astype(np.uint16)
    assert n == 0 or (np.diff(batch.seq_off.astype(np.int64)).max() <= 65535 and np.diff(batch.cigar_off.astype(np.int64)).max() <= 65535)
    values = np.unique(batch.qual)
    if 0 < len(values) <= 16:
        bits = 2 if len(values) <= 4 else 4
        per = 8 // bits
        codes = np.zeros(256, np.uint8)
        codes[values] = np.arange(len(values), dtype=np.uint8)
        c = codes[batch.qual]
        if len(c) % per:
            c = np.concatenate([c, np.zeros(per - len(c) % per, np.uint8)])
        packed = np.zeros(len(c) // per, np.uint8)
        for k in range(per):  # lowest bits = lowest base index
            packed |= c[k::per] << (bits * k)
        keep["qual"] = packed
        b.qual_bits = bits
        for i, v in enumerate(values):
            b.qual_table[i] = int(v)
    else:
        keep["qual"] = batch.qual
        b.qual_bits = 8
    b.run_contig_id, b.run_first_read = ptr(keep["run_contig"], i32p), ptr(keep["run_first"], i64p)
    b.start, b.read_len, b.n_cigar_ops = ptr(batch.start, i64p), ptr(keep["read_len"], u16p), ptr(keep["n_ops"], u16p)
    b.cigar, b.bases2 = ptr(batch.cigar, u32p), ptr(p["bases2"], u8p)
    if narrow and n:
        lens = np.diff(batch.seq_off.astype(np.int64))
        ops = np.diff(batch.cigar_off.astype(np.int64))
        base = batch.start[first[:-1]].astype(np.int64)  # the first read of a run has its lowest start
        run_of = np.searchsorted(first, np.arange(n), side="right") - 1
        rel = batch.start.astype(np.int64) - base[run_of]
        if rel.min() >= 0 and rel.max() < 2**32:
            keep["run_base"], keep["rel"] = np.ascontiguousarray(base), rel.astype(np.uint32)
            b.run_start_base, b.start_rel, b.start = ptr(keep["run_base"], i64p), ptr(keep["rel"], u32p), None
        if lens.min() == lens.max() and 0 < lens[0] <= 65535:
            b.uniform_read_len, b.read_len = int(lens[0]), None
        if ops.max() <= 255 and ops.min() >= 1:
            first_word = batch.cigar[batch.cigar_off[:-1].astype(np.int64)]
            implicit = (ops == 1) & (first_word == (lens.astype(np.uint32) << 4)) & (lens < 2**28)
            ops8 = np.where(implicit, 0, ops).astype(np.uint8)
            word_keep = np.repeat(~implicit, ops)  # per CIGAR word: does its read keep its words on the wire
            keep["ops8"], keep["wire"] = ops8, np.ascontiguousarray(batch.cigar[word_keep])
            b.n_cigar_ops8, b.n_cigar_ops, b.cigar = ptr(keep["ops8"], u8p), None, ptr(keep["wire"], u32p)
    b.n_exc, b.exc_pos, b.exc_base = p["n_exc"], ptr(p["exc_pos"], u64p), ptr(p["exc_base"], u8p)
    b.qual, b.mapq, b.flags, b.sample_id = ptr(keep["qual"], u8p), ptr(batch.mapq, u8p), ptr(batch.flags, u8p), ptr(batch.sample_id, u8p)
    if ref is not None and n:
        mm_count, mm_words, explicit = reference_relative_columns(batch, ref)
        code = np.full(256, 0, np.uint8)
        for i, ch in enumerate(b"ACGT"):
            code[ch] = i
        sym = code[explicit] if len(explicit) else np.zeros(0, np.uint8)
        if len(sym) % 4:
            sym = np.concatenate([sym, np.zeros(4 - len(sym) % 4, np.uint8)])
        packed = (sym[0::4] | (sym[1::4] << 2) | (sym[2::4] << 4) | (sym[3::4] << 6)).astype(np.uint8) if len(sym) else np.zeros(0, np.uint8)
        keep["mm_count"], keep["mm_words"] = mm_count, np.concatenate([mm_words, np.zeros(1, np.uint16)])
        keep["explicit2"] = np.concatenate([packed, np.zeros(8, np.uint8)])
        b.mm_count, b.mm_words, b.bases2 = ptr(keep["mm_count"], u8p), ptr(keep["mm_words"], u16p), ptr(keep["explicit2"], u8p)
        b.n_mm_words, b.n_explicit_bases = len(mm_words), len(explicit)
    b._keep = keep
    return b


def reference_relative_columns(batch: ReadBatch, ref):
    """(mm_count [n_reads] u8, mm_words u16, ASCII bases of the explicit reads) — see guac_read_batch_compact::mm_count: a read
    with one aligned operator over all its bases, A C G T only, wholly on A C G T reference inside its contig's window and
    shorter than 16384 bases travels as its differences (position << 2 | reference code XOR read code), at most 254 of them."""
    n = batch.n_reads
    code = np.full(256, 255, np.uint8)
    for i, ch in enumerate(b"ACGT"):
        code[ch] = i
    mm_count = np.full(n, 255, np.uint8)
    words, explicit = [], []
    seq_off, cig_off = batch.seq_off.astype(np.int64), batch.cigar_off.astype(np.int64)
    for r in range(n):
        s0, s1 = int(seq_off[r]), int(seq_off[r + 1])
        bases = batch.bases_ascii[s0:s1]
        ok = cig_off[r + 1] - cig_off[r] == 1 and s1 - s0 < 16384
        if ok:
            w = int(batch.cigar[cig_off[r]])
            ok = (w & 15) in (0, 7, 8) and (w >> 4) == s1 - s0  # M, =, X over every base
        window = ref.contigs.get(batch.contig_names[int(batch.contig_id[r])]) if ok else None
        if window is not None:
            rel = int(batch.start[r]) - window.offset
            ok = rel >= 0 and rel + (s1 - s0) <= len(window.bases)
            if ok:
                rc, bc = code[window.bases[rel:rel + s1 - s0]], code[bases]
                ok = bool((rc < 4).all() and (bc < 4).all())
                if ok:
                    pos = np.flatnonzero(rc != bc)
                    ok = len(pos) < 255
                    if ok:
                        mm_count[r] = len(pos)
                        words.append(((pos.astype(np.uint16) << 2) | (rc[pos] ^ bc[pos])).astype(np.uint16))
        else:
            ok = False
        if not ok:
            explicit.append(bases)
    return (mm_count, np.concatenate(words).astype(np.uint16) if words else np.zeros(0, np.uint16),
            np.concatenate(explicit) if explicit else np.zeros(0, np.uint8))


def _take(lib, ctx, p, dtype, n) -> np.ndarray:
    """Copies a library-owned output buffer into numpy and releases it."""
    if n == 0 or not p:
        out = np.zeros(0, dtype)
    else:
        nbytes = int(n) * np.dtype(dtype).itemsize
        addr = C.cast(p, C.c_void_p).value
        # one copy out of the pinned buffer, as bytes: numpy copies a structured dtype field by field, 15x slower
        out = np.frombuffer((C.c_char * nbytes).from_address(addr), dtype=np.uint8).copy().view(dtype)
    if p:
        lib.guac_free(ctx, C.cast(p, C.c_void_p))
    return out


class Context:
    """guac_ctx: one CUDA device, one stream, not thread-safe."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        self.ctx = C.c_void_p()
        st = self.lib.guac_ctx_create(C.c_int(device), C.byref(self.ctx))
        if st != 0:
            raise GuacError(st, "guac_ctx_create failed (no CUDA device? there is no CPU fallback)")
        self.contig_names: List[str] = []

    def close(self):
        if self.ctx:
            self.lib.guac_ctx_destroy(self.ctx)
            self.ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, st: int):
        if st != 0:
            raise GuacError(st, (self.lib.guac_last_error(self.ctx) or b"").decode())

    @property
    def kernel_launches(self) -> int:
        return int(self.lib.guac_kernel_launches(self.ctx))

    # -- reference ---------------------------------------------------------------
    def upload_reference(self, reference: Reference, contig_names: Sequence[str]):
        self.contig_names = list(contig_names)
        for cid, name in enumerate(contig_names):
            if name not in reference.contigs:
                continue
            w = reference.contigs[name]
            self._check(self.lib.guac_ref_reserve(self.ctx, C.c_int32(cid), C.c_int64(w.offset), C.c_int64(w.bases.shape[0])))
            # upload only the present runs so that absent loci stay "missing"
            present = w.bases != 0
            edges = np.flatnonzero(np.diff(np.concatenate([[0], present.view(np.int8), [0]])))
            for s, e in zip(edges[0::2], edges[1::2]):
                seg = np.ascontiguousarray(w.bases[s:e])
                self._check(self.lib.guac_ref_upload(self.ctx, C.c_int32(cid), ptr(seg, u8p), C.c_int64(e - s),
                                                     C.c_int64(w.offset + int(s))))

    # -- reads -------------------------------------------------------------------
    def upload_reads(self, batch: ReadBatch):
        b = batch_struct(batch)
        self._check(self.lib.guac_reads_upload(self.ctx, C.byref(b)))
        self._batch = batch  # keep host arrays alive

    def upload_reads_struct(self, b: ReadBatchC, n_samples: int):
        """Uploads an already packed batch (e.g. pinned buffers from export_reads_struct)."""
        self._check(self.lib.guac_reads_upload(self.ctx, C.byref(b)))
        self._batch = type("B", (), {"n_samples": n_samples})()

    def upload_reads_compact(self, batch):
        """guac_reads_upload_compact: a ReadBatch (packed here) or an already packed ReadBatchCompactC."""
        b = batch if isinstance(batch, ReadBatchCompactC) else compact_struct(batch)
        self._check(self.lib.guac_reads_upload_compact(self.ctx, C.byref(b)))
        self._batch = batch if not isinstance(batch, ReadBatchCompactC) else type("B", (), {"n_samples": b.n_samples})()

    def export_reads_compact_struct(self) -> ReadBatchCompactC:
        """Device batch -> compact wire form in library-owned pinned host buffers."""
        b = ReadBatchCompactC()
        self._check(self.lib.guac_reads_export_compact(self.ctx, C.byref(b)))
        return b

    def free_reads_compact_struct(self, b: ReadBatchCompactC):
        for name in ("run_contig_id", "run_first_read", "start", "read_len", "n_cigar_ops", "cigar", "bases2", "exc_pos", "exc_base",
                     "qual", "mapq", "flags", "sample_id", "run_start_base", "start_rel", "n_cigar_ops8", "mm_count", "mm_words"):
            v = getattr(b, name)
            if v:
                self.lib.guac_free(self.ctx, C.cast(v, C.c_void_p))

    # -- measurement support -----------------------------------------------------
    def synth_generate(self, contig_len: int, depths: Sequence[int], tumor: Sequence[bool], seed: int, contig_id: int = 0,
                       contig_name: str = "synth1", profile: int = 0):
        self.set_option("synth_profile", profile)
        d = np.asarray(depths, np.int32)
        t = np.asarray(tumor, np.uint8)
        self._check(self.lib.guac_synth_generate(self.ctx, C.c_int32(contig_id), C.c_int64(contig_len), C.c_int32(len(d)),
                                                 ptr(d, i32p), ptr(t, u8p), C.c_uint64(seed)))
        self.contig_names = [contig_name]
        self._batch = type("B", (), {"n_samples": len(d)})()

    def export_reads_struct(self) -> ReadBatchC:
        """Device batch -> library-owned pinned host buffers (still packed)."""
        b = ReadBatchC()
        self._check(self.lib.guac_reads_export(self.ctx, C.byref(b)))
        return b

    def free_reads_struct(self, b: ReadBatchC):
        for name, _ in ReadBatchC._fields_:
            v = getattr(b, name)
            if not isinstance(v, int) and v:
                self.lib.guac_free(self.ctx, C.cast(v, C.c_void_p))

    def export_reads(self, contig_names: Optional[List[str]] = None, lo: int = 0, hi: Optional[int] = None) -> ReadBatch:
        """Reads [lo, hi) of the device batch as a host ReadBatch (ASCII bases unpacked)."""
        b = self.export_reads_struct()
        try:
            n = b.n_reads
            hi = n if hi is None else min(hi, n)
            lo = max(0, min(lo, hi))

            def arr(p, dtype, count, off=0):
                if count == 0:
                    return np.zeros(0, dtype)
                a = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=((off + count) * np.dtype(dtype).itemsize,))
                return a.view(dtype)[off:off + count].copy()

            seq_off = arr(b.seq_off, np.uint64, hi - lo + 1, lo)
            cigar_off = arr(b.cigar_off, np.uint32, hi - lo + 1, lo)
            s0, s1 = int(seq_off[0]), int(seq_off[-1])
            c0, c1 = int(cigar_off[0]), int(cigar_off[-1])
            assert s0 % 8 == 0, "export window must start on an 8-base boundary"
            bases2 = arr(b.bases2, np.uint8, (s1 - s0 + 3) // 4, s0 // 4)
            codes = ((bases2[:, None] >> np.array([0, 2, 4, 6], np.uint8)[None, :]) & 3).reshape(-1)[:s1 - s0]
            ascii_bases = np.frombuffer(b"ACGT", np.uint8)[codes].copy()
            if b.n_exc:
                pos = arr(b.exc_pos, np.uint64, b.n_exc).astype(np.int64)
                eb = arr(b.exc_base, np.uint8, b.n_exc)
                m = (pos >= s0) & (pos < s1)
                ascii_bases[pos[m] - s0] = eb[m]
            return ReadBatch(n_samples=b.n_samples, contig_names=list(contig_names or self.contig_names),
                             contig_id=arr(b.contig_id, np.int32, hi - lo, lo), start=arr(b.start, np.int64, hi - lo, lo),
                             cigar_off=(cigar_off - np.uint32(c0)).astype(np.uint32), cigar=arr(b.cigar, np.uint32, c1 - c0, c0),
                             seq_off=seq_off - np.uint64(s0), bases_ascii=ascii_bases, qual=arr(b.qual, np.uint8, s1 - s0, s0),
                             mapq=arr(b.mapq, np.uint8, hi - lo, lo), flags=arr(b.flags, np.uint8, hi - lo, lo),
                             sample_id=arr(b.sample_id, np.uint8, hi - lo, lo))
        finally:
            self.free_reads_struct(b)

    def export_reference(self, contig_id: int, offset: int, length: int) -> np.ndarray:
        out = np.zeros(length, np.uint8)
        self._check(self.lib.guac_ref_export(self.ctx, C.c_int32(contig_id), C.c_int64(offset), C.c_int64(length), ptr(out, u8p)))
        return out

    @property
    def stream(self) -> int:
        self.lib.guac_stream.restype = C.c_void_p
        self.lib.guac_stream.argtypes = [C.c_void_p]
        return int(self.lib.guac_stream(self.ctx) or 0)

    def set_option(self, name: str, value: int):
        self._check(self.lib.guac_set_option(self.ctx, name.encode(), C.c_int64(value)))

    def last_stat(self, name: str) -> int:
        """guac_last_stat: a counter of the most recent caller invocation ("fallback_tiles", "heavy_loci", "sj_simple", ...)."""
        self.lib.guac_last_stat.restype = C.c_int64
        self.lib.guac_last_stat.argtypes = [C.c_void_p, C.c_char_p]
        return int(self.lib.guac_last_stat(self.ctx, name.encode()))

    @property
    def last_fallback_tiles(self) -> int:
        self.lib.guac_last_fallback_tiles.restype = C.c_int64
        self.lib.guac_last_fallback_tiles.argtypes = [C.c_void_p]
        return int(self.lib.guac_last_fallback_tiles(self.ctx))

    def last_kernel_ms(self) -> Tuple[float, int]:
        ms, n = C.c_float(), C.c_int64()
        self._check(self.lib.guac_last_kernel_ms(self.ctx, C.byref(ms), C.byref(n)))
        return float(ms.value), int(n.value)

    # -- pileup ------------------------------------------------------------------
    def pileup_counts(self, ranges: Sequence[Tuple[int, int, int]], skip_empty: bool = True):
        r = make_ranges(ranges)
        oc, ol, cnt, n = i32p(), i64p(), C.POINTER(LocusCountsC)(), C.c_int64()
        self._check(self.lib.guac_pileup_counts(self.ctx, r, C.c_int32(len(ranges)), C.c_int32(int(skip_empty)), C.byref(oc),
                                                C.byref(ol), C.byref(cnt), C.byref(n)))
        ns = self._batch.n_samples
        return (_take(self.lib, self.ctx, oc, np.int32, n.value), _take(self.lib, self.ctx, ol, np.int64, n.value),
                _take(self.lib, self.ctx, cnt, LOCUS_COUNTS_DTYPE, n.value * ns).reshape(n.value, ns))

    def pileup_elements(self, ranges: Sequence[Tuple[int, int, int]], skip_empty: bool = True):
        r = make_ranges(ranges)
        oc, ol, off, els, n = i32p(), i64p(), i64p(), C.POINTER(ElementC)(), C.c_int64()
        self._check(self.lib.guac_pileup_elements(self.ctx, r, C.c_int32(len(ranges)), C.c_int32(int(skip_empty)), C.byref(oc),
                                                  C.byref(ol), C.byref(off), C.byref(els), C.byref(n)))
        offsets = _take(self.lib, self.ctx, off, np.int64, n.value + 1)
        total = int(offsets[-1]) if offsets.shape[0] else 0
        return (_take(self.lib, self.ctx, oc, np.int32, n.value), _take(self.lib, self.ctx, ol, np.int64, n.value), offsets,
                _take(self.lib, self.ctx, els, ELEMENT_DTYPE, total))

    def germline_standard(self, ranges: Sequence[Tuple[int, int, int]], params: Optional["GSParamsC"] = None):
        """guac_germline_standard -> dict(calls, arena, active_loci [(contig, locus)], n_visited)."""
        params = params or gs_params()
        r = make_ranges(ranges)
        calls, n, arena, alen = C.c_void_p(), C.c_int64(), C.c_void_p(), C.c_int64()
        act, na, nv = C.c_void_p(), C.c_int64(), C.c_int64()
        self._check(self.lib.guac_germline_standard(self.ctx, C.byref(params), r, C.c_int32(len(ranges)), C.byref(calls), C.byref(n),
                                                    C.byref(arena), C.byref(alen), C.byref(act), C.byref(na), C.byref(nv)))
        arena_b = C.string_at(arena, alen.value) if arena else b""
        if arena:
            self.lib.guac_free(self.ctx, arena)
        active = _take(self.lib, self.ctx, act, np.int64, na.value * 2).reshape(-1, 2)
        return dict(calls=_take(self.lib, self.ctx, calls, CALLED_ALLELE_DTYPE, n.value), arena=arena_b,
                    active_loci=[(int(a), int(b)) for a, b in active], n_visited=nv.value)

    def germline_assembly(self, ranges: Sequence[Tuple[int, int, int]], params: Optional["GAParamsC"] = None):
        """guac_germline_assembly -> dict(calls, arena, windows, n_visited); assembled calls have n_genotypes == 0."""
        params = params or ga_params()
        r = make_ranges(ranges)
        calls, n, arena, alen = C.c_void_p(), C.c_int64(), C.c_void_p(), C.c_int64()
        wins, nw, nv = C.c_void_p(), C.c_int64(), C.c_int64()
        self._check(self.lib.guac_germline_assembly(self.ctx, C.byref(params), r, C.c_int32(len(ranges)), C.byref(calls), C.byref(n),
                                                    C.byref(arena), C.byref(alen), C.byref(wins), C.byref(nw), C.byref(nv)))
        arena_b = C.string_at(arena, alen.value) if arena else b""
        if arena:
            self.lib.guac_free(self.ctx, arena)
        return dict(calls=_take(self.lib, self.ctx, calls, CALLED_ALLELE_DTYPE, n.value), arena=arena_b,
                    windows=_take(self.lib, self.ctx, wins, ASSEMBLED_WINDOW_DTYPE, nw.value), n_visited=nv.value)

    def discover_paths(self, reference_sequence: str, kmer_size: int, min_occurrence: int = 1, max_paths: int = 10,
                       min_mean_kmer_quality: int = 0, sample_id: int = 0):
        """guac_discover_paths over the uploaded reads -> haplotype strings, most recently found first."""
        ref = reference_sequence.encode()
        out, n, npaths = C.c_void_p(), C.c_int64(), C.c_int32()
        self._check(self.lib.guac_discover_paths(self.ctx, C.c_int32(sample_id), ref, C.c_int32(len(ref)), C.c_int32(kmer_size),
                                                 C.c_int32(min_occurrence), C.c_int32(max_paths), C.c_int32(min_mean_kmer_quality),
                                                 C.byref(out), C.byref(n), C.byref(npaths)))
        text = C.string_at(out, n.value) if n.value else b""
        self.lib.guac_free(self.ctx, out)
        return [t.decode() for t in text.split(b"\0")[:npaths.value]]

    def read_alignment(self, sequence: str, reference: str, mismatch_probability: float = 0.0, open_gap_probability: float = 0.0,
                       close_gap_probability: float = 0.0):
        """guac_read_alignment -> (cigar, refStartIdx, refEndIdx, alignmentScore, rawScore)."""
        s, r = sequence.encode(), reference.encode()
        buf = C.create_string_buffer(4096)
        rs, re_, sc, raw = C.c_int32(), C.c_int32(), C.c_int32(), C.c_double()
        self._check(self.lib.guac_read_alignment(self.ctx, s, C.c_int32(len(s)), r, C.c_int32(len(r)), C.c_double(mismatch_probability),
                                                 C.c_double(open_gap_probability), C.c_double(close_gap_probability), buf, C.c_int32(4096),
                                                 C.byref(rs), C.byref(re_), C.byref(sc), C.byref(raw)))
        return buf.value.decode(), rs.value, re_.value, sc.value, raw.value

    def somatic_joint_records(self, loci: Sequence[Tuple[int, int, int]], params, force_loci: Sequence[Tuple[int, int, int]] = (),
                              copy: bool = True):
        """guac_somatic_joint -> (allele records, sample-evidence records, arena bytes, loci visited) as numpy arrays.

        copy=False: the two record arrays are views of the library's pinned output buffers (what a JNI / C++ consumer reads in
        place), valid until the next call with copy=False or `release_outputs()`; nothing is copied on the host."""
        from .commands import somatic_joint as sj
        self.release_outputs()
        r, f = make_ranges(loci), make_ranges(force_loci)
        al, na, sm, ns, arena, alen, nv = C.c_void_p(), C.c_int64(), C.c_void_p(), C.c_int64(), C.c_void_p(), C.c_int64(), C.c_int64()
        self._check(self.lib.guac_somatic_joint(self.ctx, C.byref(params), r, C.c_int32(len(loci)), f, C.c_int32(len(force_loci)),
                                                C.byref(al), C.byref(na), C.byref(sm), C.byref(ns), C.byref(arena), C.byref(alen),
                                                C.byref(nv)))
        arena_b = C.string_at(arena, alen.value) if arena else b""
        if arena:
            self.lib.guac_free(self.ctx, arena)
        if copy:
            alleles = _take(self.lib, self.ctx, al, sj.ALLELE_EVIDENCE_DTYPE, na.value)
            samples = _take(self.lib, self.ctx, sm, sj.SAMPLE_EVIDENCE_DTYPE, ns.value)
        else:
            def view(p, dtype, n):
                if not p or n == 0:
                    return np.zeros(0, dtype)
                return np.frombuffer((C.c_char * (int(n) * dtype.itemsize)).from_address(p.value), dtype=dtype)
            alleles, samples = view(al, sj.ALLELE_EVIDENCE_DTYPE, na.value), view(sm, sj.SAMPLE_EVIDENCE_DTYPE, ns.value)
            self._held_outputs = [x for x in (al, sm) if x]
        return alleles, samples, arena_b, nv.value

    def release_outputs(self):
        """Hands the pinned output buffers of a copy=False call back to the library (guac_free)."""
        for p in getattr(self, "_held_outputs", []):
            self.lib.guac_free(self.ctx, p)
        self._held_outputs = []

    def somatic_joint(self, loci: Sequence[Tuple[int, int, int]], params, force_loci: Sequence[Tuple[int, int, int]] = ()):
        """guac_somatic_joint -> (list of per-allele evidence dicts, loci visited); see commands/somatic_joint.py."""
        from .commands import somatic_joint as sj
        alleles, samples, arena_b, nv = self.somatic_joint_records(loci, params, force_loci)
        return sj.decode(alleles, samples, arena_b, params.n_samples), nv

    # -- other consumers of the pileup operator ---------------------------------------
    def variant_support(self, loci: Sequence[Tuple[int, int, int]]):
        """guac_variant_support -> [(sample_id, contig_id, locus, ref, alt, count)] (commands/VariantSupport.scala:100-112)."""
        r = make_ranges(loci)
        out, n, arena, alen = C.c_void_p(), C.c_int64(), C.c_void_p(), C.c_int64()
        self._check(self.lib.guac_variant_support(self.ctx, r, C.c_int32(len(loci)), C.byref(out), C.byref(n), C.byref(arena), C.byref(alen)))
        a = C.string_at(arena, alen.value) if arena else b""
        if arena:
            self.lib.guac_free(self.ctx, arena)
        rec = _take(self.lib, self.ctx, out, ALLELE_COUNT_DTYPE, n.value)
        return [(int(x["sample_id"]), int(x["contig_id"]), int(x["locus"]), a[x["ref_off"]:x["ref_off"] + x["ref_len"]].decode(),
                 a[x["alt_off"]:x["alt_off"] + x["alt_len"]].decode(), int(x["count"])) for x in rec]

    def vaf_histogram(self, loci: Sequence[Tuple[int, int, int]], bins: int = 20, min_read_depth: int = 0, min_vaf_percent: int = 0,
                      want_loci: bool = True):
        """guac_vaf_histogram -> (hist [n_samples, 101], variant loci per sample, VariantLocus records or None)
        (commands/VAFHistogram.scala:194-245)."""
        S = self._batch.n_samples
        r = make_ranges(loci)
        hist, tot = np.zeros((S, 101), np.int64), np.zeros(S, np.int64)
        out, n = C.c_void_p(), C.c_int64()
        self._check(self.lib.guac_vaf_histogram(self.ctx, r, C.c_int32(len(loci)), C.c_int32(bins), C.c_int32(min_read_depth),
                                                C.c_int32(min_vaf_percent), hist.ctypes.data_as(C.c_void_p), tot.ctypes.data_as(C.c_void_p),
                                                C.byref(out) if want_loci else None, C.byref(n)))
        return hist, tot, (_take(self.lib, self.ctx, out, VARIANT_LOCUS_DTYPE, n.value) if want_loci else None)

    # -- coverage-aware partitioning -----------------------------------------------------
    def _intervals(self, out, part, n):
        r = _take(self.lib, self.ctx, out, np.dtype([("contig_id", np.int32), ("_pad", np.int32), ("start", np.int64), ("end", np.int64)]), n)
        p = _take(self.lib, self.ctx, part, np.int32, n)
        return [(int(a["contig_id"]), int(a["start"]), int(a["end"]), int(b)) for a, b in zip(r, p)]

    def take_loci(self, coverages, max_regions_per_partition: int, trim_ranges: bool = False):
        """guac_take_loci: [(contig_id, locus, depth, starts)] -> [(contig_id, start, end, partition)] (loci/set/TakeLociIterator.scala)."""
        c = np.asarray([x[0] for x in coverages], np.int32)
        l = np.asarray([x[1] for x in coverages], np.int64)
        d = np.asarray([x[2] for x in coverages], np.int32)
        s = np.asarray([x[3] for x in coverages], np.int32)
        out, part, n, npart = C.c_void_p(), C.c_void_p(), C.c_int64(), C.c_int32()
        self._check(self.lib.guac_take_loci(self.ctx, c.ctypes.data_as(C.c_void_p), l.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p),
                                            s.ctypes.data_as(C.c_void_p), C.c_int64(len(c)), C.c_int32(max_regions_per_partition),
                                            C.c_int32(int(trim_ranges)), C.byref(out), C.byref(part), C.byref(n), C.byref(npart)))
        return self._intervals(out, part, n.value), npart.value

    def capped_partitioning(self, loci: Sequence[Tuple[int, int, int]], half_window: int = 0, max_reads_per_partition: int = 100000,
                            trim_ranges: bool = False):
        """guac_capped_partitioning -> ([(contig_id, start, end, partition)], number of partitions)
        (loci/partitioning/CappedRegionsPartitioner.scala:66-88)."""
        r = make_ranges(loci)
        out, part, n, npart = C.c_void_p(), C.c_void_p(), C.c_int64(), C.c_int32()
        self._check(self.lib.guac_capped_partitioning(self.ctx, r, C.c_int32(len(loci)), C.c_int32(half_window),
                                                      C.c_int32(max_reads_per_partition), C.c_int32(int(trim_ranges)), C.byref(out),
                                                      C.byref(part), C.byref(n), C.byref(npart)))
        return self._intervals(out, part, n.value), npart.value

    # -- callers -----------------------------------------------------------------
    def somatic_standard(self, ranges: Sequence[Tuple[int, int, int]], params: Optional[SSParamsC] = None,
                         want_summary: bool = False):
        params = params or ss_params()
        r = make_ranges(ranges)
        calls, n, arena, alen = C.POINTER(SSCallC)(), C.c_int64(), C.c_char_p(), C.c_int64()
        summ, nv = C.POINTER(LocusSummaryC)(), C.c_int64()
        arena_p = C.c_void_p()
        self._check(self.lib.guac_somatic_standard(self.ctx, C.byref(params), r, C.c_int32(len(ranges)), C.byref(calls),
                                                   C.byref(n), C.byref(arena_p), C.byref(alen),
                                                   C.byref(summ) if want_summary else None, C.byref(nv)))
        arena_b = C.string_at(arena_p, alen.value) if arena_p else b""
        if arena_p:
            self.lib.guac_free(self.ctx, arena_p)
        out = dict(calls=_take(self.lib, self.ctx, calls, SS_CALL_DTYPE, n.value), arena=arena_b, n_visited=nv.value)
        if want_summary:
            out["summary"] = _take(self.lib, self.ctx, summ, LOCUS_SUMMARY_DTYPE, nv.value * 2).reshape(nv.value, 2)
        return out
