"""ctypes binding of the library's host stages (include/guac_b200.h, "host stages either side of the device path"): the
C++ read packer with the ingest filters, the somatic genotype filter chain and the two VCF row writers.  None of them needs a
GPU; a Scala shim binds the same entry points."""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import numpy as np

from .native import ReadBatchCompactC, RangeC, load_library, make_ranges
from .reads import FLAG_DUPLICATE, FLAG_FAILED_QC, FLAG_PAIRED, FLAG_POSITIVE_STRAND, ReadBatch


class SamRecordsC(C.Structure):
    _fields_ = [("n_records", C.c_int64), ("n_samples", C.c_int32), ("_pad", C.c_int32), ("sam_flags", C.c_void_p),
                ("reference_index", C.c_void_p), ("alignment_start", C.c_void_p), ("mapping_quality", C.c_void_p), ("cigar_off", C.c_void_p),
                ("cigar", C.c_void_p), ("seq_off", C.c_void_p), ("bases_ascii", C.c_void_p), ("base_qualities", C.c_void_p),
                ("sample_id", C.c_void_p)]


class ReadFiltersC(C.Structure):
    _fields_ = [("non_duplicate", C.c_int32), ("passed_vendor_quality_checks", C.c_int32), ("is_paired", C.c_int32),
                ("min_alignment_quality", C.c_int32), ("overlaps_loci", C.c_void_p), ("n_overlaps_loci", C.c_int32), ("_pad", C.c_int32)]


class SSFilterParamsC(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("min_likelihood", "min_vaf", "min_average_mapping_quality", "min_average_base_quality",
                                          "min_tumor_read_depth", "min_normal_read_depth", "max_tumor_read_depth",
                                          "min_tumor_alternate_read_depth", "maximum_median_mismatches", "_pad")]


def ss_filter_params(**kw) -> SSFilterParamsC:
    """SomaticGenotypeFilterArguments defaults (filters/somatic/SomaticGenotypeFilter.scala:34-80); -1 = not set."""
    p = SSFilterParamsC(-1, 0, 20, 20, 5, 5, 2**31 - 1, -1, -1, 0)
    for k, v in kw.items():
        if not hasattr(p, k):
            raise AttributeError(k)
        setattr(p, k, -1 if v is None else v)
    return p


def sam_records_of(batch: ReadBatch, extra_flags: Optional[np.ndarray] = None):
    """The SAM-record columns a shim would hand over for the reads of a ReadBatch (1-based starts, SAM FLAG bits);
    extra_flags is OR-ed in (e.g. 0x4 to mark records unmapped).  Returns (struct, keep-alive list)."""
    fl = batch.flags.astype(np.uint16)
    sam = (np.where(fl & FLAG_POSITIVE_STRAND, 0, 0x10) | np.where(fl & FLAG_PAIRED, 0x1, 0) | np.where(fl & FLAG_DUPLICATE, 0x400, 0) |
           np.where(fl & FLAG_FAILED_QC, 0x200, 0)).astype(np.uint16)
    if extra_flags is not None:
        sam |= extra_flags.astype(np.uint16)
    cols = dict(sam_flags=sam, reference_index=batch.contig_id.astype(np.int32), alignment_start=(batch.start + 1).astype(np.int64),
                mapping_quality=batch.mapq.astype(np.uint8), cigar_off=batch.cigar_off.astype(np.uint32), cigar=batch.cigar.astype(np.uint32),
                seq_off=batch.seq_off.astype(np.uint64), bases_ascii=batch.bases_ascii.astype(np.uint8), base_qualities=batch.qual.astype(np.uint8),
                sample_id=batch.sample_id.astype(np.uint8))
    cols = {k: np.ascontiguousarray(v) for k, v in cols.items()}
    s = SamRecordsC()
    s.n_records, s.n_samples = batch.n_reads, batch.n_samples
    for k, v in cols.items():
        setattr(s, k, v.ctypes.data)
    return s, list(cols.values())


class HostReferenceC(C.Structure):
    """guac_host_reference: one contig window of the reference in host memory."""
    _fields_ = [("contig_id", C.c_int32), ("_pad", C.c_int32), ("offset", C.c_int64), ("len", C.c_int64), ("bases", C.c_void_p)]


def host_references(ref, contig_names: Sequence[str]):
    """(array of guac_host_reference, keep-alive list) for the windows of a `Reference` whose contigs are in contig_names."""
    windows = [(i, ref.contigs[c]) for i, c in enumerate(contig_names) if c in ref.contigs]
    arr = (HostReferenceC * max(1, len(windows)))()
    keep = []
    for k, (i, w) in enumerate(windows):
        bases = np.ascontiguousarray(w.bases, np.uint8)
        keep.append(bases)
        arr[k] = HostReferenceC(i, 0, int(w.offset), int(bases.shape[0]), bases.ctypes.data)
    return arr, len(windows), keep


def pack_reads(records: SamRecordsC, ctx=None, non_duplicate=False, passed_vendor_quality_checks=False, is_paired=False,
               min_alignment_quality=0, overlaps_loci: Sequence[Tuple[int, int, int]] = (), n_threads: int = 0, ref=None,
               contig_names: Sequence[str] = ()):
    """guac_pack_reads -> (ReadBatchCompactC, reads kept).  ctx: a native.Context (pinned outputs) or None (C heap).
    ref (a `Reference`, with contig_names giving the contig ids): guac_pack_reads_ref — reference-relative bases."""
    lib = load_library()
    f = ReadFiltersC(int(non_duplicate), int(passed_vendor_quality_checks), int(is_paired), int(min_alignment_quality), None, 0, 0)
    ranges = make_ranges(overlaps_loci) if overlaps_loci else None
    if ranges is not None:
        f.overlaps_loci, f.n_overlaps_loci = C.cast(ranges, C.c_void_p), len(overlaps_loci)
    out, n = ReadBatchCompactC(), C.c_int64()
    if ref is not None:
        refs, n_refs, keep = host_references(ref, contig_names)
        lib.guac_pack_reads_ref.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]
        st = lib.guac_pack_reads_ref(ctx.ctx if ctx is not None else None, C.byref(records), C.byref(f), refs, C.c_int32(n_refs),
                                     C.c_int32(n_threads), C.byref(out), C.byref(n))
        del keep
    else:
        lib.guac_pack_reads.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]
        st = lib.guac_pack_reads(ctx.ctx if ctx is not None else None, C.byref(records), C.byref(f), C.c_int32(n_threads), C.byref(out),
                                 C.byref(n))
    if st != 0:
        raise RuntimeError(f"guac_pack_reads failed ({st})" + (": " + lib.guac_last_error(ctx.ctx).decode() if ctx is not None else ""))
    return out, n.value


def free_packed(out: ReadBatchCompactC, ctx=None):
    lib = load_library()
    lib.guac_host_free.argtypes = [C.c_void_p]
    lib.guac_host_free.restype = None
    for name, _ in ReadBatchCompactC._fields_:
        v = getattr(out, name)
        if not isinstance(v, int) and not isinstance(v, C.Array) and v:
            p = C.cast(v, C.c_void_p)
            if ctx is not None:
                lib.guac_free(ctx.ctx, p)
            else:
                lib.guac_host_free(p)


def _text_call(fn, *args) -> str:
    need = C.c_int64()
    fn(*args, None, C.c_int64(0), C.byref(need))
    buf = C.create_string_buffer(need.value + 1)
    st = fn(*args, buf, C.c_int64(need.value), C.byref(need))
    if st != 0:
        raise RuntimeError(f"text call failed ({st})")
    return buf.raw[:need.value].decode()


def _names(names: Sequence[str]):
    arr = (C.c_char_p * len(names))(*[n.encode() for n in names])
    return arr


def ss_genotype_filter(calls: np.ndarray, params: Optional[SSFilterParamsC] = None) -> np.ndarray:
    """guac_ss_genotype_filter -> boolean keep mask (filters/somatic/SomaticGenotypeFilter.scala:82-130)."""
    lib = load_library()
    params = params or ss_filter_params()
    calls = np.ascontiguousarray(calls)
    keep = np.zeros(len(calls), np.uint8)
    st = lib.guac_ss_genotype_filter(C.byref(params), calls.ctypes.data_as(C.c_void_p), C.c_int64(len(calls)), keep.ctypes.data_as(C.c_void_p))
    assert st == 0
    return keep.astype(bool)


def ss_vcf_rows(result: dict, contig_names: Sequence[str], keep: Optional[np.ndarray] = None, sample_name: str = "tumor") -> List[str]:
    lib = load_library()
    calls = np.ascontiguousarray(result["calls"])
    k = None if keep is None else np.ascontiguousarray(keep.astype(np.uint8))
    text = _text_call(lib.guac_ss_vcf_rows, calls.ctypes.data_as(C.c_void_p), C.c_int64(len(calls)),
                      k.ctypes.data_as(C.c_void_p) if k is not None else None, C.c_char_p(result["arena"]), _names(contig_names),
                      C.c_char_p(sample_name.encode()))
    return text.rstrip("\n").split("\n")


def sj_vcf_rows(alleles: np.ndarray, samples: np.ndarray, arena: bytes, params, contig_names: Sequence[str], sample_names: Sequence[str],
                include_pooled_normal: int = -1, include_pooled_tumor: int = -1, policy: int = 0) -> List[str]:
    lib = load_library()
    alleles, samples = np.ascontiguousarray(alleles), np.ascontiguousarray(samples)
    text = _text_call(lib.guac_sj_vcf_rows, C.byref(params), alleles.ctypes.data_as(C.c_void_p), C.c_int64(len(alleles)),
                      samples.ctypes.data_as(C.c_void_p), C.c_char_p(arena), _names(contig_names), _names(sample_names),
                      C.c_int32(include_pooled_normal), C.c_int32(include_pooled_tumor), C.c_int32(policy))
    return text.rstrip("\n").split("\n")


def format_double(kind: int, x: float) -> str:
    lib = load_library()
    buf = C.create_string_buffer(64)
    assert lib.guac_format_double(C.c_int32(kind), C.c_double(x), buf, C.c_int32(64)) == 0
    return buf.value.decode()
