"""guac_pileup_counts_noalloc: the pileup-counting entry point for hosts that are not allowed to
allocate GPU memory when they call (their framework owns the whole device pool).  It works out of
a fixed array inside the library, so this test
checks exactly what such a caller relies on: given device buffers it already owns, two calls in a
row — the second on different, smaller input, so stale working memory would show — must each
return what the allocating entry point guac_pileup_counts returns for the same reads, bit for bit.

Collected last on purpose (file name): if it fails under `-x`, every other test has already run.
"""
import ctypes as C

import numpy as np
import pytest

from guacamole_b200 import synth
from guacamole_b200.native import LOCUS_COUNTS_DTYPE, ReadBatchC, batch_struct, load_library

pytestmark = pytest.mark.gpu


def device_batch(torch, batch):
    """The caller's own device copies of the read columns (allocated by the caller, not the library)."""
    host = batch_struct(batch)
    keep = []
    dev = ReadBatchC()
    dev.n_reads, dev.n_samples, dev.n_exc = host.n_reads, host.n_samples, host.n_exc
    cols = dict(contig_id=batch.contig_id, start=batch.start, cigar_off=batch.cigar_off, cigar=batch.cigar, seq_off=batch.seq_off,
                bases2=batch.packed["bases2"], nmask=batch.packed["nmask"], exc_pos=batch.packed["exc_pos"],
                exc_base=batch.packed["exc_base"], qual=batch.qual, mapq=batch.mapq, flags=batch.flags, sample_id=batch.sample_id)
    for name, arr in cols.items():
        raw = np.ascontiguousarray(arr).view(np.uint8).reshape(-1)
        t = torch.from_numpy(raw.copy()).cuda()  # exactly the size guac_read_batch states: no padding is required
        keep.append(t)
        ftype = dict(ReadBatchC._fields_)[name]
        setattr(dev, name, C.cast(C.c_void_p(t.data_ptr()), ftype))
    return dev, keep


def run_noalloc(torch, lib, batch, ref_bases, contig_len, lo, hi):
    dev, keep = device_batch(torch, batch)
    ref_t = torch.from_numpy(np.ascontiguousarray(ref_bases)).cuda()
    out_t = torch.full(((hi - lo) * batch.n_samples * LOCUS_COUNTS_DTYPE.itemsize,), 0xAB, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    st = lib.guac_pileup_counts_noalloc(C.c_int(0), C.byref(dev), C.c_void_p(ref_t.data_ptr()), C.c_int64(0), C.c_int64(contig_len),
                                        C.c_int32(0), C.c_int64(lo), C.c_int64(hi), C.c_void_p(out_t.data_ptr()), C.c_void_p(stream))
    assert st == 0, lib.guac_noalloc_last_error().decode()
    torch.cuda.synchronize()
    return out_t.cpu().numpy().view(LOCUS_COUNTS_DTYPE).reshape(hi - lo, batch.n_samples)


def test_no_alloc_matches_allocating_entry_point_twice(gpu_ctx):
    import torch
    lib = load_library()
    lib.guac_noalloc_last_error.restype = C.c_char_p
    lib.guac_noalloc_max_loci.restype = C.c_int64
    lib.guac_noalloc_max_reads.restype = C.c_int64

    cases = [(30000, 21, 500, 29500), (9000, 33, 0, 8000)]  # (contig length, seed, range): big first, then smaller
    for contig_len, seed, lo, hi in cases:
        ref, batch = synth.make_tumor_normal(contig_len=contig_len, seed=seed)
        assert batch.n_reads <= lib.guac_noalloc_max_reads()
        got = run_noalloc(torch, lib, batch, ref.contigs["synth1"].bases, contig_len, lo, hi)
        gpu_ctx.upload_reference(ref, batch.contig_names)
        gpu_ctx.upload_reads(batch)
        _, loci, want = gpu_ctx.pileup_counts([(0, lo, hi)], skip_empty=False)
        assert list(loci) == list(range(lo, hi))
        assert np.array_equal(got, want)  # integer counts: bit-exact, the tolerance of the other pileup tests
        assert int(got["depth"].sum()) > 0

    # sizes beyond the fixed working memory are refused up front, with a message that names the limit
    ref, batch = synth.make_tumor_normal(contig_len=2000, seed=3)
    dev, keep = device_batch(torch, batch)
    too_many = lib.guac_noalloc_max_loci() + 1
    st = lib.guac_pileup_counts_noalloc(C.c_int(0), C.byref(dev), C.c_void_p(keep[0].data_ptr()), C.c_int64(0), C.c_int64(2000),
                                        C.c_int32(0), C.c_int64(0), C.c_int64(too_many), C.c_void_p(keep[0].data_ptr()), None)
    assert st == -5 and b"loci exceed the fixed working memory" in lib.guac_noalloc_last_error()
