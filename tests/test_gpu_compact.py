"""The compact wire form (guac_read_batch_compact) is lossless: a batch uploaded through it is, on the device, the very
batch guac_reads_upload would have produced, and the callers return the same records."""
import ctypes as C

import numpy as np
import pytest

from guacamole_b200 import synth
from guacamole_b200.native import GuacError, compact_struct, ss_params
from guacamole_b200.reads import ReadBatch, make_read
from tests.golden import fixtures

pytestmark = pytest.mark.gpu

COLUMNS = ["contig_id", "start", "cigar_off", "cigar", "seq_off", "bases_ascii", "qual", "mapq", "flags", "sample_id"]


def assert_same_batch(a: ReadBatch, b: ReadBatch):
    assert a.n_reads == b.n_reads and a.n_samples == b.n_samples
    for name in COLUMNS:
        assert np.array_equal(getattr(a, name), getattr(b, name)), name


def device_copy(ctx, names):
    return ctx.export_reads(names)


@pytest.fixture(params=["narrow", "wide"])
def narrow(request, gpu_ctx):
    """Both layouts of the wire form: with the narrow start / operator-count / read-length columns, and without."""
    gpu_ctx.set_option("compact_narrow", 1 if request.param == "narrow" else 0)
    yield request.param == "narrow"
    gpu_ctx.set_option("compact_narrow", 1)


@pytest.mark.parametrize("name", ["tough", "tiny_sjc", "testrna"])
def test_real_reads_round_trip(gpu_ctx, name, narrow):
    """Real reads: > 16 distinct qualities (one byte per base on the wire), N bases, soft clips, indels, several contigs."""
    batch, ref, _ = fixtures.load(name)
    if ref is not None:  # the RNA fixture has no reference; uploads do not need one
        gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    plain = device_copy(gpu_ctx, batch.contig_names)
    c = compact_struct(batch, narrow=narrow)
    assert bool(c.start_rel) == narrow and bool(c.n_cigar_ops8) == narrow
    gpu_ctx.upload_reads_compact(c)
    assert_same_batch(device_copy(gpu_ctx, batch.contig_names), plain)
    assert_same_batch(plain, batch)
    e = gpu_ctx.export_reads_compact_struct()  # the device's own packing of the same batch, and back in
    try:
        assert bool(e.start_rel) == narrow and bool(e.n_cigar_ops8) == narrow and e.uniform_read_len == c.uniform_read_len
        gpu_ctx.upload_reads_compact(e)
        assert_same_batch(device_copy(gpu_ctx, batch.contig_names), batch)
    finally:
        gpu_ctx.free_reads_compact_struct(e)


def test_binned_qualities_travel_as_codes(gpu_ctx, narrow):
    contig_len, seed = 30000, 4005
    ref = synth.make_reference(contig_len, seed)
    batch = synth.generate_reads(contig_len, (30, 60), (False, True), seed)
    c = compact_struct(batch, narrow=narrow)
    assert c.qual_bits == 2 and sorted(c.qual_table[:4]) == [10, 20, 30, 40]  # four quality bins: 2 bits per base
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    want = gpu_ctx.somatic_standard([(0, 0, contig_len)], ss_params())
    gpu_ctx.upload_reads_compact(c)
    assert_same_batch(device_copy(gpu_ctx, batch.contig_names), batch)
    got = gpu_ctx.somatic_standard([(0, 0, contig_len)], ss_params())
    assert len(want["calls"]) > 0 and got["n_visited"] == want["n_visited"]

    def records(out):  # the allele strings sit in the arena in the order the CTAs claimed their slices: compare them by value
        a, c = out["arena"], out["calls"].copy()
        strings = [(a[r["ref_off"]:r["ref_off"] + r["ref_len"]], a[r["alt_off"]:r["alt_off"] + r["alt_len"]]) for r in c]
        c["ref_off"] = c["alt_off"] = 0
        return strings, c.tobytes()

    assert records(got) == records(want)
    # and the way back: the device batch exported in compact form is the host packer's output
    e = gpu_ctx.export_reads_compact_struct()
    try:
        assert (e.n_reads, e.n_samples, e.n_contig_runs, e.qual_bits, e.n_exc) == (c.n_reads, c.n_samples, c.n_contig_runs, 2, c.n_exc)
        assert list(e.qual_table[:4]) == list(c.qual_table[:4])
        n, n_bases = int(c.n_reads), int(batch.seq_off[-1])
        if narrow:  # 100-base reads, nearly all with the single operator 100M: 58 bytes per read on the wire
            assert e.uniform_read_len == c.uniform_read_len == 100 and not e.read_len and not e.start and not e.n_cigar_ops
            n_wire = int(c._keep["ops8"].sum())
            assert 0 < n_wire < len(batch.cigar) // 4
            columns = (("run_start_base", 1, 8), ("start_rel", n, 4), ("n_cigar_ops8", n, 1), ("cigar", n_wire, 4))
            assert c.wire_bytes(n_bases, n_wire) < 59 * n
        else:
            columns = (("start", n, 8), ("read_len", n, 2), ("n_cigar_ops", n, 2), ("cigar", len(batch.cigar), 4))
        for field, count, width in columns + (("bases2", (n_bases + 3) // 4, 1), ("qual", (n_bases + 3) // 4, 1), ("mapq", n, 1),
                                              ("flags", n, 1), ("sample_id", n, 1)):
            a = C.string_at(C.cast(getattr(e, field), C.c_void_p), count * width)
            b = C.string_at(C.cast(getattr(c, field), C.c_void_p), count * width)
            assert a == b, field
        gpu_ctx.upload_reads_compact(e)  # pinned buffers straight back in, as the e2e bench does
        assert_same_batch(device_copy(gpu_ctx, batch.contig_names), batch)
    finally:
        gpu_ctx.free_reads_compact_struct(e)


def _u8(p, n):
    return np.frombuffer(C.string_at(C.cast(p, C.c_void_p), n), np.uint8)


@pytest.mark.parametrize("name", ["tough", "tiny_sjc", "synth0", "synth1", "odd"])
def test_reference_relative_bases(gpu_ctx, name):
    """Reads as differences from the reference (mm_count / mm_words): the numpy packer's form and the device's own export of it
    both expand, on the device, to the very batch the plain upload gives."""
    if name.startswith("synth"):
        contig_len, seed = 30000, 4100
        ref = synth.make_reference(contig_len, seed)
        batch = synth.generate_reads(contig_len, (30, 60), (False, True), seed, profile=int(name[-1]))
    elif name == "odd":  # read lengths that are no multiple of 4: every read shares its first and last byte with a neighbour
        contig_len, seed = 4000, 4101
        ref = synth.make_reference(contig_len, seed)
        (w,) = ref.contigs.values()
        rng = np.random.default_rng(5)
        reads = []
        for i in range(300):
            ln, st = int(rng.integers(1, 40)), int(rng.integers(0, contig_len - 50))
            seq = bytearray(w.bases[st:st + ln].tobytes())
            for k in rng.integers(0, ln, size=int(rng.integers(0, 3))):
                seq[k] = b"ACGT"[(b"ACGT".index(seq[k]) + 1 + int(rng.integers(0, 3))) % 4]
            cigar = f"{ln}M" if i % 5 or ln < 3 else f"1M1I{ln - 2}M"  # every fifth read travels explicitly
            reads.append(make_read(seq.decode(), cigar, start=st, sample_id=i % 2, chr=w.name))
        batch = ReadBatch.from_reads(reads, contig_names=list(ref.contigs))
    else:
        batch, ref, _ = fixtures.load(name)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    c = compact_struct(batch, ref=ref)
    n, n_bases = batch.n_reads, int(batch.seq_off[-1])
    mm = np.ctypeslib.as_array(c.mm_count, (n,)).copy()
    assert (mm != 255).sum() > 0.5 * n or name in ("tough", "tiny_sjc")  # most synthetic reads travel as differences
    assert c.wire_bytes(n_bases, 0) < compact_struct(batch).wire_bytes(n_bases, 0)
    gpu_ctx.upload_reads_compact(c)
    assert_same_batch(device_copy(gpu_ctx, batch.contig_names), batch)
    # the device's own packing of the same batch: a read with more differences than a slot of the read pass holds stays
    # explicit there, every other read agrees with the numpy packer
    gpu_ctx.set_option("compact_ref_relative", 1)
    try:
        e = gpu_ctx.export_reads_compact_struct()
    finally:
        gpu_ctx.set_option("compact_ref_relative", 0)
    try:
        assert bool(e.mm_count)
        emm = _u8(e.mm_count, n)
        few = mm <= 8
        # (the device may keep a few more reads explicit: it tests the reference for A C G T in groups of 16 loci)
        same = emm[few] == mm[few]
        assert (same | (emm[few] == 255)).all() and (emm[~few] == 255).all()
        assert same.mean() > 0.5 if name in ("tough", "tiny_sjc") else same.all()  # (partial-FASTA fixtures: windows with holes)
        if few.all() and same.all():
            assert e.n_mm_words == c.n_mm_words and e.n_explicit_bases == c.n_explicit_bases
            assert C.string_at(C.cast(e.mm_words, C.c_void_p), 2 * int(e.n_mm_words)) == C.string_at(C.cast(c.mm_words, C.c_void_p), 2 * int(c.n_mm_words))
            nb = (int(e.n_explicit_bases) + 3) // 4
            assert C.string_at(C.cast(e.bases2, C.c_void_p), nb) == C.string_at(C.cast(c.bases2, C.c_void_p), nb)
        gpu_ctx.upload_reads_compact(e)
        assert_same_batch(device_copy(gpu_ctx, batch.contig_names), batch)
    finally:
        gpu_ctx.free_reads_compact_struct(e)


def test_reference_relative_bases_need_the_reference_first(gpu_ctx):
    contig_len, seed = 5000, 4102
    ref = synth.make_reference(contig_len, seed)
    batch = synth.generate_reads(contig_len, (10, 10), (False, True), seed)
    c = compact_struct(batch, ref=ref)
    fresh = type(gpu_ctx)(0)
    try:
        with pytest.raises(GuacError, match="upload the reference window"):
            fresh.upload_reads_compact(c)
        fresh.upload_reference(ref, batch.contig_names)
        c._keep["mm_words"][0] = (200 << 2) | 1  # a position beyond the read
        with pytest.raises(GuacError, match="difference word out of range"):
            fresh.upload_reads_compact(c)
        c2 = compact_struct(batch, ref=ref)
        c2.n_mm_words -= 1
        with pytest.raises(GuacError, match="do not match mm_count"):
            fresh.upload_reads_compact(c2)
    finally:
        fresh.close()


def test_odd_base_count_and_empty_batch(gpu_ctx):
    # 4 distinct qualities (2-bit codes), then 7 (4-bit codes), neither filling the last byte
    for quals in ([7, 9, 7, 9, 11], [7, 9, 12, 13, 14]):
        reads = [make_read("ACGTA", "5M", start=3, quality_scores=quals), make_read("TTGCATG", "3M1I3M", start=4, sample_id=1),
                 make_read("GGNAC", "5M", start=9, quality_scores=[2, 2, 2, 2, 2] if len(set(quals)) > 3 else None)]
        batch = ReadBatch.from_reads(reads)
        c = compact_struct(batch)
        assert c.qual_bits == (2 if len(np.unique(batch.qual)) <= 4 else 4)
        gpu_ctx.upload_reads_compact(c)
        assert_same_batch(device_copy(gpu_ctx, batch.contig_names), batch)
    empty = ReadBatch.from_reads([], contig_names=["chr1"], n_samples=1)
    gpu_ctx.upload_reads_compact(empty)
    assert device_copy(gpu_ctx, ["chr1"]).n_reads == 0


def test_ordering_contract(gpu_ctx):
    batch = ReadBatch.from_reads([make_read("ACGT", "4M", start=10), make_read("ACGT", "4M", start=20)])
    c = compact_struct(batch, narrow=False)
    c._keep["start_swapped"] = np.array([20, 10], np.int64)
    c.start = c._keep["start_swapped"].ctypes.data_as(C.POINTER(C.c_int64))
    with pytest.raises(GuacError, match="Regions must be sorted by start locus"):  # windowing/SlidingWindow.scala:36-37
        gpu_ctx.upload_reads_compact(c)
    c = compact_struct(batch, narrow=True)  # the same through the narrow start column
    c._keep["rel_swapped"] = np.array([10, 0], np.uint32)
    c.start_rel = c._keep["rel_swapped"].ctypes.data_as(C.POINTER(C.c_uint32))
    with pytest.raises(GuacError, match="Regions must be sorted by start locus"):
        gpu_ctx.upload_reads_compact(c)


@pytest.mark.parametrize("with_ref", [False, True])
def test_host_packer_feeds_the_device(gpu_ctx, with_ref):
    """SAM-record columns -> guac_pack_reads (pinned outputs of the context) -> guac_reads_upload_compact -> the device batch is
    byte-identical to the plain upload of the same (filtered) reads, and the caller gives the same records."""
    from guacamole_b200 import host_stage as hs
    from guacamole_b200 import partitioning
    from guacamole_b200.reads import FLAG_DUPLICATE
    batch, ref, _ = fixtures.load("tough")
    fl = batch.flags.copy()
    fl[::7] |= FLAG_DUPLICATE
    batch.flags = fl
    records, keepalive = hs.sam_records_of(batch)
    packed, kept = hs.pack_reads(records, ctx=gpu_ctx, non_duplicate=True, ref=ref if with_ref else None, contig_names=batch.contig_names)
    want_batch = partitioning.take_reads(batch, np.flatnonzero((fl & FLAG_DUPLICATE) == 0))
    assert kept == want_batch.n_reads and bool(packed.mm_count) == with_ref
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads_compact(packed)
    gpu_ctx._batch = want_batch
    got = gpu_ctx.export_reads(batch.contig_names)
    for f in ["contig_id", "start", "cigar_off", "cigar", "seq_off", "bases_ascii", "qual", "mapq", "flags", "sample_id"]:
        assert np.array_equal(getattr(got, f), getattr(want_batch, f)), f
    from guacamole_b200.commands.somatic_standard import covered_ranges
    loci = covered_ranges(want_batch)
    a = gpu_ctx.somatic_standard(loci)
    gpu_ctx.upload_reads(want_batch)
    b = gpu_ctx.somatic_standard(loci)
    from guacamole_b200.commands.somatic_standard import alleles_of
    assert len(a["calls"]) == len(b["calls"]) > 0 and alleles_of(a) == alleles_of(b)  # (arena offsets are handed out in any order)
    for f in ("start", "somatic_log_odds", "tumor_variant_evidence", "normal_reference_evidence"):
        assert np.array_equal(a["calls"][f], b["calls"][f]), f
    hs.free_packed(packed, ctx=gpu_ctx)
