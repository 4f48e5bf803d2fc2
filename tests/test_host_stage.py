"""The library's host stages (C++ behind the C ABI, no GPU needed): the two golden VCFs of the reference's end-to-end suites
through the C row writers, the genotype filter chain and the read packer against their Python twins, and the JVM-style number
formatting."""
import math
import random

import numpy as np
import pytest

from guacamole_b200 import host_stage as hs
from guacamole_b200.commands import somatic_joint as sj
from guacamole_b200.commands import somatic_standard as ss
from guacamole_b200.native import compact_struct
from guacamole_b200.reads import FLAG_DUPLICATE, FLAG_PAIRED, ReadBatch, filter_reads
from oracle import oracle
from tests.golden import fixtures


def test_tough_golden_vcf_through_the_c_rows():
    """SomaticStandardEndToEndSuite.scala:13-37: filter chain + rows in C++ == tough.golden.vcf."""
    batch, ref, extra = fixtures.load("tough")
    res = oracle.somatic_standard(batch, ref, ss.covered_ranges(batch))
    keep = hs.ss_genotype_filter(res["calls"])
    assert np.array_equal(res["calls"][keep], ss.somatic_genotype_filter(res["calls"]))
    rows = hs.ss_vcf_rows(res, batch.contig_names, keep, sample_name="tumor")
    assert rows == [str(x) for x in extra["golden_vcf"]]
    # every optional filter against the Python chain
    for kw in (dict(min_likelihood=30), dict(min_vaf=20), dict(min_tumor_alternate_read_depth=6), dict(maximum_median_mismatches=1),
               dict(min_average_mapping_quality=50), dict(max_tumor_read_depth=40), dict(min_normal_read_depth=30)):
        want = ss.somatic_genotype_filter(res["calls"], ss.SomaticGenotypeFilterArguments(**kw))
        assert np.array_equal(res["calls"][hs.ss_genotype_filter(res["calls"], hs.ss_filter_params(**kw))], want), kw


def test_tiny_sjc_golden_row_through_the_c_rows():
    """SomaticJointCallerEndToEndSuite.scala:20-38: the data row of tiny-sjc-output/all.vcf, rendered in C++."""
    batch, ref, extra = fixtures.load("tiny_sjc")
    kinds = [sj.NORMAL_DNA, sj.TUMOR_DNA, sj.TUMOR_DNA]
    loci = sj.loci_from_vcf_lines([str(x) for x in extra["force_call_vcf"]], batch.contig_names)
    p = sj.sj_params(kinds)
    alleles, samples, arena, _ = oracle.somatic_joint(batch, ref, loci, p, force_loci=loci, raw_records=True)
    rows = hs.sj_vcf_rows(alleles, samples, arena, p, batch.contig_names, ["normal", "primary", "recurrence"])
    assert rows == [str(x) for x in extra["golden_vcf"]]
    # ... and on a seeded batch with many records, against the Python renderer (both policies, pooled columns forced on)
    from guacamole_b200 import synth
    contig_len, depths, tumor, kinds, seed = 20000, (30, 30, 60, 60), (False, False, True, True), [0, 0, 1, 1], 4005
    sref, sbatch = synth.make_reference(contig_len, seed), synth.generate_reads(contig_len, depths, tumor, seed)
    p = sj.sj_params(kinds, include_filtered=1)
    alleles, samples, arena, _ = oracle.somatic_joint(sbatch, sref, [(0, 1, contig_len)], p, raw_records=True)
    evs = sj.decode(alleles, samples, arena, 4)
    names = ["n1", "n2", "t1", "t2"]
    assert len(evs) > 20
    for policy, pname in ((0, "presence"), (1, "trigger")):
        want = sj.vcf_rows(evs, sbatch.contig_names, names, kinds, somatic_genotype_policy=pname)
        assert hs.sj_vcf_rows(alleles, samples, arena, p, sbatch.contig_names, names, policy=policy) == want


def test_number_formatting_like_the_jvm():
    rnd = random.Random(7)
    xs = [0.0, 1.0, -1.0, 0.5, 0.05, 0.95, 1e-4, 9.99999995e-5, 99999999.5, 1e8, 123456789.0, -0.00012345678, 1e-21, 1e-20, 0.01, 0.00999,
          0.999, 0.9995, 1.005, 2.675, float("nan"), -41.26341, -1.7976931348623157e308]
    xs += [rnd.uniform(-300, 5) for _ in range(300)] + [10 ** rnd.uniform(-25, 12) * rnd.choice((-1, 1)) for _ in range(300)]
    for x in xs:
        assert hs.format_double(0, x) == sj.java_g8(x), x
        if not math.isnan(x):
            assert hs.format_double(1, x) == sj.format_vcf_double(x), x
    # ties go up (HALF_UP on the exact value), where C's printf would go to even
    assert hs.format_double(2, 0.125) == "0.13" and hs.format_double(2, 0.375) == "0.38" and hs.format_double(2, 2.5) == "2.50"
    assert hs.format_double(0, 1.00000005) == sj.java_g8(1.00000005)


def _columns(c):
    """The wire columns of a ReadBatchCompactC as numpy arrays (for comparison)."""
    import ctypes as C
    n = int(c.n_reads)
    def arr(p, dtype, count):
        if not p or count == 0:
            return None
        return np.frombuffer((C.c_char * (count * np.dtype(dtype).itemsize)).from_address(C.cast(p, C.c_void_p).value), dtype=dtype).copy()
    n_bases = n * int(c.uniform_read_len) if c.uniform_read_len > 0 else int(arr(c.read_len, np.uint16, n).sum(dtype=np.int64)) if n else 0
    ops8 = arr(c.n_cigar_ops8, np.uint8, n)
    n_cigar = int(ops8.sum(dtype=np.int64)) if ops8 is not None else (int(arr(c.n_cigar_ops, np.uint16, n).sum(dtype=np.int64)) if n else 0)
    per = 8 // int(c.qual_bits)
    return dict(n=n, runs=int(c.n_contig_runs), run_contig=arr(c.run_contig_id, np.int32, int(c.n_contig_runs)),
                run_first=arr(c.run_first_read, np.int64, int(c.n_contig_runs) + 1), start=arr(c.start, np.int64, n),
                run_base=arr(c.run_start_base, np.int64, int(c.n_contig_runs)), start_rel=arr(c.start_rel, np.uint32, n),
                read_len=arr(c.read_len, np.uint16, n), uniform=int(c.uniform_read_len), ops16=arr(c.n_cigar_ops, np.uint16, n), ops8=ops8,
                cigar=arr(c.cigar, np.uint32, n_cigar),
                bases2=arr(c.bases2, np.uint8, ((int(c.n_explicit_bases) if c.mm_count else n_bases) + 3) // 4),
                mm_count=arr(c.mm_count, np.uint8, n) if c.mm_count else 0, n_mm_words=int(c.n_mm_words), n_explicit=int(c.n_explicit_bases),
                mm_words=arr(c.mm_words, np.uint16, int(c.n_mm_words)) if c.mm_count and c.n_mm_words else 0, n_exc=int(c.n_exc),
                exc_pos=arr(c.exc_pos, np.uint64, int(c.n_exc)), exc_base=arr(c.exc_base, np.uint8, int(c.n_exc)), qual_bits=int(c.qual_bits),
                qual_table=bytes(c.qual_table), qual=arr(c.qual, np.uint8, (n_bases + per - 1) // per), mapq=arr(c.mapq, np.uint8, n),
                flags=arr(c.flags, np.uint8, n), sample=arr(c.sample_id, np.uint8, n))


def _assert_same_columns(a, b):
    for k in a:
        x, y = a[k], b[k]
        if isinstance(x, np.ndarray) or isinstance(y, np.ndarray):
            assert x is not None and y is not None and np.array_equal(x, y), k
        else:
            assert x == y, k


@pytest.mark.parametrize("name", ["tiny_sjc", "tough", "gatk_mini", "synthetic"])
def test_pack_reads_equals_the_numpy_packer(name):
    """Read.apply + ReadSets.filterRDD + packer in C++: the kept reads and every wire column equal the Python twins'."""
    if name == "synthetic":
        from guacamole_b200 import synth
        batch = synth.generate_reads(30000, (30, 60), (False, True), 11, profile=1)
    else:
        batch, _, _ = fixtures.load(name)
    rng = np.random.default_rng(5)
    # a few records unmapped, a few not paired / duplicates: the filters have something to drop
    fl = batch.flags.copy()
    fl[rng.random(batch.n_reads) < 0.05] |= FLAG_DUPLICATE
    fl[rng.random(batch.n_reads) < 0.05] &= ~np.uint8(FLAG_PAIRED)
    batch.flags = fl
    unmapped = rng.random(batch.n_reads) < 0.03
    records, keepalive = hs.sam_records_of(batch, extra_flags=np.where(unmapped, 0x4, 0))
    from guacamole_b200 import partitioning
    for kw in (dict(), dict(non_duplicate=True, is_paired=True), dict(min_alignment_quality=30, passed_vendor_quality_checks=True)):
        got, n_kept = hs.pack_reads(records, **kw)
        keep = ~unmapped
        if kw.get("non_duplicate"):
            keep &= (fl & FLAG_DUPLICATE) == 0
        if kw.get("is_paired"):
            keep &= (fl & FLAG_PAIRED) != 0
        if kw.get("min_alignment_quality"):
            keep &= batch.mapq >= kw["min_alignment_quality"]
        if kw.get("passed_vendor_quality_checks"):
            keep &= (fl & 8) == 0
        want_batch = partitioning.take_reads(batch, np.flatnonzero(keep))
        assert n_kept == want_batch.n_reads > 0
        want = compact_struct(want_batch)
        _assert_same_columns(_columns(got), _columns(want))
        hs.free_packed(got)
    # overlapsLoci: only reads over the given ranges
    c0 = int(batch.contig_id[0])
    lo = int(batch.start[batch.n_reads // 2])
    got, n_kept = hs.pack_reads(records, overlaps_loci=[(c0, lo, lo + 50)])
    ends = batch.ends()
    assert n_kept == int(np.count_nonzero(~unmapped & (batch.contig_id == c0) & (batch.start < lo + 50) & (ends > lo)))
    hs.free_packed(got)


@pytest.mark.parametrize("name", ["tiny_sjc", "tough", "synthetic", "synthetic_plain"])
def test_pack_reads_with_the_reference_equals_the_numpy_packer(name):
    """guac_pack_reads_ref: reads that can travel as their differences from the reference do, the others' bases are packed
    on their own — every wire column equals the numpy twin's (native.compact_struct(batch, ref=...))."""
    if name.startswith("synthetic"):
        from guacamole_b200 import synth
        ref = synth.make_reference(30000, 12)
        batch = synth.generate_reads(30000, (30, 60), (False, True), 12, profile=0 if name.endswith("plain") else 1)
    else:
        batch, ref, _ = fixtures.load(name)
    fl = batch.flags.copy()
    fl[::9] |= FLAG_DUPLICATE
    batch.flags = fl
    records, keepalive = hs.sam_records_of(batch)
    from guacamole_b200 import partitioning
    for threads in (1, 5):
        got, n_kept = hs.pack_reads(records, non_duplicate=True, ref=ref, contig_names=batch.contig_names, n_threads=threads)
        want_batch = partitioning.take_reads(batch, np.flatnonzero((fl & FLAG_DUPLICATE) == 0))
        assert n_kept == want_batch.n_reads > 0
        want = compact_struct(want_batch, ref=ref)
        cols = _columns(got)
        assert isinstance(cols["mm_count"], np.ndarray) and (cols["mm_count"] != 255).sum() > 0.5 * n_kept
        _assert_same_columns(cols, _columns(want))
        hs.free_packed(got)


def test_reference_relative_columns_expand_to_the_reads():
    """The wire semantics without a GPU: the bases rebuilt from guac_pack_reads_ref's columns — the reference window patched
    with the difference words, or the explicit stream — are the reads' bases (what refrel_expand_kernel does on the device)."""
    from guacamole_b200 import synth
    ref = synth.make_reference(20000, 31)
    batch = synth.generate_reads(20000, (20, 40), (False, True), 31, profile=1)
    (window,) = ref.contigs.values()
    records, keepalive = hs.sam_records_of(batch)
    got, n_kept = hs.pack_reads(records, ref=ref, contig_names=batch.contig_names, n_threads=3)
    assert n_kept == batch.n_reads
    c = _columns(got)
    mm, words = c["mm_count"], c["mm_words"]
    explicit = np.repeat(c["bases2"], 4)
    explicit = (explicit >> np.tile(np.array([0, 2, 4, 6], np.uint8), len(c["bases2"]))) & 3  # one 2-bit code per entry
    code = np.full(256, 255, np.uint8)
    for i, ch in enumerate(b"ACGT"):
        code[ch] = i
    w = e = 0
    for r in range(batch.n_reads):
        s0, s1 = int(batch.seq_off[r]), int(batch.seq_off[r + 1])
        want = code[batch.bases_ascii[s0:s1]]
        if mm[r] == 255:
            have = explicit[e:e + s1 - s0]
            e += s1 - s0
            n_mask = want > 3  # (not A C G T: travels in the exception list, any code here)
            assert np.array_equal(have[~n_mask], want[~n_mask]), r
        else:
            rel = int(batch.start[r]) - window.offset
            have = code[window.bases[rel:rel + s1 - s0]].copy()
            for word in words[w:w + int(mm[r])]:
                assert word & 3, r
                have[int(word) >> 2] ^= int(word) & 3
            w += int(mm[r])
            assert np.array_equal(have, want), r
    assert w == c["n_mm_words"] and e == c["n_explicit"]
    hs.free_packed(got)


def test_pack_reads_ref_edge_cases():
    """Which reads may travel as differences: hand-made reads at the window's edges, on a reference base that is not A C G T,
    with an N, with = / X operators, with two operators, on a contig without a window — C++ packer and numpy twin agree, and the
    expected reads are the explicit ones."""
    from guacamole_b200.io import ContigWindow, Reference
    from guacamole_b200.reads import ReadBatch, make_read
    window = np.frombuffer(b"ACGTACGTNNACGTACGTACGTTTGGCCAAACGT", np.uint8).copy()  # offset 100: loci 100..133, N at 108, 109
    ref = Reference({"c1": ContigWindow("c1", 100, window, 1000)})
    def seq(lo, hi):
        return window[lo - 100:hi - 100].tobytes().decode()
    reads = [
        make_read(seq(100, 108), "8M", start=100, chr="c1"),                      # 0: first base of the window: relative
        make_read("T" + seq(101, 108), "8M", start=100, chr="c1"),                # 1: one difference at position 0
        make_read(seq(104, 112).replace("N", "A"), "8M", start=104, chr="c1"),    # 2: over the N bases of the reference: explicit
        make_read(seq(110, 118), "8=", start=110, chr="c1"),                      # 3: '=' operator: relative
        make_read(seq(110, 114) + "A" + seq(115, 118), "4=1X3=", start=110, chr="c1"),  # 4: three operators: explicit
        make_read(seq(112, 116) + "N" + seq(117, 120), "8M", start=112, chr="c1"),      # 5: an N in the read: explicit
        make_read(seq(126, 134), "8M", start=126, chr="c1"),                      # 6: ends at the window's last base: relative
        make_read(seq(127, 134) + "A", "8M", start=127, chr="c1"),                # 7: one base past the window: explicit
        make_read("ACGTACGT", "8M", start=5, chr="c2"),                           # 8: a contig without a window: explicit
        make_read("ACGTAC", "2S4M", start=120, chr="c1"),                         # 9: soft clip: explicit
        make_read("ACGTACGT"[::-1], "8X", start=114, chr="c1"),                   # 10: 'X' operator over all bases: relative
    ]
    batch = ReadBatch.from_reads(reads, contig_names=["c1", "c2"])
    order = sorted(range(len(reads)), key=lambda i: (0 if reads[i].contig == "c1" else 1, reads[i].start, i))
    records, keepalive = hs.sam_records_of(batch)
    got, n_kept = hs.pack_reads(records, ref=ref, contig_names=batch.contig_names, n_threads=2)
    want = compact_struct(batch, ref=ref)
    cols = _columns(got)
    _assert_same_columns(cols, _columns(want))
    explicit_reads = {2, 4, 5, 7, 8, 9}
    assert [int(m) == 255 for m in cols["mm_count"]] == [i in explicit_reads for i in order]
    by_read = dict(zip(order, cols["mm_count"]))
    assert by_read[0] == 0 and by_read[1] == 1 and by_read[3] == 0 and by_read[6] == 0 and by_read[10] > 0
    hs.free_packed(got)


def test_pack_reads_sorts_a_shuffled_partition():
    batch, _, _ = fixtures.load("tough")
    from guacamole_b200 import partitioning
    perm = np.random.default_rng(3).permutation(batch.n_reads)
    shuffled = partitioning.take_reads(batch, perm)
    records, keepalive = hs.sam_records_of(shuffled)
    got, n_kept = hs.pack_reads(records)
    order = np.lexsort((np.arange(batch.n_reads), shuffled.start, shuffled.contig_id))  # stable by (contig, start)
    want_batch = partitioning.take_reads(shuffled, order)  # (kept alive: the struct points into its arrays)
    want = compact_struct(want_batch)
    _assert_same_columns(_columns(got), _columns(want))
    hs.free_packed(got)
