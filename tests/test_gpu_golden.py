"""GPU parity on the reference's own fixtures, through the C ABI: golden VCF rows, and every
PileupElement of the PileupSuite / PileupFlatMapUtilsSuite inputs bit-exact against the oracle."""
import numpy as np
import pytest

from guacamole_b200.commands import somatic_standard as ss
from guacamole_b200.io import make_reference
from guacamole_b200.native import GuacError
from guacamole_b200.reads import ReadBatch, make_read
from oracle import oracle
from tests.golden import fixtures
from tests.test_oracle_golden import FLATMAP_REF, flatmap_reads

pytestmark = pytest.mark.gpu


def check_pileups(ctx, batch, ref, ranges):
    ctx.upload_reference(ref, batch.contig_names)
    ctx.upload_reads(batch)
    for skip in (True, False):
        gc, gl, gcnt = ctx.pileup_counts(ranges, skip_empty=skip)
        oc, ol, ocnt = oracle.pileup_counts(batch, ref, ranges, skip_empty=skip)
        assert np.array_equal(gc, oc) and np.array_equal(gl, ol)
        assert np.array_equal(gcnt, ocnt)
    gc, gl, goff, gel = ctx.pileup_elements(ranges)
    oc, ol, ooff, oel, _ = oracle.pileup_elements(batch, ref, ranges)
    assert np.array_equal(gl, ol) and np.array_equal(goff, ooff)
    assert np.array_equal(gel, oel)
    return gl


@pytest.mark.parametrize("impl", ["tile", "gather", "stream"])
def test_tough_golden_vcf_on_gpu(gpu_ctx, impl):
    batch, ref, extra = fixtures.load("tough")
    gpu_ctx.set_option("ss_impl", {"tile": 0, "gather": 1, "stream": 2}[impl])
    try:
        gpu_ctx.upload_reference(ref, batch.contig_names)
        gpu_ctx.upload_reads(batch)
        ranges = ss.covered_ranges(batch)
        out = gpu_ctx.somatic_standard(ranges, want_summary=True)
    finally:
        gpu_ctx.set_option("ss_impl", 2)
    rows = ss.vcf_rows(out, batch.contig_names, ss.somatic_genotype_filter(out["calls"]))
    assert rows == [str(x) for x in extra["golden_vcf"]]
    want = oracle.somatic_standard(batch, ref, ranges, want_summary=True)
    assert np.array_equal(out["summary"], want["summary"])
    np.testing.assert_allclose(out["calls"]["somatic_log_odds"], want["calls"]["somatic_log_odds"], rtol=1e-9)
    for ev in ("tumor_variant_evidence", "normal_reference_evidence"):
        for f in ("mean_mapping_quality", "median_mapping_quality", "mean_base_quality", "median_base_quality",
                  "median_mismatches_per_read"):
            np.testing.assert_allclose(out["calls"][ev][f], want["calls"][ev][f], rtol=1e-9, equal_nan=True)


def test_tough_pileups(gpu_ctx):
    batch, ref, _ = fixtures.load("tough")
    check_pileups(gpu_ctx, batch, ref, ss.covered_ranges(batch))


def test_same_start_reads(gpu_ctx):
    batch, _, _ = fixtures.load("same_start_reads")
    ref = make_reference((batch.contig_names[0], 0, "A" * 100))
    check_pileups(gpu_ctx, batch, ref, [(0, 0, 80)])


def test_different_start_reads(gpu_ctx):
    """29M10D31M, 5M5I39M10D16M, 10M10I10D40M, 5M4=1X5=, 4=1D4=4S, 4=1N4=4H (PileupSuite.scala:249-381)."""
    batch, _, _ = fixtures.load("different_start_reads")
    ref = make_reference((batch.contig_names[0], 0, "A" * 34 + "G" * 10 + "A" * 5 + "G" * 15 + "A" * 15 + "ACGT" * 10 + "A" * 200))
    check_pileups(gpu_ctx, batch, ref, [(0, 0, 120)])


def test_rna_reads_with_introns(gpu_ctx):
    batch, _, _ = fixtures.load("testrna")
    lo = int(batch.start.min())
    ref = make_reference((batch.contig_names[0], lo, "ACGT" * ((int(batch.ends().max()) - lo) // 4 + 4)), contig_length=249250621)
    loci = check_pileups(gpu_ctx, batch, ref, [(0, 229580590, 229580720)])
    assert 229580594 in loci


def test_flatmap_three_samples_with_non_acgt_bases(gpu_ctx):
    """PileupFlatMapUtilsSuite.scala:179-256: reads XYX / XZX exercise the non-ACGT side table, the
    reference has N runs, chr0 has no reference at all and is never visited."""
    ref = make_reference(*FLATMAP_REF)
    batch = flatmap_reads(3)
    loci = check_pileups(gpu_ctx, batch, ref, [(1, 1, 500), (2, 10, 20)])
    assert list(loci) == [1, 2, 3, 4, 5, 6, 7, 8, 99, 100, 101, 102, 103, 104, 105, 106, 107]
    # with skipEmpty, a contig that has neither reads nor a reference is simply never visited
    _, gl, _ = gpu_ctx.pileup_counts([(0, 0, 1000), (1, 1, 500), (2, 10, 20)], skip_empty=True)
    _, ol, _ = oracle.pileup_counts(batch, ref, [(0, 0, 1000), (1, 1, 500), (2, 10, 20)], skip_empty=True)
    assert np.array_equal(gl, ol)


def test_insertion_at_contig_start_and_errors(gpu_ctx):
    ref = make_reference(("chr1", 0, "CTCGATCGACG"))
    batch = ReadBatch.from_reads([make_read("AAAAAACGT", "5I4M", 0, "chr1")], ["chr1"], 1)
    check_pileups(gpu_ctx, batch, ref, [(0, 0, 4)])  # PileupSuite.scala:163-167
    # a locus whose reference base is missing must fail like MapBackedReferenceSequence.apply
    batch = ReadBatch.from_reads([make_read("CTCGATCGACGAAAA", "15M", 0, "chr1")], ["chr1"], 1)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    with pytest.raises(GuacError) as e:
        gpu_ctx.pileup_counts([(0, 0, 15)])
    assert e.value.status == -1 and "missing from reference" in e.value.message
    # unsorted reads are rejected with the reference's wording (SlidingWindow.scala:37)
    bad = ReadBatch.from_reads([make_read("CTCG", "4M", 5, "chr1"), make_read("CTCG", "4M", 1, "chr1")], ["chr1"], 1)
    bad.start = bad.start[::-1].copy()
    with pytest.raises(GuacError) as e:
        gpu_ctx.upload_reads(bad)
    assert "sorted by start" in e.value.message
