"""GPU parity for somatic-joint, through the C ABI: the golden VCF row of the reference's end-to-end
suite, and record-by-record agreement with the oracle on seeded 3- and 4-sample batches."""
import math

import numpy as np
import pytest

from guacamole_b200 import synth
from guacamole_b200.commands import somatic_joint as sj
from guacamole_b200.io import make_reference
from guacamole_b200.reads import ReadBatch, make_read
from oracle import oracle
from tests.golden import fixtures

pytestmark = pytest.mark.gpu

INT_KEYS = ["contig_id", "start", "ref", "alt", "is_germline_call", "is_somatic_call", "germline_genotype", "tumor_rna_expressed",
            "insufficient_normal", "failing_filters", "forced", "site_index"]


def assert_evidences_equal(got, want):
    assert len(got) == len(want)
    for g, w in zip(got, want):
        for k in INT_KEYS:
            assert g[k] == w[k], (k, g["start"], g[k], w[k])
        assert len(g["samples"]) == len(w["samples"])
        for gs, ws in zip(g["samples"], w["samples"]):
            for k in ("kind", "total_depth", "allelic_depths", "strand_bias", "sb_filtered", "triggered"):
                assert gs[k] == ws[k], (k, g["start"], gs[k], ws[k])
            # FP64 log10 likelihoods / posteriors within 1e-9 relative (north star)
            for k in ("log_likelihoods", "posteriors"):
                a, b = np.array(gs[k]), np.array(ws[k])
                huge = np.abs(b) > 1e300  # Double.MaxValue priors
                np.testing.assert_allclose(a[~huge], b[~huge], rtol=1e-9, atol=1e-12)
                assert np.array_equal(np.sign(a[huge]), np.sign(b[huge]))
            assert math.isclose(gs["alt_vaf"], ws["alt_vaf"], rel_tol=1e-12, abs_tol=0)
            assert math.isclose(gs["sb_phred"], ws["sb_phred"], rel_tol=1e-9, abs_tol=1e-12)


def test_tiny_sjc_golden_vcf_on_gpu(gpu_ctx, sj_impl):
    batch, ref, extra = fixtures.load("tiny_sjc")
    kinds = [sj.NORMAL_DNA, sj.TUMOR_DNA, sj.TUMOR_DNA]
    loci = sj.loci_from_vcf_lines([str(x) for x in extra["force_call_vcf"]], batch.contig_names)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    evs, visited = gpu_ctx.somatic_joint(loci, sj.sj_params(kinds), force_loci=loci)
    assert visited == 1
    rows = sj.vcf_rows(evs, batch.contig_names, ["normal", "primary", "recurrence"], kinds)
    assert rows == [str(x) for x in extra["golden_vcf"]]
    want, _ = oracle.somatic_joint(batch, ref, loci, sj.sj_params(kinds), force_loci=loci)
    assert_evidences_equal(evs, want)


def test_tiny_sjc_all_covered_loci(gpu_ctx, sj_impl):
    """Beyond the single golden site: every covered locus of the fixture, no forcing, filtered calls included."""
    from guacamole_b200.commands.somatic_standard import covered_ranges
    batch, ref, _ = fixtures.load("tiny_sjc")
    kinds = [sj.NORMAL_DNA, sj.TUMOR_DNA, sj.TUMOR_DNA]
    loci = [(c, s + 1, e - 120) for c, s, e in covered_ranges(batch) if e - 120 > s + 1]
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    p = sj.sj_params(kinds, include_filtered=1)
    got, nv_g = gpu_ctx.somatic_joint(loci, p)
    want, nv_o = oracle.somatic_joint(batch, ref, loci, p)
    assert nv_g == nv_o
    assert_evidences_equal(got, want)


@pytest.mark.parametrize("depths,tumor,kinds,seed", [
    ((30, 30, 60, 60), (False, False, True, True), [0, 0, 1, 1], 4005),   # BASELINE configs[3] shape: 2 normal + 2 tumor
    ((30, 60, 40), (False, True, True), [0, 1, 2], 4105),                   # with a tumor RNA sample
])
def test_synthetic_multi_sample(gpu_ctx, sj_impl, depths, tumor, kinds, seed):
    contig_len = 20000
    ref = synth.make_reference(contig_len, seed)
    batch = synth.generate_reads(contig_len, depths, tumor, seed)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    loci = [(0, 1, contig_len)]
    force = [(0, 5000, 5003), (0, 12000, 12001)]
    for p, f in ((sj.sj_params(kinds), ()), (sj.sj_params(kinds, include_filtered=1), force), (sj.sj_params(kinds, only_somatic=1), ())):
        got, nv_g = gpu_ctx.somatic_joint(loci, p, force_loci=f)
        want, nv_o = oracle.somatic_joint(batch, ref, loci, p, force_loci=f)
        assert nv_g == nv_o
        assert len(want) > 0 or p.only_somatic
        assert_evidences_equal(got, want)
    # partition invariance (PileupFlatMapUtilsSuite.scala:214-241): many small ranges == one range
    whole, _ = gpu_ctx.somatic_joint(loci, sj.sj_params(kinds))
    edges = list(range(1, contig_len, 613)) + [contig_len]
    parts, _ = gpu_ctx.somatic_joint([(0, a, b) for a, b in zip(edges[:-1], edges[1:])], sj.sj_params(kinds))
    assert_evidences_equal(parts, whole)  # integer fields exact; likelihood sums are order-dependent in the last bits


def test_indel_alleles_and_non_standard_bases(gpu_ctx, sj_impl):
    """ReadSubsequenceSuite / PileupStatsSuite reads: shared insertion alleles, an SNV, an N base."""
    ref = make_reference(("chr1", 0, "NTCGATCGACGTTACGGATCCA"))
    reads = []
    for s in (0, 1):
        reads += [make_read("TCGATCGACGTTACG", "15M", 1, sample_id=s) for _ in range(6)]
        reads += [make_read("TCGACCCTCGACGTTACG", "4M3I11M", 1, sample_id=s) for _ in range(3)]
        reads += [make_read("TCGAGCGACGTTACG", "15M", 1, sample_id=s, is_positive_strand=False) for _ in range(3)]
        reads += [make_read("TCGATCGCGTTACG", "7M1D7M", 1, sample_id=s) for _ in range(3)]
        reads += [make_read("TNGATCGACGTTACG", "15M", 1, sample_id=s)]
    batch = ReadBatch.from_reads(reads, ["chr1"], 2)
    kinds = [sj.NORMAL_DNA, sj.TUMOR_DNA]
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    p = sj.sj_params(kinds, include_filtered=1)
    force = [(0, 1, 14)]  # force-calling emits every site's best allele, called or not
    got, _ = gpu_ctx.somatic_joint([(0, 1, 14)], p, force_loci=force)
    want, _ = oracle.somatic_joint(batch, ref, [(0, 1, 14)], p, force_loci=force)
    assert {(e["ref"], e["alt"]) for e in want} >= {("A", "ACCC"), ("T", "G"), ("GA", "G")}
    assert_evidences_equal(got, want)


@pytest.fixture(params=[(2, 1), (2, 0), (1, 1), (0, 1)], ids=["stream", "stream-no-prune", "single-kernel-locus-pass", "generic-kernels"])
def sj_impl(request, gpu_ctx):
    """Every parity case runs through the default form (read pass + classifying, pruning locus pass + fast site kernel), the
    same without the in-tile prune (every simple candidate reaches the site kernel), the single-kernel locus pass, and the
    generic per-locus kernels alone."""
    gpu_ctx.set_option("sj_impl", request.param[0])
    gpu_ctx.set_option("sj_prune", request.param[1])
    yield request.param
    gpu_ctx.set_option("sj_impl", 2)
    gpu_ctx.set_option("sj_prune", 1)


def test_repeat_runs_are_bit_identical(gpu_ctx):
    """The likelihood sums are taken in a fixed order (per-lane partial sums combined by fixed trees; the generic kernel adds
    its terms in read order), so two runs over the same batch return the same bytes — rendered VCF rows cannot flip."""
    contig_len, depths, tumor, kinds, seed = 600_000, (30, 60, 30), (False, True, True), [0, 1, 1], 4205
    ref = synth.make_reference(contig_len, seed)
    batch = synth.generate_reads(contig_len, depths, tumor, seed)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    loci = [(0, 1, contig_len)]
    p = sj.sj_params(kinds, include_filtered=1)
    for impl in (2, 0):
        gpu_ctx.set_option("sj_impl", impl)
        try:
            a1, s1, arena1, nv1 = gpu_ctx.somatic_joint_records(loci, p)
            a2, s2, arena2, nv2 = gpu_ctx.somatic_joint_records(loci, p)
        finally:
            gpu_ctx.set_option("sj_impl", 2)
        assert nv1 == nv2 and len(a1) > 100
        assert a1.tobytes() == a2.tobytes() and s1.tobytes() == s2.tobytes() and arena1 == arena2
    # the records read in place in the library's pinned buffers (what a JNI consumer does) are the same bytes
    a1, s1, arena1, nv1 = gpu_ctx.somatic_joint_records(loci, p)
    a3, s3, arena3, nv3 = gpu_ctx.somatic_joint_records(loci, p, copy=False)
    assert nv3 == nv1 and a3.tobytes() == a1.tobytes() and s3.tobytes() == s1.tobytes() and arena3 == arena1
    del a3, s3
    gpu_ctx.release_outputs()
    # and the fast site kernel took nearly every candidate of this SNV-only workload
    gpu_ctx.somatic_joint_records(loci, p)
    simple, bail, heavy = gpu_ctx.last_stat("sj_simple"), gpu_ctx.last_stat("sj_bail"), gpu_ctx.last_stat("heavy_loci")
    assert simple > 0 and bail < 0.05 * simple and heavy < 0.15 * simple, (simple, bail, heavy)
