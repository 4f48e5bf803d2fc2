"""GPU parity for the germline toy caller (callPileupVariant over all possible genotypes), through the C ABI."""
import numpy as np
import pytest

from guacamole_b200 import synth
from guacamole_b200.commands.somatic_standard import alleles_of, covered_ranges
from guacamole_b200.native import gs_params
from oracle import oracle
from tests.golden import fixtures

pytestmark = pytest.mark.gpu

EV_INT = ["read_depth", "allele_read_depth", "forward_depth", "allele_forward_depth"]
EV_F = ["probability", "mean_mapping_quality", "median_mapping_quality", "mean_base_quality", "median_base_quality",
        "median_mismatches_per_read"]


def assert_germline_equal(got, want):
    assert got["n_visited"] == want["n_visited"]
    assert got["active_loci"] == want["active_loci"]
    assert np.array_equal(got["calls"]["start"], want["calls"]["start"])
    assert np.array_equal(got["calls"]["contig_id"], want["calls"]["contig_id"])
    assert np.array_equal(got["calls"]["n_genotypes"], want["calls"]["n_genotypes"])
    assert alleles_of(got) == alleles_of(want)
    # normalised natural-log genotype probabilities, FP64 within 1e-9 relative (tiny absolute slack near 0)
    np.testing.assert_allclose(got["calls"]["genotype_log_probability"], want["calls"]["genotype_log_probability"], rtol=1e-9, atol=1e-12)
    for f in EV_INT:
        assert np.array_equal(got["calls"]["evidence"][f], want["calls"]["evidence"][f]), f
    for f in EV_F:
        np.testing.assert_allclose(got["calls"]["evidence"][f], want["calls"]["evidence"][f], rtol=1e-9, atol=1e-300, equal_nan=True)


@pytest.mark.parametrize("respect_active,min_phred", [(0, 0), (1, 0), (0, 20)])
def test_synthetic_30x_single_sample(gpu_ctx, respect_active, min_phred):
    """BASELINE configs[1] shape: one 30x sample, planted het / hom SNVs, 1 % indel reads."""
    contig_len, seed = 40000, 2000
    ref = synth.make_reference(contig_len, seed)
    batch = synth.generate_reads(contig_len, (30,), (False,), seed)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    p = gs_params(respect_active_regions=respect_active, min_phred_scaled_likelihood=min_phred)
    got = gpu_ctx.germline_standard([(0, 0, contig_len)], p)
    want = oracle.germline_standard(batch, ref, [(0, 0, contig_len)], p)
    if respect_active:
        assert len(want["active_loci"]) > 20  # most candidate regions of this noisy batch would go to the assembler
    else:
        assert len(want["calls"]) > 20
    assert_germline_equal(got, want)
    # the same through the tile kernel's whole-tile fallback
    gpu_ctx.set_option("ss_force_fallback_mod", 2)
    try:
        again = gpu_ctx.germline_standard([(0, 0, contig_len)], p)
    finally:
        gpu_ctx.set_option("ss_force_fallback_mod", 0)
    assert_germline_equal(again, want)


def test_tumor_sample_of_the_tough_fixture(gpu_ctx):
    """Real reads (indels, soft clips): the toy caller over the tumor sample of the somatic-standard fixture."""
    batch, ref, _ = fixtures.load("tough")
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    ranges = covered_ranges(batch)
    for active in (0, 1):
        p = gs_params(sample_id=1, respect_active_regions=active)
        got = gpu_ctx.germline_standard(ranges, p)
        want = oracle.germline_standard(batch, ref, ranges, p)
        assert len(want["calls"]) > 0
        assert_germline_equal(got, want)
