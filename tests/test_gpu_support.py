"""GPU parity of guac_variant_support and guac_vaf_histogram (commands/VariantSupport.scala:100-112,
commands/VAFHistogram.scala:37-58,194-245) through the C ABI: VariantSupportSuite's vectors on the device, and
record-by-record agreement with the oracle on real reads (one and two samples) and on a seeded synthetic batch."""
import numpy as np
import pytest

from guacamole_b200 import synth
from guacamole_b200.commands.somatic_standard import covered_ranges
from oracle import oracle
from tests.golden import fixtures
from tests.test_oracle_support import suite_cases

pytestmark = pytest.mark.gpu


def test_variant_support_suite_vectors_on_gpu(gpu_ctx):
    for batch, ref, cases in suite_cases():
        c20 = batch.contig_names.index("20")
        gpu_ctx.upload_reference(ref, batch.contig_names)
        gpu_ctx.upload_reads(batch)
        for locus, want in cases:
            got = gpu_ctx.variant_support([(c20, locus, locus + 1)])
            assert sorted((r, a, n) for _, _, _, r, a, n in got) == sorted(want), locus


@pytest.mark.parametrize("name", ["gatk_mini", "tough"])
def test_real_reads_match_the_oracle(gpu_ctx, name):
    batch, ref, _ = fixtures.load(name)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    loci = [(c, s, min(e, s + 3000)) for c, s, e in covered_ranges(batch)][:6]
    assert gpu_ctx.variant_support(loci) == oracle.variant_support(batch, ref, loci)
    for bins, depth, vaf in ((20, 0, 0), (7, 3, 10), (100, 0, 0)):
        h, t, recs = gpu_ctx.vaf_histogram(loci, bins, depth, vaf)
        ho, to, ro = oracle.vaf_histogram(batch, ref, loci, bins, depth, vaf)
        assert np.array_equal(h, ho) and np.array_equal(t, to) and int(t.sum()) > 0
        assert [(int(r["sample_id"]), int(r["contig_id"]), int(r["locus"]), np.float32(r["variant_allele_frequency"])) for r in recs] == ro


def test_synthetic_two_samples(gpu_ctx):
    ref, batch = synth.make_tumor_normal(contig_len=6000, seed=77)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    loci = [(0, 100, 2100), (0, 3000, 3500)]
    assert gpu_ctx.variant_support(loci) == oracle.variant_support(batch, ref, loci)
    h, t, recs = gpu_ctx.vaf_histogram(loci, 20, 0, 0)
    ho, to, ro = oracle.vaf_histogram(batch, ref, loci, 20, 0, 0)
    assert np.array_equal(h, ho) and np.array_equal(t, to) and len(recs) == len(ro)
    with pytest.raises(Exception):
        gpu_ctx.vaf_histogram(loci, 0, 0, 0)  # "Bins should be between 1 and 100"
