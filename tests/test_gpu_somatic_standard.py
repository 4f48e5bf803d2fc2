"""GPU parity: CUDA path (through the C ABI) vs the CPU oracle on the same seeded inputs."""
import numpy as np
import pytest

from guacamole_b200 import synth
from guacamole_b200.native import Context, GuacError, ss_params
from oracle import oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(params=["tile", "gather", "tile+fallback", "stream", "stream+fallback", "stream+bitmaporder"])
def impl(request, gpu_ctx):
    """Every parity case runs through the tile kernel, the generic gather kernel, the tile kernel with every third
    tile forced through its fallback path, the stream implementation (read pass as a kernel of its own), and the stream
    implementation with the calls ranked through the bitmap over the loci (the path taken beyond 32768 calls)."""
    gpu_ctx.set_option("ss_impl", {"gather": 1, "stream": 2, "stream+fallback": 2, "stream+bitmaporder": 2}.get(request.param, 0))
    gpu_ctx.set_option("ss_force_fallback_mod", 3 if request.param.endswith("+fallback") else 0)
    gpu_ctx.set_option("ss_rank_max", 0 if request.param.endswith("+bitmaporder") else -1)
    yield request.param
    gpu_ctx.set_option("ss_impl", 2)  # the default
    gpu_ctx.set_option("ss_force_fallback_mod", 0)
    gpu_ctx.set_option("ss_rank_max", -1)

EV_INT = ["read_depth", "allele_read_depth", "forward_depth", "allele_forward_depth"]
EV_F = ["probability", "mean_mapping_quality", "median_mapping_quality", "mean_base_quality", "median_base_quality",
        "median_mismatches_per_read"]


def alleles(out):
    a = out["arena"]
    return [(a[c["ref_off"]:c["ref_off"] + c["ref_len"]], a[c["alt_off"]:c["alt_off"] + c["alt_len"]]) for c in out["calls"]]


def assert_ss_equal(got, want):
    assert got["n_visited"] == want["n_visited"]
    if "summary" in want:
        assert np.array_equal(got["summary"], want["summary"])  # bit-exact depths / counts / flags
    assert np.array_equal(got["calls"]["contig_id"], want["calls"]["contig_id"])
    assert np.array_equal(got["calls"]["start"], want["calls"]["start"])
    assert alleles(got) == alleles(want)
    for f in ["somatic_log_odds", "tumor_ref_log_prob", "tumor_alt_log_prob", "normal_ref_log_prob", "normal_alt_log_prob"]:
        # FP64 log-likelihoods within 1e-9 relative (north star); tiny absolute slack for values ~0
        np.testing.assert_allclose(got["calls"][f], want["calls"][f], rtol=1e-9, atol=1e-12)
    for ev in ["tumor_variant_evidence", "normal_reference_evidence"]:
        for f in EV_INT:
            assert np.array_equal(got["calls"][ev][f], want["calls"][ev][f]), (ev, f)
        for f in EV_F:
            np.testing.assert_allclose(got["calls"][ev][f], want["calls"][ev][f], rtol=1e-9, atol=0, equal_nan=True)


@pytest.mark.parametrize("seed,contig_len", [(12, 20000), (11, 60000)])
def test_synthetic_tumor_normal(gpu_ctx, impl, seed, contig_len):
    ref, batch = synth.make_tumor_normal(contig_len=contig_len, seed=seed)
    ranges = [(0, 0, contig_len)]
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    got = gpu_ctx.somatic_standard(ranges, want_summary=True)
    want = oracle.somatic_standard(batch, ref, ranges, want_summary=True)
    assert len(want["calls"]) > 0
    assert_ss_equal(got, want)
    if impl.endswith("+fallback"):
        assert gpu_ctx.last_fallback_tiles > 0
    elif impl in ("tile", "stream"):
        assert gpu_ctx.last_fallback_tiles == 0


def test_partition_invariance(gpu_ctx, impl):
    """UniformPartitioner(1) vs many partitions must give identical output
    (reference: PileupFlatMapUtilsSuite.scala:214-241)."""
    ref, batch = synth.make_tumor_normal(contig_len=30000, seed=5)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    one = gpu_ctx.somatic_standard([(0, 0, 30000)], want_summary=True)
    edges = list(range(0, 30000, 777)) + [30000]
    many = gpu_ctx.somatic_standard([(0, a, b) for a, b in zip(edges[:-1], edges[1:])], want_summary=True)
    assert_ss_equal(many, one)


def test_thresholds_and_params(gpu_ctx, impl):
    ref, batch = synth.make_tumor_normal(contig_len=20000, seed=13)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    p = ss_params(tumor_odds_threshold=2, normal_odds_threshold=1, max_normal_alternate_read_depth=2,
                  min_tumor_variant_allele_frequency_percent=7)
    got = gpu_ctx.somatic_standard([(0, 100, 19000)], p, want_summary=True)
    want = oracle.somatic_standard(batch, ref, [(0, 100, 19000)], p, want_summary=True)
    assert_ss_equal(got, want)


def test_pileup_counts_and_elements(gpu_ctx):
    ref, batch = synth.make_tumor_normal(contig_len=8000, seed=21)
    ranges = [(0, 0, 3000), (0, 3500, 8000)]
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    for skip in (True, False):
        gc, gl, gcnt = gpu_ctx.pileup_counts(ranges, skip_empty=skip)
        oc, ol, ocnt = oracle.pileup_counts(batch, ref, ranges, skip_empty=skip)
        assert np.array_equal(gl, ol) and np.array_equal(gc, oc)
        assert np.array_equal(gcnt, ocnt)
    gc, gl, goff, gel = gpu_ctx.pileup_elements(ranges)
    oc, ol, ooff, oel, _ = oracle.pileup_elements(batch, ref, ranges)
    assert np.array_equal(gl, ol) and np.array_equal(goff, ooff)
    assert np.array_equal(gel, oel)  # every PileupElement cursor field, bit-exact
