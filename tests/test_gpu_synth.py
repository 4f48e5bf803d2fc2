"""The CUDA generator (guac_synth_generate) must be bit-identical to its numpy specification,
so that the oracle can be run on any window of the full-size benchmark data."""
import numpy as np
import pytest

from guacamole_b200 import synth
from oracle import oracle

pytestmark = pytest.mark.gpu

FIELDS = ["start", "cigar_off", "cigar", "seq_off", "bases_ascii", "qual", "mapq", "flags", "sample_id", "contig_id"]


@pytest.mark.parametrize("profile", [0, 1])
@pytest.mark.parametrize("contig_len,depths,tumor,seed", [(30000, (30, 60), (False, True), 3000),
                                                          (20000, (30, 30, 60, 60), (False, False, True, True), 4001)])
def test_device_generator_matches_numpy(gpu_ctx, contig_len, depths, tumor, seed, profile):
    gpu_ctx.synth_generate(contig_len, depths, tumor, seed, profile=profile)
    got = gpu_ctx.export_reads(["synth1"])
    want = synth.generate_reads(contig_len, depths, tumor, seed, profile=profile)
    for f in FIELDS:
        assert np.array_equal(getattr(got, f), getattr(want, f)), f
    assert np.array_equal(gpu_ctx.export_reference(0, 0, contig_len), synth.reference_bases(seed, 0, contig_len))


def test_window_of_large_batch_matches_oracle(gpu_ctx):
    """Full-size-style check: generate a multi-Mbp batch on the device, run the fused caller over the
    whole contig, and compare a window of it with the oracle run on the same window of reads."""
    contig_len, seed = 3_000_000, 3000
    depths, tumor = (30, 60), (False, True)
    gpu_ctx.synth_generate(contig_len, depths, tumor, seed)
    lo, hi = 1_500_000, 1_540_000
    got = gpu_ctx.somatic_standard([(0, lo, hi)], want_summary=True)
    j_lo, j_hi = synth.read_window_for_loci(contig_len, depths, lo, hi)
    batch = oracle.synth_generate(contig_len, depths, tumor, seed, j_lo, j_hi)
    from guacamole_b200.io import ContigWindow, Reference
    ref = Reference({"synth1": ContigWindow("synth1", lo - 300, oracle.synth_reference(seed, lo - 300, hi + 300), contig_len)})
    want = oracle.somatic_standard(batch, ref, [(0, lo, hi)], want_summary=True)
    assert got["n_visited"] == want["n_visited"] == hi - lo
    assert np.array_equal(got["summary"], want["summary"])
    assert np.array_equal(got["calls"]["start"], want["calls"]["start"])
    np.testing.assert_allclose(got["calls"]["somatic_log_odds"], want["calls"]["somatic_log_odds"], rtol=1e-9)
    # whole-contig property: every locus below the margin is covered, depth sums match the read count
    full = gpu_ctx.somatic_standard([(0, 0, contig_len)], want_summary=True)
    assert full["n_visited"] >= contig_len - 2 * synth.MARGIN
    sel = (full["calls"]["start"] >= lo) & (full["calls"]["start"] < hi)
    assert np.array_equal(full["calls"]["start"][sel], want["calls"]["start"])
