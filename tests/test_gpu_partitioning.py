"""GPU parity of the coverage-aware partitioner (guac_take_loci, guac_capped_partitioning) through the C ABI:
TakeLociIteratorSuite's vectors on the device, and the whole partitioner against the oracle on seeded reads."""
import numpy as np
import pytest

from guacamole_b200 import partitioning, synth
from oracle import oracle
from tests.test_oracle_partitioning import CONTIGS, TAKE_LOCI_CASES, as_strings, coverages_of

pytestmark = pytest.mark.gpu


def sets_of(intervals):
    out = {}
    for c, a, b, p in intervals:
        out.setdefault(p, []).append((c, a, b))
    return [out[p] for p in sorted(out)]


@pytest.mark.parametrize("name,trim,inputs,expected", TAKE_LOCI_CASES, ids=[c[0] for c in TAKE_LOCI_CASES])
def test_take_loci_suite_on_gpu(gpu_ctx, name, trim, inputs, expected):
    names = {v: k for k, v in CONTIGS.items()}
    intervals, n_parts = gpu_ctx.take_loci(coverages_of(inputs), 15, trim)
    assert as_strings(sets_of(intervals), names) == expected and n_parts == len(expected)


@pytest.mark.parametrize("half_window,max_reads,trim", [(0, 400, False), (0, 150, True), (7, 300, False), (0, 70, False), (3, 95, True)])
def test_capped_partitioning_matches_oracle(gpu_ctx, half_window, max_reads, trim):
    """Two samples at 30x + 60x over 6 kbp, profile 1 (uneven depth): caps low enough that loci are dropped (depth ~90 > 70)."""
    contig_len, seed = 6000, 3300
    ref = synth.make_reference(contig_len, seed)
    batch = synth.generate_reads(contig_len, (30, 60), (False, True), seed, profile=1)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    loci = [(0, 50, 2000), (0, 2100, 2105), (0, 2500, 5900)]
    got, n_parts = gpu_ctx.capped_partitioning(loci, half_window, max_reads, trim)
    want = oracle.capped_partitioning(batch, loci, half_window, max_reads, trim)
    assert got == want and n_parts == len({p for *_, p in want}) > 3


def test_partitions_feed_the_shards(gpu_ctx):
    """The loci sets, dealt to ranks in contiguous blocks, cover the same visited loci as the uncut ranges and give the same
    calls (every locus is in exactly one loci set: coverage-balanced instead of length-balanced shards)."""
    ref, batch = synth.make_tumor_normal(contig_len=30000, seed=35)
    loci = [(0, 0, 30000)]
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    whole = gpu_ctx.somatic_standard(loci)
    intervals, n_parts = gpu_ctx.capped_partitioning(loci, 0, 500)  # ~55 loci sets of at most 500 reads
    by_rank = partitioning.assign_partitions(intervals, n_parts, 4)
    parts = []
    for rank in range(4):
        ranges = by_rank[rank]
        gpu_ctx.upload_reads(partitioning.take_reads(batch, partitioning.reads_for_ranges(batch, ranges)))
        parts.append(gpu_ctx.somatic_standard(ranges))
    merged = partitioning.gather_calls(parts)
    assert merged["n_visited"] == whole["n_visited"] and np.array_equal(merged["calls"]["start"], whole["calls"]["start"])
    reads_per_rank = [len(partitioning.reads_for_ranges(batch, by_rank[r])) for r in range(4)]
    assert n_parts > 40 and max(reads_per_rank) < 1.15 * min(reads_per_rank), reads_per_rank
