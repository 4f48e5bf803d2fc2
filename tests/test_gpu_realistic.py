"""The fast paths on reads that look like real ones (synth profile 1: depth that varies like a Poisson pileup's, quality-2
tails on ~15 % of reads, 10 % multi-operator reads, 1 % of reads with 10..20 mismatches) and on the reference's real-read
fixtures: parity with the oracle, and the counters that say how much work left the fast path."""
import numpy as np
import pytest

from guacamole_b200 import synth
from guacamole_b200.commands import somatic_joint as sj
from guacamole_b200.commands.somatic_standard import covered_ranges
from oracle import oracle
from tests.golden import fixtures
from tests.test_gpu_somatic_joint import assert_evidences_equal

pytestmark = pytest.mark.gpu


def test_somatic_standard_on_realistic_reads(gpu_ctx):
    contig_len, seed = 400_000, 3100
    ref = synth.make_reference(contig_len, seed)
    batch = synth.generate_reads(contig_len, (30, 60), (False, True), seed, profile=1)
    assert 0.1 < np.mean(batch.qual.reshape(-1, 100).min(axis=1) < 4) < 0.2  # reads with a base below quality 4
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    loci = [(0, 0, contig_len)]
    got = gpu_ctx.somatic_standard(loci, want_summary=True)
    fallback, heavy = gpu_ctx.last_stat("fallback_tiles"), gpu_ctx.last_stat("heavy_loci")
    want = oracle.somatic_standard(batch, ref, loci, want_summary=True, n_threads=8)
    assert got["n_visited"] == want["n_visited"] and len(want["calls"]) > 20
    assert np.array_equal(got["summary"], want["summary"])
    assert np.array_equal(got["calls"]["start"], want["calls"]["start"])
    np.testing.assert_allclose(got["calls"]["somatic_log_odds"], want["calls"]["somatic_log_odds"], rtol=1e-9)
    print(f"realistic reads: fallback tiles {fallback} of {(contig_len + 511) // 512}, heavy loci {heavy} of {got['n_visited']}")
    assert fallback <= 0.01 * ((contig_len + 511) // 512) and heavy <= 0.01 * got["n_visited"], (fallback, heavy)


def test_two_contigs_of_realistic_reads(gpu_ctx):
    """A batch of two contigs: the read pass runs once per contig (each with its own reference) and both share the overflow
    arena of the reads with more events than their slot holds (the 1 % of reads with 10..20 mismatches)."""
    from guacamole_b200.io import Reference
    from guacamole_b200.reads import ReadBatch
    lens, seeds, names = (60_000, 90_000), (3201, 3202), ("ctgA", "ctgB")
    refs = [synth.make_reference(n, s) for n, s in zip(lens, seeds)]
    parts = [synth.generate_reads(n, (30, 60), (False, True), s, profile=1, contig_name=c) for n, s, c in zip(lens, seeds, names)]
    windows = {}
    for c, r in zip(names, refs):
        (w,) = r.contigs.values()
        w.name = c
        windows[c] = w
    ref = Reference(windows)
    a, b = parts
    batch = ReadBatch(
        n_samples=2, contig_names=list(names), contig_id=np.concatenate([a.contig_id, b.contig_id + 1]), start=np.concatenate([a.start, b.start]),
        cigar_off=np.concatenate([a.cigar_off, b.cigar_off[1:] + a.cigar_off[-1]]).astype(np.uint32), cigar=np.concatenate([a.cigar, b.cigar]),
        seq_off=np.concatenate([a.seq_off, b.seq_off[1:] + a.seq_off[-1]]).astype(np.uint64), bases_ascii=np.concatenate([a.bases_ascii, b.bases_ascii]),
        qual=np.concatenate([a.qual, b.qual]), mapq=np.concatenate([a.mapq, b.mapq]), flags=np.concatenate([a.flags, b.flags]),
        sample_id=np.concatenate([a.sample_id, b.sample_id]))
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    loci = [(0, 0, lens[0]), (1, 0, lens[1])]
    got = gpu_ctx.somatic_standard(loci, want_summary=True)
    want = oracle.somatic_standard(batch, ref, loci, want_summary=True, n_threads=8)
    assert got["n_visited"] == want["n_visited"] and len(want["calls"]) > 5
    assert gpu_ctx.last_stat("fallback_tiles") == 0
    assert np.array_equal(got["summary"], want["summary"])
    assert np.array_equal(got["calls"]["contig_id"], want["calls"]["contig_id"]) and np.array_equal(got["calls"]["start"], want["calls"]["start"])
    np.testing.assert_allclose(got["calls"]["somatic_log_odds"], want["calls"]["somatic_log_odds"], rtol=1e-9)
    # and the second contig alone, its reads in the middle of the batch's arena
    part = gpu_ctx.somatic_standard([(1, 20_000, 50_000)])
    inside = (want["calls"]["contig_id"] == 1) & (want["calls"]["start"] >= 20_000) & (want["calls"]["start"] < 50_000)
    assert np.array_equal(part["calls"]["start"], want["calls"]["start"][inside])


def test_somatic_joint_on_realistic_reads(gpu_ctx):
    contig_len, depths, tumor, kinds, seed = 150_000, (30, 30, 60, 60), (False, False, True, True), [0, 0, 1, 1], 4500
    ref = synth.make_reference(contig_len, seed)
    batch = synth.generate_reads(contig_len, depths, tumor, seed, profile=1)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    loci = [(0, 1, contig_len)]
    p = sj.sj_params(kinds)
    got, nv = gpu_ctx.somatic_joint(loci, p)
    stats = {s: gpu_ctx.last_stat(s) for s in ("fallback_tiles", "heavy_loci", "sj_simple", "sj_bail", "sj_emitted_sites")}
    want, nv_o = oracle.somatic_joint(batch, ref, loci, p)
    assert nv == nv_o and len(want) > 50
    assert_evidences_equal(got, want)
    print("realistic reads, joint:", stats)


@pytest.mark.parametrize("name", ["tough", "real_syn1fp", "real_complexvar", "real_simplefp"])
def test_counters_on_real_read_fixtures(gpu_ctx, name):
    """Where do the reference's own real reads leave the fast path?  (Partial FASTA windows: a tile that reaches outside the
    reference window is a fallback tile by construction, so the fallback count mostly says how the fixture was cut.)"""
    batch, ref, _ = fixtures.load(name)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    loci = covered_ranges(batch)
    got = gpu_ctx.somatic_standard(loci)
    fallback, heavy = gpu_ctx.last_stat("fallback_tiles"), gpu_ctx.last_stat("heavy_loci")
    want = oracle.somatic_standard(batch, ref, loci)
    assert np.array_equal(got["calls"]["start"], want["calls"]["start"]) and got["n_visited"] == want["n_visited"]
    n_tiles = sum((e - s + 511) // 512 for _, s, e in loci)
    print(f"{name}: fallback tiles {fallback} of {n_tiles}, heavy loci {heavy} of {got['n_visited']} visited")
    assert heavy <= got["n_visited"]
