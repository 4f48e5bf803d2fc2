"""The multi-GPU contract on the GPU path itself: loci-range shards with halo reads
(readsets/rdd/PartitionedRegions.scala:186-208 over loci/partitioning/UniformPartitioner.scala:48-86), each uploaded and run
through the device callers, gathered on the host — equal to the unsharded device result and to the oracle
(PileupFlatMapUtilsSuite.scala:214-241: the number of partitions must not change the output)."""
import threading

import numpy as np
import pytest

from guacamole_b200 import partitioning, synth
from guacamole_b200.commands import somatic_joint as sj
from guacamole_b200.commands import somatic_standard as ss
from guacamole_b200.native import Context, ga_params
from oracle import oracle

pytestmark = pytest.mark.gpu


def test_somatic_standard_shards(gpu_ctx):
    ref, batch = synth.make_tumor_normal(contig_len=40000, seed=15)
    loci = [(0, 0, 40000)]
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    whole = gpu_ctx.somatic_standard(loci)
    parts = []
    for rank, (ranges, mine) in partitioning.shard(batch, loci, 4).items():
        gpu_ctx.upload_reads(mine)
        parts.append(gpu_ctx.somatic_standard(ranges))
    merged = partitioning.gather_calls(parts)
    want = oracle.somatic_standard(batch, ref, loci)
    for other in (whole, want):
        assert merged["n_visited"] == other["n_visited"] and len(merged["calls"]) == len(other["calls"]) > 0
        assert np.array_equal(merged["calls"]["start"], other["calls"]["start"])
        assert ss.alleles_of(merged) == ss.alleles_of(other)
        np.testing.assert_allclose(merged["calls"]["somatic_log_odds"], other["calls"]["somatic_log_odds"], rtol=1e-9)


@pytest.mark.parametrize("profile", [0, 1])
def test_loci_subsets_of_a_resident_batch(gpu_ctx, profile):
    """The whole batch stays on the device and the callers are asked for some of its loci only (what `bench.py --scaling
    strong` does): the read pass then runs over the reads near those loci, and the records are the whole call's records at
    those loci."""
    contig_len, seed = 300000, 5151
    ref = synth.make_reference(contig_len, seed)
    batch = synth.generate_reads(contig_len, (30, 60), (False, True), seed, profile=profile)
    assert batch.n_reads > 100 * 1024  # (many entries in the host's sampled index of the batch)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    whole = gpu_ctx.somatic_standard([(0, 0, contig_len)])
    w_start, w_alleles = whole["calls"]["start"], ss.alleles_of(whole)
    assert len(w_start) > 20
    p = sj.sj_params([0, 1])
    whole_j, _ = gpu_ctx.somatic_joint([(0, 1, contig_len)], p)
    for ranges in ([(0, 0, 1)], [(0, 123457, 123458)], [(0, 1000, 1999), (0, 2050, 2051), (0, 250000, 299999)], [(0, 299000, contig_len)],
                   [(0, 150000, 150700)]):
        part = gpu_ctx.somatic_standard(ranges)
        inside = np.zeros(len(w_start), bool)
        for _, s, e in ranges:
            inside |= (w_start >= s) & (w_start < e)
        assert part["n_visited"] <= sum(e - s for _, s, e in ranges)
        assert np.array_equal(part["calls"]["start"], w_start[inside])
        assert ss.alleles_of(part) == [a for a, k in zip(w_alleles, inside) if k]
        assert np.array_equal(part["calls"]["somatic_log_odds"], whole["calls"]["somatic_log_odds"][inside])
        part_j, _ = gpu_ctx.somatic_joint([(c, max(1, s), e) for c, s, e in ranges], p)
        want_j = [e_ for e_ in whole_j if any(s <= e_["start"] - 1 < e for _, s, e in ranges)]  # (pileup locus = allele start - 1)
        key = lambda e_: (e_["start"], e_["ref"], e_["alt"], e_["is_germline_call"], e_["is_somatic_call"],
                          tuple(tuple(s_["log_likelihoods"]) for s_ in e_["samples"]))
        assert [key(e_) for e_ in part_j] == [key(e_) for e_ in want_j]


def test_somatic_joint_shards(gpu_ctx):
    contig_len, depths, tumor, kinds, seed = 40000, (30, 30, 60, 60), (False, False, True, True), [0, 0, 1, 1], 4401
    ref = synth.make_reference(contig_len, seed)
    batch = synth.generate_reads(contig_len, depths, tumor, seed)
    loci = [(0, 1, contig_len)]
    p = sj.sj_params(kinds)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    whole, nv = gpu_ctx.somatic_joint(loci, p)
    merged, nv_parts = [], 0
    for rank, (ranges, mine) in partitioning.shard(batch, loci, 4, half_window=1).items():  # pileups sit one locus before the alleles
        gpu_ctx.upload_reads(mine)
        evs, v = gpu_ctx.somatic_joint(ranges, p)
        merged += evs
        nv_parts += v
    want, nv_o = oracle.somatic_joint(batch, ref, loci, p)
    key = lambda e: (e["start"], e["ref"], e["alt"], e["is_germline_call"], e["is_somatic_call"], tuple(s["total_depth"] for s in e["samples"]))
    assert nv_parts == nv == nv_o and len(want) > 10
    assert [key(e) for e in merged] == [key(e) for e in whole] == [key(e) for e in want]
    for a, b in zip(merged, want):
        for sa, sb in zip(a["samples"], b["samples"]):
            np.testing.assert_allclose(sa["log_likelihoods"], sb["log_likelihoods"], rtol=1e-9, atol=1e-12)


def test_germline_assembly_shards(gpu_ctx):
    """The assembly closure carries state along a partition (lastCalledLocus, the window jump), so the reference's own result
    depends on the partitioning: the sharded device run is compared with the oracle run over the same shards, halo =
    assembly_window_range."""
    contig_len, seed = 30000, 2300
    ref = synth.make_reference(contig_len, seed)
    batch = synth.generate_reads(contig_len, (30,), (False,), seed)
    p = ga_params(kmer_size=47, assembly_window_range=120, min_occurrence=5, min_mean_kmer_quality=20, min_area_vaf_percent=10)
    loci = [(0, 200, contig_len - 200)]
    gpu_ctx.upload_reference(ref, batch.contig_names)
    from tests.test_gpu_assembly import assert_assembly_equal
    n_calls = 0
    for rank, (ranges, mine) in partitioning.shard(batch, loci, 3, half_window=p.assembly_window_range).items():
        gpu_ctx.upload_reads(mine)
        got = gpu_ctx.germline_assembly(ranges, p)
        want = oracle.germline_assembly(mine, ref, ranges, p)
        assert_assembly_equal(got, want)
        n_calls += len(got["calls"])
    assert n_calls > 0


def test_two_contexts_on_two_devices_from_two_threads():
    """The Spark local[N] shape of INTEGRATION.md: one process, one context per GPU, each driven by its own thread."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    ref, batch = synth.make_tumor_normal(contig_len=60000, seed=25)
    want = oracle.somatic_standard(batch, ref, [(0, 0, 60000)])
    contig_len, depths, tumor, kinds, seed = 20000, (30, 30, 60, 60), (False, False, True, True), [0, 0, 1, 1], 4402
    jref, jbatch = synth.make_reference(contig_len, seed), synth.generate_reads(contig_len, depths, tumor, seed)
    jwant, _ = oracle.somatic_joint(jbatch, jref, [(0, 1, contig_len)], sj.sj_params(kinds))
    results, errors = {}, []

    def work(device):
        try:
            ctx = Context(device)
            for _ in range(3):
                ctx.upload_reference(ref, batch.contig_names)
                ctx.upload_reads(batch)
                out = ctx.somatic_standard([(0, 0, 60000)])
                ctx.upload_reference(jref, jbatch.contig_names)
                ctx.upload_reads(jbatch)
                evs, _ = ctx.somatic_joint([(0, 1, contig_len)], sj.sj_params(kinds))
            results[device] = (out, evs)
        except Exception as e:  # noqa: BLE001
            errors.append((device, repr(e)))
    threads = [threading.Thread(target=work, args=(d,)) for d in (0, 1)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for d in (0, 1):
        out, evs = results[d]
        assert np.array_equal(out["calls"]["start"], want["calls"]["start"]) and out["n_visited"] == want["n_visited"]
        assert [(e["start"], e["alt"]) for e in evs] == [(e["start"], e["alt"]) for e in jwant]
