"""somatic-joint at BASELINE configs[3]'s depths (2 normal + 2 tumor, 30/30/60/60) over a megabase: the default device path
(read pass, classifying and pruning locus pass, fast site kernel) against the generic per-locus kernels and against the
oracle, record by record.  Small fixtures cannot show a one-in-a-thousand disagreement of the call bound; this can."""
import numpy as np
import pytest

from guacamole_b200 import synth
from guacamole_b200.commands import somatic_joint as sj
from oracle import oracle
from tests.test_gpu_somatic_joint import assert_evidences_equal

pytestmark = pytest.mark.gpu


def keys(evs):
    return [(e["start"], e["ref"], e["alt"], e["is_germline_call"], e["is_somatic_call"]) for e in evs]


def test_megabase_four_samples(gpu_ctx):
    contig_len, depths, tumor, kinds, seed = 1_000_000, (30, 30, 60, 60), (False, False, True, True), [0, 0, 1, 1], 4301
    ref = synth.make_reference(contig_len, seed)
    batch = synth.generate_reads(contig_len, depths, tumor, seed)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    loci = [(0, 1, contig_len)]
    p = sj.sj_params(kinds)
    results = {}
    try:
        for name, impl, prune in (("default", 2, 1), ("no-prune", 2, 0), ("generic", 0, 1)):
            gpu_ctx.set_option("sj_impl", impl)
            gpu_ctx.set_option("sj_prune", prune)
            results[name], nv = gpu_ctx.somatic_joint(loci, p)
            if name == "default":
                stats = {s: gpu_ctx.last_stat(s) for s in ("heavy_loci", "sj_simple", "sj_bail", "sj_sites", "sj_emitted_sites")}
    finally:
        gpu_ctx.set_option("sj_impl", 2)
        gpu_ctx.set_option("sj_prune", 1)
    print("stats", stats)
    for name in ("no-prune", "default"):
        a, b = set(keys(results[name])), set(keys(results["generic"]))
        assert a == b, (name, sorted(a - b)[:10], sorted(b - a)[:10])
    want, nv_o = oracle.somatic_joint(batch, ref, loci, p)
    assert nv == nv_o and len(want) > 900
    assert keys(results["generic"]) == keys(want)
    assert_evidences_equal(results["default"], want)
    # the locus pass keeps few sites beyond the ones that are called, and the generic kernels see few loci
    assert stats["sj_simple"] < 3 * stats["sj_emitted_sites"] and stats["heavy_loci"] < stats["sj_emitted_sites"], stats


def test_device_paths_agree_on_ten_megabases(gpu_ctx):
    """Default path against the generic kernels alone on a device-generated 10 Mbp batch (no oracle: the generic kernels are
    pinned to it above and in test_gpu_somatic_joint.py)."""
    n, depths, tumor, kinds = 10_000_000, (30, 30, 60, 60), (False, False, True, True), [0, 0, 1, 1]
    gpu_ctx.synth_generate(n, depths, tumor, 4000)
    p = sj.sj_params(kinds)
    out = {}
    try:
        for name, impl, prune in (("default", 2, 1), ("no-prune", 2, 0), ("generic", 0, 1)):
            gpu_ctx.set_option("sj_impl", impl)
            gpu_ctx.set_option("sj_prune", prune)
            a, smp, arena, nv = gpu_ctx.somatic_joint_records([(0, 0, n)], p)
            out[name] = (a, smp, arena)
    finally:
        gpu_ctx.set_option("sj_impl", 2)
        gpu_ctx.set_option("sj_prune", 1)
    g = out["generic"]
    for name in ("no-prune", "default"):
        d = out[name]
        sd, sg = set(d[0]["start"].tolist()), set(g[0]["start"].tolist())
        assert sd == sg, (name, sorted(sd - sg)[:10], sorted(sg - sd)[:10])
        assert len(d[0]) == len(g[0]) and d[2] == g[2]
        for f in ("start", "is_germline_call", "is_somatic_call", "failing_filters", "germline_genotype"):
            assert np.array_equal(d[0][f], g[0][f]), (name, f)
        for f in ("total_depth", "sb_variant_forward", "sb_total_forward", "triggered"):
            assert np.array_equal(d[1][f], g[1][f]), (name, f)
        np.testing.assert_allclose(d[1]["log_likelihoods"], g[1]["log_likelihoods"], rtol=1e-9, atol=1e-12)
