"""CPU-side tests: the C-ABI library loads and exports every declared symbol, host helpers,
the generator twins, partitioning, and the world_size-2 (gloo) shard / gather logic."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from guacamole_b200 import partitioning, synth
from guacamole_b200.commands import somatic_standard as ss
from guacamole_b200.io import make_reference
from guacamole_b200.reads import ReadBatch, make_read, parse_cigar, cigar_to_string
from oracle import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = []
    inc = os.path.join(ROOT, "include")
    for f in os.listdir(inc):
        if f.endswith(".h"):
            text = re.sub(r"/\*.*?\*/", "", open(os.path.join(inc, f)).read(), flags=re.S)
            names += re.findall(r"\b(guac_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


def test_library_exports_every_declared_symbol(built):
    from guacamole_b200.native import load_library
    lib = load_library()
    syms = declared_symbols()
    assert len(syms) >= 15
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/*.h but not exported"


def test_no_cpu_fallback_without_device(built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from guacamole_b200.native import Context, GuacError
    with pytest.raises(GuacError) as e:
        Context(0)
    assert e.value.status == -3  # GUAC_E_NO_DEVICE: the product fails loudly, it never computes on the CPU


def test_pack_bases_round_trip(built):
    from guacamole_b200.native import pack_bases
    rng = np.random.default_rng(1)
    ascii_bases = np.frombuffer(b"ACGTNRYX", np.uint8)[rng.choice(8, 1003, p=[.24, .24, .24, .24, .01, .01, .01, .01])]
    p = pack_bases(ascii_bases)
    codes = ((p["bases2"][:, None] >> np.array([0, 2, 4, 6], np.uint8)[None, :]) & 3).reshape(-1)[:1003]
    out = np.frombuffer(b"ACGT", np.uint8)[codes].copy()
    mask = ((p["nmask"][:, None] >> np.arange(8, dtype=np.uint8)[None, :]) & 1).reshape(-1)[:1003].astype(bool)
    assert np.array_equal(mask, ~np.isin(ascii_bases, (65, 67, 71, 84)))
    assert p["n_exc"] == mask.sum()
    out[p["exc_pos"][:p["n_exc"]].astype(np.int64)] = p["exc_base"][:p["n_exc"]]
    assert np.array_equal(out, ascii_bases)


def test_cigar_codec():
    assert cigar_to_string(parse_cigar("5M4=1X5=3I2D7N4S2H1P")) == "5M4=1X5=3I2D7N4S2H1P"
    r = make_read("A" * 20, "5M3I2D7M1P5S", start=10)
    assert r.end == 10 + 5 + 2 + 7 + 1  # getPaddedReferenceLength counts M D N = X P (MappedRead.scala:75)


def test_generator_twins_agree():
    for (L, depths, tumor, seed, lo, hi) in [(20000, (30, 60), (False, True), 12, 0, 3000),
                                             (50_000_000, (30, 60), (False, True), 3000, 1_000_000, 1_002_000),
                                             (125_000_000, (30, 30, 60, 60), (False, False, True, True), 4001, 200_000_000, 200_001_000)]:
        for profile in (0, 1):
            a = synth.generate_reads(L, depths, tumor, seed, lo, hi, profile=profile)
            b = oracle.synth_generate(L, depths, tumor, seed, lo, hi, profile=profile)
            for f in ["start", "cigar_off", "cigar", "seq_off", "bases_ascii", "qual", "mapq", "flags", "sample_id"]:
                assert np.array_equal(getattr(a, f), getattr(b, f)), (f, profile)
            assert np.all(np.diff(a.start) >= 0)
            if profile == 1 and hi - lo >= 2000:  # the realistic variant's shares: multi-operator reads, quality-2 tails
                n_ops = np.diff(a.cigar_off.astype(np.int64))
                assert 0.06 < np.mean(n_ops > 1) < 0.14
                tails = (a.qual.reshape(-1, 100)[:, -1] == 2)
                assert 0.10 < tails.mean() < 0.20
    assert np.array_equal(synth.reference_bases(5, 100, 1000), oracle.synth_reference(5, 100, 1000))


def test_uniform_partitioner_strings():  # UniformPartitionerSuite.scala:12-42
    names = ["chrM"]
    assert partitioning.partitioning_to_string(partitioning.uniform_partition([(0, 0, 16571)], 4), names) == \
        "chrM:0-4143=0,chrM:4143-8286=1,chrM:8286-12428=2,chrM:12428-16571=3"
    assert partitioning.partitioning_to_string(partitioning.uniform_partition([(0, 0, 10)], 3), names) == "chrM:0-3=0,chrM:3-7=1,chrM:7-10=2"
    assert partitioning.partitioning_to_string(partitioning.uniform_partition([(0, 0, 3)], 4), names) == "chrM:0-1=0,chrM:1-2=1,chrM:2-3=2"
    assert partitioning.uniform_partition([(0, 10, 10)], 4) == []
    p = partitioning.uniform_partition([(0, 1000, 1100)], 100)
    assert p == [(0, 1000 + i, 1001 + i, i) for i in range(100)]
    two = partitioning.uniform_partition([(1, 0, 15), (1, 100, 121), (2, 100, 200)], 2)
    assert sum(e - s for _, s, e, q in two if q == 0) == sum(e - s for _, s, e, q in two if q == 1) == 68


def test_filter_and_vcf_rows_shape():
    ref, batch = synth.make_tumor_normal(contig_len=20000, seed=12)
    out = oracle.somatic_standard(batch, ref, [(0, 0, 20000)])
    kept = ss.somatic_genotype_filter(out["calls"])
    rows = ss.vcf_rows(out, batch.contig_names, kept)
    assert rows[0].startswith("#CHROM") and len(rows) == len(kept) + 1
    assert all(r.split("\t")[8] == "GT:AD:DP:GQ" for r in rows[1:])
    strict = ss.somatic_genotype_filter(out["calls"], ss.SomaticGenotypeFilterArguments(min_tumor_alternate_read_depth=10 ** 6))
    assert len(strict) == 0


def test_sharding_is_result_invariant():
    """Loci-range shards with halo reads give exactly the unsharded result (the multi-GPU contract)."""
    ref, batch = synth.make_tumor_normal(contig_len=30000, seed=5)
    loci = [(0, 0, 30000)]
    whole = oracle.somatic_standard(batch, ref, loci)
    shards = partitioning.shard(batch, loci, 4)
    assert sum(len(b.start) for _, b in shards.values()) > batch.n_reads  # halo reads are duplicated
    parts = [oracle.somatic_standard(b, ref, rs) for rs, b in shards.values()]
    merged = partitioning.gather_calls(parts)
    assert merged["n_visited"] == whole["n_visited"]
    assert np.array_equal(merged["calls"]["start"], whole["calls"]["start"])
    assert ss.alleles_of(merged) == ss.alleles_of(whole)
    np.testing.assert_allclose(merged["calls"]["somatic_log_odds"], whole["calls"]["somatic_log_odds"], rtol=1e-12)


GLOO_WORKER = r'''
import os, sys, pickle
sys.path.insert(0, os.environ["GUAC_ROOT"])
import numpy as np
import torch.distributed as dist
from guacamole_b200 import partitioning, synth
from oracle import oracle
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
ref, batch = synth.make_tumor_normal(contig_len=24000, seed=5)
loci = [(0, 0, 24000)]
ranges, mine = partitioning.shard(batch, loci, world)[rank]
# the per-shard compute stands in for the GPU path (tests may use the oracle); no data-path collective
res = oracle.somatic_standard(mine, ref, ranges)
gathered = [None] * world
dist.all_gather_object(gathered, res)   # host-side gather of sparse call records only
if rank == 0:
    merged = partitioning.gather_calls(gathered)
    whole = oracle.somatic_standard(batch, ref, loci)
    assert merged["n_visited"] == whole["n_visited"]
    assert np.array_equal(merged["calls"]["start"], whole["calls"]["start"])
    print("GLOO_OK", len(whole["calls"]))
dist.barrier()
dist.destroy_process_group()
'''


@pytest.mark.timeout(300)
def test_world_size_2_gloo_shard_and_gather(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(GLOO_WORKER)
    env = dict(os.environ, GUAC_ROOT=ROOT)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
                          "127.0.0.1", "--master-port", "29517", str(script)], capture_output=True, text=True, env=env, timeout=280)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "GLOO_OK" in out.stdout


def test_python_mirrors_match_the_header_layout(tmp_path):
    """The ctypes structures and numpy record types of native.py are hand-written twins of include/guac_b200.h: their
    sizes are checked against what a C compiler makes of the header (a field added on one side only shows up here, on
    the CPU, instead of as garbage on the GPU box)."""
    import ctypes as C
    import subprocess
    import numpy as np
    from guacamole_b200 import native
    pairs = {
        "guac_read_batch": C.sizeof(native.ReadBatchC), "guac_range": C.sizeof(native.RangeC),
        "guac_locus_counts": C.sizeof(native.LocusCountsC), "guac_element": C.sizeof(native.ElementC),
        "guac_allele_evidence": C.sizeof(native.AlleleEvidenceC), "guac_ss_params": C.sizeof(native.SSParamsC),
        "guac_ss_call": C.sizeof(native.SSCallC), "guac_locus_summary": C.sizeof(native.LocusSummaryC),
        "guac_gs_params": C.sizeof(native.GSParamsC), "guac_ga_params": C.sizeof(native.GAParamsC),
        "guac_read_batch_compact": C.sizeof(native.ReadBatchCompactC),
        "guac_called_allele": native.CALLED_ALLELE_DTYPE.itemsize, "guac_assembled_window": native.ASSEMBLED_WINDOW_DTYPE.itemsize,
    }
    from guacamole_b200.commands import somatic_joint as sj
    pairs.update({"guac_sj_params": C.sizeof(sj.SJParamsC), "guac_sj_sample_evidence": sj.SAMPLE_EVIDENCE_DTYPE.itemsize,
                  "guac_sj_allele_evidence": sj.ALLELE_EVIDENCE_DTYPE.itemsize})
    assert native.SS_CALL_DTYPE.itemsize == C.sizeof(native.SSCallC) and native.ELEMENT_DTYPE.itemsize == C.sizeof(native.ElementC)
    assert native.LOCUS_SUMMARY_DTYPE.itemsize == C.sizeof(native.LocusSummaryC)
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include "guac_b200.h"\nint main(void) {\n' +
                   "".join(f'  printf("{name} %zu\\n", sizeof({name}));\n' for name in pairs) + "  return 0;\n}\n")
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    out = dict(line.split() for line in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines())
    for name, size in pairs.items():
        assert int(out[name]) == size, (name, out[name], size)
    # the narrow columns sit at the end of the compact batch, after the wide ones
    assert native.ReadBatchCompactC.start_rel.offset > native.ReadBatchCompactC.sample_id.offset


def test_reference_arm_prints_the_contract_line(built):
    """`bench.py --impl reference` (the CPU restatement timed on the host cores, no GPU involved) prints one JSON line with
    the keys the driver reads; run here on a small bounded sample."""
    import json
    import subprocess
    import sys
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                          "--contig-mbp", "5", "--reference-step-seconds", "0.3"], check=True, capture_output=True, text=True, timeout=300)
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "loci/sec (pileup+likelihood)" and line["unit"] == "loci/s"
    assert line["higher_is_better"] is True and line["value"] > 0 and line["steps"] == 1 and line["n_gpus"] == 1
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1 and line["cpu_baseline"]["value"] == line["value"]
    assert line["e2e"] == {"value": line["value"], "unit": "loci/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
