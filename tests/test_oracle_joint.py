"""Pins the oracle's somatic-joint restatement against the reference's vectors (no GPU needed):
  SomaticJointCallerEndToEndSuite.scala:20-38 -> tiny-sjc-output/all.vcf (tests/golden/tiny_sjc.npz)
  ReadSubsequenceSuite.scala:28-73, PileupStatsSuite.scala:28-61 (inline reads)
"""
import numpy as np

from guacamole_b200.commands import somatic_joint as sj
from guacamole_b200.io import make_reference
from guacamole_b200.reads import ReadBatch, make_read
from oracle import oracle
from tests.golden import fixtures

KINDS = [sj.NORMAL_DNA, sj.TUMOR_DNA, sj.TUMOR_DNA]
SAMPLES = ["normal", "primary", "recurrence"]


def test_tiny_sjc_golden_vcf():
    batch, ref, extra = fixtures.load("tiny_sjc")
    loci = sj.loci_from_vcf_lines([str(x) for x in extra["force_call_vcf"]], batch.contig_names)
    evs, visited = oracle.somatic_joint(batch, ref, loci, sj.sj_params(KINDS), force_loci=loci)
    assert visited == 1 and len(evs) == 1
    rows = sj.vcf_rows(evs, batch.contig_names, SAMPLES, KINDS)
    assert rows == [str(x) for x in extra["golden_vcf"]]  # header line and the data row, character for character
    ev = evs[0]
    assert (ev["ref"], ev["alt"]) == ("G", "N") and ev["forced"] and not ev["is_germline_call"] and not ev["is_somatic_call"]
    assert [sum(c for _, c in s["allelic_depths"]) for s in ev["samples"]] == [23, 56, 65, 23, 121]  # DP column
    assert [s["total_depth"] for s in ev["samples"]] == [23, 59, 66, 23, 125]  # incl. reads contributing no allele


def test_java_format_helpers():
    assert sj.java_g8(-0.0072313716) == "-0.0072313716"
    assert sj.java_g8(-10.92369) == "-10.923690"
    assert sj.java_g8(-86.749695) == "-86.749695"
    assert sj.java_g8(1234567.891) == "1234567.9"
    assert sj.java_g8(0.00001234) == "1.2340000e-05"
    assert sj.format_vcf_double(0.0) == "0.00" and sj.format_vcf_double(0.375) == "0.375" and sj.format_vcf_double(1.5) == "1.50"
    assert sj.format_vcf_double(0.005) == "5.000e-03"


def subseqs(reads, locus, length):
    ref = make_reference(("chr1", 0, "NTCGATCGACG"))
    out = []
    for r in reads:
        b = ReadBatch.from_reads([r], ["chr1"], 1)
        out.append(oracle.read_subsequences(b, ref, "chr1", locus, length)[0])
    return out


def test_of_fixed_reference_length():  # ReadSubsequenceSuite.scala:28-45
    reads = [make_read("TCGATCGA", "8M", 1), make_read("TCGACCCTCGA", "4M3I4M", 1), make_read("TCGAGCGA", "8M", 1),
             make_read("TNGAGCGA", "8M", 1)]
    assert subseqs(reads[:1], 1, 1)[0][0] == "C"
    assert subseqs(reads[:1], 1, 2)[0][0] == "CG"
    assert subseqs(reads[1:2], 1, 6)[0][0] == "CGACCCTCG"
    assert subseqs(reads[3:4], 1, 1)[0][0] == "N"  # not all standard bases


def test_of_next_alt_allele():  # ReadSubsequenceSuite.scala:47-73
    assert subseqs([make_read("TCGATCGA", "8M", 1)], 1, 1)[0][1] is None
    assert subseqs([make_read("TCGAGCGA", "8M", 1)], 4, 1)[0][1] == "G"
    assert subseqs([make_read("TCGACCCTCGA", "4M3I4M", 1)], 3, 1)[0][1] == "ACCC"
    assert subseqs([make_read("TCGGCCCTCGA", "4M3I4M", 1)], 3, 1)[0][1] == "GCCC"
    assert subseqs([make_read("TCAGCCCTCGA", "4M3I4M", 1)], 2, 1)[0][1] == "AGCCC"
    assert subseqs([make_read("TNGAGCGA", "8M", 1)], 1, 1)[0][1] == "N"


def test_pileup_stats_likelihoods():  # PileupStatsSuite.scala:28-61
    ref = make_reference(("chr1", 0, "NTCGATCGA"))
    reads = [make_read("TCGATCGA", "8M", quality_scores=[10] * 8), make_read("TCGATCGA", "8M", quality_scores=[20] * 8)] + \
            [make_read("TCGCTCGA", "8M", quality_scores=[50] * 8) for _ in range(3)] + \
            [make_read("TCGACCCTCGA", "4M3I4M", quality_scores=[30] * 11)]
    batch = ReadBatch.from_reads(reads, ["chr1"], 1)
    total, depths, ll = oracle.pileup_stats(batch, ref, "chr1", 2, "G", [[("G", 1.0)], [("G", .99), ("C", .01)], [("T", 1.0)]])
    assert total == 6 and depths == [("G", 6)]
    assert ll[0] > ll[1] > ll[2]
    total, depths, ll = oracle.pileup_stats(batch, ref, "chr1", 3, "A", [[("A", .5), ("C", .5)], [("A", 1.0)], [("C", 1.0)]])
    assert dict(depths) == {"A": 2, "C": 3, "ACCC": 1} and [a for a, _ in depths][:1] == ["C"]
    assert ll[0] > ll[1] and ll[2] > ll[1]
    total, depths, _ = oracle.pileup_stats(batch, ref, "chr1", 4, "T")
    assert total == 6 and depths == [("T", 2)]  # reads with an SNV at position 4 don't count


def test_joint_calls_on_synthetic_four_samples_are_partition_invariant():
    from guacamole_b200 import synth
    contig_len, seed = 20000, 4005
    depths, tumor = (30, 30, 60, 60), (False, False, True, True)
    ref = synth.make_reference(contig_len, seed)
    batch = synth.generate_reads(contig_len, depths, tumor, seed)
    kinds = [sj.NORMAL_DNA, sj.NORMAL_DNA, sj.TUMOR_DNA, sj.TUMOR_DNA]
    whole, nv = oracle.somatic_joint(batch, ref, [(0, 1, contig_len)], sj.sj_params(kinds))
    assert len(whole) > 5 and any(e["is_somatic_call"] for e in whole) and any(e["is_germline_call"] for e in whole)
    edges = list(range(1, contig_len, 997)) + [contig_len]
    parts = []
    for a, b in zip(edges[:-1], edges[1:]):
        parts += oracle.somatic_joint(batch, ref, [(0, a, b)], sj.sj_params(kinds))[0]
    key = lambda e: (e["start"], e["ref"], e["alt"], e["is_germline_call"], e["is_somatic_call"])
    assert [key(e) for e in parts] == [key(e) for e in whole]


def test_germline_toy_caller_finds_planted_variants():
    """callPileupVariant (GermlineAssemblyCaller.scala:233-261) on a seeded 30x sample: planted germline SNVs
    (1 per kbp) are called, and the called genotype probability is a normalised probability."""
    from guacamole_b200 import synth
    from guacamole_b200.commands.somatic_standard import alleles_of
    contig_len, seed = 30000, 2000
    ref = synth.make_reference(contig_len, seed)
    batch = synth.generate_reads(contig_len, (30,), (False,), seed)
    out = oracle.germline_standard(batch, ref, [(0, 0, contig_len)])
    gk, _, _, _ = synth.site_info(seed, np.arange(contig_len))
    planted = set(np.flatnonzero(gk > 0)[np.flatnonzero(gk > 0) < contig_len - 300].tolist())
    called = set(int(x) for x in out["calls"]["start"])
    assert len(planted) > 10 and len(planted - called) <= 2  # a het site can lose by chance at 30x
    assert np.all(out["calls"]["genotype_log_probability"] <= 1e-12)
    assert all(len(r) == 1 and len(a) >= 1 for r, a in alleles_of(out) if len(r) == 1)
