"""Pins the CPU oracle against the reference's own golden vectors (no GPU needed).

Sources, all in the reference checkout's src/test:
  SomaticStandardEndToEndSuite.scala:13-37  -> tough.golden.vcf        (tests/golden/tough.npz)
  LikelihoodSuite.scala:155-273             -> likelihood vectors, eps 1e-12
  PileupSuite.scala:36-438                  -> alignment / cursor vectors
  PileupFlatMapUtilsSuite.scala:124-256     -> loci visited, element strings, partition invariance
"""
import math

import numpy as np
import pytest

from guacamole_b200.commands import somatic_standard as ss
from guacamole_b200.io import make_reference
from guacamole_b200.reads import ReadBatch, make_read
from oracle import oracle
from tests.golden import fixtures

EPS = 1e-12  # LikelihoodSuite.scala:22


def test_tough_golden_vcf():
    batch, ref, extra = fixtures.load("tough")
    ranges = ss.covered_ranges(batch)
    out = oracle.somatic_standard(batch, ref, ranges)
    kept = ss.somatic_genotype_filter(out["calls"])
    rows = ss.vcf_rows(out, batch.contig_names, kept)
    assert rows == [str(x) for x in extra["golden_vcf"]]  # VCFCmpTest.checkVCFs: all non-## lines equal


def test_tough_multithreaded_partitions_agree():
    batch, ref, _ = fixtures.load("tough")
    ranges = ss.covered_ranges(batch)
    one = oracle.somatic_standard(batch, ref, ranges, n_threads=1, want_summary=True)
    many = oracle.somatic_standard(batch, ref, ranges, n_threads=7, want_summary=True)
    assert np.array_equal(one["summary"], many["summary"])
    assert np.array_equal(one["calls"]["start"], many["calls"]["start"])
    np.testing.assert_allclose(one["calls"]["somatic_log_odds"], many["calls"]["somatic_log_odds"], rtol=1e-12)


# ---- LikelihoodSuite ----------------------------------------------------------------------------
E30, E40 = 10 ** -3.0, 10 ** -4.0
S30, S40 = 1 - E30, 1 - E40
HALF = 0.5


def lik_fixture(bases_quals):
    ref = make_reference(("chr1", 1, "C"))
    reads = [make_read(b, "1M", 1, "chr1", quality_scores=[q]) for b, q in bases_quals]
    return ReadBatch.from_reads(reads, ["chr1"], 1), ref


def genotype(a, b):
    alleles = list(dict.fromkeys([a, b]))
    return [(("C", x), 1.0 / len(alleles)) for x in alleles]


@pytest.mark.parametrize("reads,expected", [
    ([("C", 30), ("C", 40), ("C", 30)], {("C", "C"): S30 * S40 * S30, ("C", "A"): HALF ** 3, ("A", "C"): HALF ** 3,
                                         ("A", "A"): E30 * E40 * E30, ("A", "T"): E30 * E40 * E30}),
    ([("C", 30), ("C", 40), ("A", 30)], {("C", "C"): S30 * S40 * E30, ("C", "A"): HALF ** 3, ("A", "C"): HALF ** 3,
                                         ("A", "A"): E30 * E40 * S30, ("A", "T"): E30 * E40 * HALF, ("T", "T"): E30 * E40 * E30}),
    ([("C", 30), ("A", 40), ("A", 30)], {("C", "C"): S30 * E40 * E30, ("C", "A"): HALF ** 3, ("A", "C"): HALF ** 3,
                                         ("A", "A"): E30 * S40 * S30, ("A", "T"): E30 * HALF * HALF, ("T", "T"): E30 * E40 * E30}),
    ([("A", 30), ("A", 40), ("A", 30)], {("C", "C"): E30 * E40 * E30, ("C", "A"): HALF ** 3, ("A", "C"): HALF ** 3,
                                         ("A", "A"): S30 * S40 * S30, ("A", "T"): HALF ** 3, ("T", "T"): E30 * E40 * E30}),
])
def test_likelihood_of_genotypes(reads, expected):  # LikelihoodSuite.scala:155-200
    batch, ref = lik_fixture(reads)
    gts = list(expected)
    got = oracle.likelihoods_of_genotypes(batch, ref, "chr1", 1, [genotype(*g) for g in gts])
    for g, v in zip(gts, got):
        assert abs(v - expected[g]) <= EPS, g


@pytest.mark.parametrize("log_space", [False, True])
@pytest.mark.parametrize("reads,expected", [
    ([("C", 30), ("C", 40), ("C", 30)], {("C", "C"): S30 * S40 * S30}),
    ([("C", 30), ("C", 40), ("A", 30)], {("C", "C"): S30 * S40 * E30, ("A", "C"): 1 / 8.0, ("A", "A"): E30 * E40 * S30}),
    ([("A", 30), ("A", 40), ("A", 30)], {("A", "A"): S30 * S40 * S30}),
])
def test_all_possible_genotypes_normalised(reads, expected, log_space):  # LikelihoodSuite.scala:202-273
    batch, ref = lik_fixture(reads)
    got = oracle.all_genotypes(batch, ref, "chr1", 1, log_space=log_space)
    total = sum(expected.values())
    assert len(got) == len(expected)
    for (r1, a1), (r2, a2), v in got:
        key = (a1, a2) if (a1, a2) in expected else (a2, a1)
        want = expected[key] / total
        assert abs(v - (math.log(want) if log_space else want)) <= EPS


# ---- PileupSuite -----------------------------------------------------------------------------------
PILEUP_REF = [("chr1", 0, "CTCGATCGACG"), ("chr2", 0, "AATTG"), ("chr3", 0, "AAATTT"), ("chr4", 0, "AATTGCAATTG"),
              ("artificial", 0, "A" * 34 + "G" * 10 + "A" * 5 + "G" * 15 + "A" * 15 + "ACGT" * 10)]
MATCH, MISMATCH, INSERTION, DELETION, MID_DELETION, CLIPPED = range(6)


def alleles_at(reads, contig, locus, contigs=None):
    ref = make_reference(*PILEUP_REF)
    batch = ReadBatch.from_reads(reads, contigs or [c for c, _, _ in PILEUP_REF], 1)
    return oracle.pileup_alleles(batch, ref, contig, locus)


def test_long_insert_reads():  # PileupSuite.scala:36-118
    quals = [10, 15, 20, 25, 10, 15, 20, 25]
    reads = [make_read("TCGATCGA", "8M", 1, "chr1", quality_scores=quals), make_read("TCGATCGA", "8M", 1, "chr1", quality_scores=quals),
             make_read("TCGACCCTCGA", "4M3I4M", 1, "chr1", quality_scores=[10, 15, 20, 25, 5, 5, 5, 10, 15, 20, 25])]
    first = alleles_at(reads, "chr1", 1)
    assert all(k == MATCH for _, _, k, _ in first)
    at4 = alleles_at(reads, "chr1", 4)
    assert [(r, a, k) for r, a, k, _ in at4] == [("A", "A", MATCH), ("A", "A", MATCH), ("A", "ACCC", INSERTION)]
    assert [q for _, _, _, q in at4] == [25, 25, 5]  # insertion quality = min(25, 5, 5, 5)
    assert all(k == MATCH and q == 10 for _, _, k, q in alleles_at(reads, "chr1", 5))
    assert all(a == "G" and k == MATCH for _, a, k, _ in alleles_at(reads, "chr1", 7))
    assert all(a == "A" and k == MATCH and q == 25 for _, a, k, q in alleles_at(reads, "chr1", 8))


def test_default_quality_is_31():  # ReadsUtil.scala:74 ('@' = Phred 31), PileupSuite.scala:44-46
    reads = [make_read("TCGATCGA", "8M", 1, "chr1"), make_read("TCGACCCTCGA", "4M3I4M", 1, "chr1")]
    assert all(q == 31 for _, _, _, q in alleles_at(reads, "chr1", 4))


def test_insertion_at_contig_start():  # PileupSuite.scala:163-167
    got = alleles_at([make_read("AAAAAACGT", "5I4M", 0, "chr1")], "chr1", 0)
    assert got == [("A", "AAAAAA", INSERTION, 31)]


def test_deletion_cursor():  # PileupSuite.scala:183-207
    read = [make_read("AATTGAATTG", "5M1D5M", 0, "chr4")]
    assert alleles_at(read, "chr4", 0)[0][2] == MATCH
    assert alleles_at(read, "chr4", 4)[0][:3] == ("GC", "G", DELETION)
    assert alleles_at(read, "chr4", 5)[0][:3] == ("C", "", MID_DELETION)
    assert alleles_at(read, "chr4", 6)[0][2] == MATCH
    assert alleles_at(read, "chr4", 9)[0][2] == MATCH


def test_pileup_in_the_middle_of_a_deletion():  # PileupSuite.scala:421-438
    reads = [make_read("TCGAAAAGCT", "5M6D5M", 0, "chr1") for _ in range(3)]
    assert {a[:2] for a in alleles_at(reads, "chr1", 4)} == {("ATCGACG", "A")}
    assert {a[:2] for a in alleles_at(reads, "chr1", 5)} == {("T", "")}


def test_same_start_reads_fixture():  # PileupSuite.scala:120-130,207-232
    batch, _, _ = fixtures.load("same_start_reads")
    c = batch.contig_names[0]
    ref = make_reference((c, 0, "A" * 100))
    assert len(oracle.pileup_alleles(batch, ref, c, 0)) == 10
    for i in range(1, 60):
        assert len(oracle.pileup_alleles(batch, ref, c, i)) == 10
    at9 = oracle.pileup_alleles(batch, ref, c, 9)
    assert sum(1 for r, a, k, _ in at9 if k == DELETION and r == "A" * 11) == 5
    for i in range(10, 20):
        assert sum(1 for _, _, k, _ in oracle.pileup_alleles(batch, ref, c, i) if k == MID_DELETION) == 5
    for i in range(60, 70):
        assert len(oracle.pileup_alleles(batch, ref, c, i)) == 5


def test_different_start_reads_fixture():  # PileupSuite.scala:249-381
    batch, _, _ = fixtures.load("different_start_reads")
    contig = batch.contig_names[0]
    ref = make_reference((contig, 0, "A" * 34 + "G" * 10 + "A" * 5 + "G" * 15 + "A" * 15 + "ACGT" * 10 + "A" * 200))

    # the fixture is sorted by start; the suite indexes reads in file order (read1..read7)
    file_order = {0: 0, 1: 1, 2: 3, 3: 4, 4: 2, 5: 5, 6: 6}

    def read_alleles(i, locus):  # element of the suite's testAdamRecords(i) at locus
        sub = ReadBatch.from_reads([batch.read(file_order[i])], batch.contig_names, 1)
        return oracle.pileup_alleles(sub, ref, contig, locus)

    # read1: starts at 5, CIGAR 29M10D31M
    assert read_alleles(0, 5)[0][1] == "A"
    assert read_alleles(0, 5 + 28)[0][:3] == ("AGGGGGGGGGG", "A", DELETION)
    assert read_alleles(0, 5 + 29)[0][1] == "" and read_alleles(0, 5 + 38)[0][1] == ""
    assert read_alleles(0, 5 + 39)[0][1] == "A"
    assert read_alleles(0, 74) and not read_alleles(0, 75) and not read_alleles(0, 4)
    # read3
    assert [read_alleles(2, x)[0][1] for x in (15, 16, 17, 18)] == ["A", "T", "C", "G"]
    # read4: 10M10I10D40M, ACGT repeated
    for i in range(2):
        assert [read_alleles(3, 20 + i * 4 + k)[0][1][0] for k in range(4)] == list("ACGT")
    at29 = read_alleles(3, 29)[0]
    assert at29[2] == INSERTION and at29[1] == "CGTACGTACGT"
    # read5: 5M4=1X5= over [10, 25)
    assert [read_alleles(4, x)[0][1] for x in (10, 14, 18, 19, 20, 21, 22, 24)] == list("AAACGTAG")
    # read6 / read7: 4=1D4=4S (the file has D where the suite's comment says N) and 4=1N4=4H
    for i, kind in ((5, MID_DELETION), (6, CLIPPED)):
        assert [read_alleles(i, x)[0][1] for x in (99, 100, 101, 102, 103, 104, 107)] == ["A", "C", "G", "T", "", "A", "T"]
        assert read_alleles(i, 103)[0][2] == kind


def test_rna_read_with_intron():  # PileupSuite.scala:383-405
    start = 229538779
    ref = make_reference(("1", 229538778, "A" * 2000), contig_length=249250621)
    read = make_read("ACGTAGCCTAGGCCTTCGACACTGGGGGGCTGAGGGAAGGGGCACCTGTC", "7M1000N43M", start, "1")
    b = ReadBatch.from_reads([read], ["1"], 1)
    assert oracle.pileup_alleles(b, ref, "1", start + 1)[0][1] == "C"
    assert oracle.pileup_alleles(b, ref, "1", start + 2)[0][1] == "G"
    assert oracle.pileup_alleles(b, ref, "1", start + 100)[0][:3] == ("", "", CLIPPED)
    assert oracle.pileup_alleles(b, ref, "1", start + 1049)[0][1] == "C"


def test_rna_reads_fixture_depths():  # PileupSuite.scala:407-419
    batch, _, _ = fixtures.load("testrna")
    lo = int(batch.start.min())
    ref = make_reference(("1", lo, "A" * (int(batch.ends().max()) - lo + 10)), contig_length=249250621)
    _, loci, cnt = oracle.pileup_counts(batch, ref, [(0, 229580594, 229580595), (0, 229580706, 229580708)])
    assert list(loci) == [229580594, 229580706, 229580707]
    assert [int(c["depth"]) for c in cnt[:, 0]] == [94, 4, 1]


# ---- PileupFlatMapUtilsSuite ---------------------------------------------------------------------------
FLATMAP_REF = [("chr1", 0, "ATCGATCGA" + "N" * 90), ("chr1", 99, "ACGTACGTACGT" + "N" * 500), ("chr2", 10, "N" * 10)]


def flatmap_reads(n_samples):
    r1 = [make_read("TCGATCGA", "8M", 1, sample_id=0)] * 3 + [make_read("GGGGGGGG", "8M", 100, sample_id=0)] * 3
    r2 = [make_read(s, "8M", 1, sample_id=1) for s in ("AAAAAAAA", "CCCCCCCC", "TTTTTTTT")] + [make_read("XYX", "3M", 99, sample_id=1)]
    r3 = [make_read(s, "8M", 1, sample_id=2) for s in ("AAGGCCTT", "GGAATTCC", "GGGGGGGG")] + [make_read("XZX", "3M", 99, sample_id=2)]
    return ReadBatch.from_reads((r1 + r2 + r3)[:6 + 4 * (n_samples - 1)], ["chr0", "chr1", "chr2"], n_samples)


def test_flatmap_skip_empty_loci():  # PileupFlatMapUtilsSuite.scala:124-138
    ref = make_reference(*FLATMAP_REF)
    batch = ReadBatch.from_reads([make_read("TCGATCGA", "8M", 1)] * 3, ["chr0", "chr1", "chr2"], 1)
    ranges = [(0, 5, 10), (1, 0, 100), (2, 0, 1000), (2, 5000, 6000)]
    _, loci, _ = oracle.pileup_counts(batch, ref, ranges, skip_empty=True)
    assert list(loci) == [1, 2, 3, 4, 5, 6, 7, 8]
    _, loci, cnt = oracle.pileup_counts(batch, ref, [(1, 1, 9)], skip_empty=False)
    assert len(loci) == 8 and all(int(c["reference_depth"]) == 3 for c in cnt[:, 0])  # :80-100


def test_flatmap_two_samples_loci():  # PileupFlatMapUtilsSuite.scala:140-177
    ref = make_reference(*FLATMAP_REF)
    batch = flatmap_reads(2)
    _, loci, _ = oracle.pileup_counts(batch, ref, [(0, 0, 1000), (1, 1, 500), (2, 10, 20)])
    assert list(loci) == [1, 2, 3, 4, 5, 6, 7, 8, 99, 100, 101, 102, 103, 104, 105, 106, 107]


def element_strings(batch, ref, ranges, skip_empty=True):
    _, loci, off, els, order = oracle.pileup_elements(batch, ref, ranges, skip_empty=skip_empty)
    out = []
    for i in range(len(loci)):
        e = els[off[i]:off[i + 1]]
        out.append([[chr(b) for b, s in zip(e["base"], e["sample_id"]) if s == k] for k in range(batch.n_samples)])
    return out


def test_flatmap_three_samples_element_strings_and_partition_invariance():  # PileupFlatMapUtilsSuite.scala:179-256
    ref = make_reference(*FLATMAP_REF)
    batch = flatmap_reads(3)
    plain = element_strings(batch, ref, [(1, 1, 500), (2, 10, 20)])
    # UniformPartitioner(800): one range per locus
    fine = element_strings(batch, ref, [(1, x, x + 1) for x in range(1, 500)] + [(2, x, x + 1) for x in range(10, 20)])
    assert plain == fine
    assert plain[0] == [["T", "T", "T"], ["A", "C", "T"], ["A", "G", "G"]]
    assert plain[1] == [["C", "C", "C"], ["A", "C", "T"], ["A", "G", "G"]]
    assert plain[2] == [["G", "G", "G"], ["A", "C", "T"], ["G", "A", "G"]]
    assert plain[3] == [["A", "A", "A"], ["A", "C", "T"], ["G", "A", "G"]]
    assert plain[8] == [[], ["X"], ["X"]]
    assert plain[9] == [["G", "G", "G"], ["Y"], ["Z"]]
    assert plain[10] == [["G", "G", "G"], ["X"], ["X"]]
    with_empty = element_strings(batch, ref, [(1, 1, 500), (2, 10, 20)], skip_empty=False)
    assert with_empty[:4] == plain[:4] and with_empty[35] == [[], [], []]
