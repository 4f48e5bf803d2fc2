"""Vectors of the reference's caller unit suites, shared by the oracle (CPU) and device (GPU) tests:
  SomaticStandardCallerSuite.scala:17-207   findPotentialVariantAtLocus on hand-made pileups (odds thresholds 2 / 2)
  SomaticJointCallerSuite.scala:33-57,112-130   makeCalls on the cancer-wgs1 BAMs: the two tests whose loci lie inside the
                                            tiny BAMs of the checkout (tests/golden/tiny_sjc.npz covers chr12:65856761-65857220);
                                            the suite's other tests read normal.bam / primary.bam / recurrence.bam and
                                            chr22.fa.gz, of which only the .bai files are present
`run` is either the oracle's or the device's entry point with the same signature.
"""
from guacamole_b200.commands import somatic_joint as sj
from guacamole_b200.commands.somatic_standard import alleles_of
from guacamole_b200.io import make_reference
from guacamole_b200.native import ss_params
from guacamole_b200.reads import ReadBatch, make_read

# ReferenceUtil.makeReference(sc, ("chr1", 0, ...), ...) of SomaticStandardCallerSuite.scala:17-24
SS_REFERENCE = (("chr1", 0, "TCGATCGACG"), ("chr2", 0, "TCGAAGCTTCG"), ("chr3", 10, "TCGAATCGATCGATCGA"), ("chr4", 0, "TCGAAGCTTCGAAGCT"))

# (name, contig, normal (sequence, cigar, start), tumor (sequence, cigar, start), [(locus, expected (ref, alt) or None)])
SS_CASES = [
    ("no indels", "chr1", ("TCGATCGA", "8M", 0), ("TCGGTCGA", "8M", 0), [(2, None)]),
    ("single-base deletion", "chr1", ("TCGATCGA", "8M", 0), ("TCGTCGA", "3M1D4M", 0), [(2, ("GA", "G"))]),
    ("multiple-base deletion", "chr4", ("TCGAAGCTTCGAAGCT", "16M", 0), ("TCGAAAAGCT", "5M6D5M", 0), [(4, ("AGCTTCG", "A"))]),
    ("single-base insertion", "chr1", ("TCGATCGA", "8M", 0), ("TCGAGTCGA", "4M1I4M", 0), [(3, ("A", "AG"))]),
    ("multiple-base insertion", "chr1", ("TCGATCGA", "8M", 0), ("TCGAGGTCTCGA", "4M4I4M", 0), [(3, ("A", "AGGTC"))]),
    ("insertions and deletions", "chr3", ("TCGAATCGATCGATCGA", "17M", 10), ("TCATCTCAAAAGAGATCGA", "2M2D1M2I2M4I2M2D6M", 10),
     [(11, ("CGA", "C")), (14, ("A", "ATC")), (16, ("C", "CAAAA")), (18, ("ATC", "A"))]),
]


def ss_case_inputs(case):
    _, contig, normal, tumor, expectations = case
    reads = [make_read(normal[0], normal[1], start=normal[2], chr=contig, sample_id=0) for _ in range(3)]
    reads += [make_read(tumor[0], tumor[1], start=tumor[2], chr=contig, sample_id=1) for _ in range(3)]
    names = [c for c, _, _ in SS_REFERENCE]
    return ReadBatch.from_reads(reads, names, 2), make_reference(*SS_REFERENCE), names.index(contig), expectations


def check_ss_case(case, run):
    """run(batch, reference, ranges, params) -> result dict with calls / arena."""
    batch, ref, cid, expectations = ss_case_inputs(case)
    p = ss_params(normal_odds_threshold=2, tumor_odds_threshold=2)
    for locus, expected in expectations:
        got = alleles_of(run(batch, ref, [(cid, locus, locus + 1)], p))
        want = [] if expected is None else [(expected[0].encode(), expected[1].encode())]
        assert got == want, (case[0], locus, got)


# ---- SomaticJointCallerSuite over tests/golden/tiny_sjc.npz ------------------------------------------------------------
KINDS3 = [sj.NORMAL_DNA, sj.TUMOR_DNA, sj.TUMOR_DNA]


def check_joint_cases(batch, ref, run):
    """run(batch, reference, loci, params, force_loci) -> (evidence dicts, loci visited)."""
    names = batch.contig_names
    # "force-call a non-variant locus" (:33-57)
    loci = [(names.index("chr12"), 65857040, 65857041)]
    evs, _ = run(batch, ref, loci, sj.sj_params(KINDS3), loci)
    assert len(evs) == 1 and (evs[0]["start"], evs[0]["ref"]) == (65857040, "G")
    # "don't call variants with N as the reference base" (:112-130)
    loci = [(names.index("chr12"), 65857030, 65857080)]
    n_ref = make_reference(("chr12", 65856930, "N" * 250), contig_length=70000000)
    evs, _ = run(batch, n_ref, loci, sj.sj_params(KINDS3), loci)
    assert evs == []


# ---- SomaticStandardCallerRealDataSuite.scala:79-150 over tests/golden/real_*.npz ---------------------------------------
# fixture -> positions where a variant must be found; every other stored position must yield none
REAL_POSITIVE = {"syn1fp": [], "complexvar": [82949713, 130919744], "simplefp": []}


def test_somatic_genotype_filter(calls, min_tumor_read_depth=8, max_tumor_read_depth=200, min_normal_read_depth=4,
                                 min_tumor_alternate_read_depth=3, min_vaf=5):
    """filters/somatic/TestSomaticGenotypeFilter.scala:11-40 with the suite's constants (minLikelihood = 0: skipped)."""
    import numpy as np
    t, n = calls["tumor_variant_evidence"], calls["normal_reference_evidence"]
    keep = (t["read_depth"] >= min_tumor_read_depth) & (t["read_depth"] < max_tumor_read_depth) & (n["read_depth"] >= min_normal_read_depth)
    vaf = t["allele_read_depth"].astype(np.float32) / t["read_depth"].astype(np.float32)
    keep &= vaf.astype(np.float64) * 100.0 > min_vaf
    keep &= t["allele_read_depth"] >= min_tumor_alternate_read_depth
    return calls[keep]


test_somatic_genotype_filter.__test__ = False  # a helper, not a test


def check_real_data_case(name, batch, ref, positions, run):
    """run(batch, reference, ranges, params) -> result dict; logOddsThreshold = 7 for both samples (:27,55-60)."""
    p = ss_params(normal_odds_threshold=7, tumor_odds_threshold=7)
    cid = int(batch.contig_id[0])
    assert (batch.contig_id == cid).all()
    for pos in [int(x) for x in positions]:
        result = run(batch, ref, [(cid, pos, pos + 1)], p)
        found = len(test_somatic_genotype_filter(result["calls"])) > 0
        assert found == (pos in REAL_POSITIVE[name]), (name, pos, alleles_of(result))


# ---- AlleleEvidenceSuite.scala:14-58 through callPileupVariant ----------------------------------------------------------
# reads (sequence, cigar, mapping quality) at start 1 over "NTAGATCGA"; evidence of A -> C at locus 2:
# (meanMappingQuality, medianMappingQuality, medianMismatchesPerRead).  The suite's third case (no supporting read: NaN
# everywhere) cannot come out of a caller, which only reports alleles some read carries.
ALLELE_EVIDENCE_CASES = [
    ([("TCGATCGA", "8M", 30), ("TCGATCGA", "8M", 30), ("TCGACCCTCGA", "4M3I4M", 60)], (40.0, 30.0, 1.0)),
    ([("TAGATCGA", "8M", 30), ("TCGATCGA", "8M", 60), ("TAGACCCTCGA", "4M3I4M", 60)], (60.0, 60.0, 1.0)),
]


def check_allele_evidence_case(case, run):
    """run(batch, reference, ranges, params) -> germline_standard result."""
    from guacamole_b200.native import gs_params
    reads, (mean_mq, median_mq, median_mm) = case
    batch = ReadBatch.from_reads([make_read(s, c, start=1, alignment_quality=q) for s, c, q in reads], ["chr1"], 1)
    out = run(batch, make_reference(("chr1", 0, "NTAGATCGA")), [(0, 2, 3)], gs_params(min_alt_reads=0))
    assert alleles_of(out) == [(b"A", b"C")]
    ev = out["calls"]["evidence"][0]
    assert (ev["mean_mapping_quality"], ev["median_mapping_quality"], ev["median_mismatches_per_read"]) == (mean_mq, median_mq, median_mm)
