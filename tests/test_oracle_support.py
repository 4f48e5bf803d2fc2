"""The oracle's restatement of VariantSupport.pileupsToAlleleCounts and of the VAFHistogram pipeline against the vectors
VariantSupportSuite.scala:84-122 holds (gatk_mini_bundle_extract.bam; fixture made by tests/golden/make_fixtures.py)."""
import numpy as np
import pytest

from guacamole_b200.reads import FLAG_DUPLICATE as GUAC_READ_DUPLICATE, ReadBatch
from oracle import oracle
from tests.golden import fixtures

# (locus, expected [(ref, alt, count)]) — 0-based loci as the suite writes them
SUITE_ALL_READS = [
    (10008951, [("CACACACACACA", "C", 1), ("C", "C", 4)]),   # "read evidence for simple snvs"
    (10006822, [("C", "", 6), ("C", "C", 2)]),               # "read evidence for mid-deletion"
    (10000624, [("T", "T", 6), ("T", "C", 1)]),              # "read evidence for simple snvs 2"
    (9999900, []),                                           # "... no filters": empty
    (9999995, [("A", "ACT", 9)]),
    (10007174, [("C", "T", 5), ("C", "C", 3)]),
    (10007175, [("T", "T", 8)]),
]
SUITE_NON_DUPLICATE = [(9999995, [("A", "ACT", 8)]), (10006822, [("C", "", 5), ("C", "C", 1)])]  # "... duplicate filtering"


def subset(batch: ReadBatch, keep: np.ndarray) -> ReadBatch:
    from guacamole_b200.reads import ReadBatch as RB
    idx = np.flatnonzero(keep)
    seq = [batch.bases_ascii[batch.seq_off[i]:batch.seq_off[i + 1]] for i in idx]
    qual = [batch.qual[batch.seq_off[i]:batch.seq_off[i + 1]] for i in idx]
    cig = [batch.cigar[batch.cigar_off[i]:batch.cigar_off[i + 1]] for i in idx]
    so = np.concatenate([[0], np.cumsum([len(x) for x in seq])]).astype(np.uint64)
    co = np.concatenate([[0], np.cumsum([len(x) for x in cig])]).astype(np.uint32)
    return RB(n_samples=batch.n_samples, contig_names=batch.contig_names, contig_id=batch.contig_id[idx], start=batch.start[idx],
              cigar_off=co, cigar=np.concatenate(cig).astype(np.uint32) if cig else np.zeros(0, np.uint32), seq_off=so,
              bases_ascii=np.concatenate(seq) if seq else np.zeros(0, np.uint8), qual=np.concatenate(qual) if qual else np.zeros(0, np.uint8),
              mapq=batch.mapq[idx], flags=batch.flags[idx], sample_id=batch.sample_id[idx])


def suite_cases():
    batch, ref, _ = fixtures.load("gatk_mini")
    nondup = subset(batch, (batch.flags & GUAC_READ_DUPLICATE) == 0)
    return [(batch, ref, SUITE_ALL_READS), (nondup, ref, SUITE_NON_DUPLICATE)]


def test_variant_support_suite_vectors():
    for batch, ref, cases in suite_cases():
        c20 = batch.contig_names.index("20")
        for locus, want in cases:
            got = oracle.variant_support(batch, ref, [(c20, locus, locus + 1)])
            assert sorted((r, a, n) for _, _, _, r, a, n in got) == sorted(want), locus


def test_vaf_histogram_arithmetic():
    """VariantLocus / roundToBin on the same reads: the frequencies follow from the suite's allele counts, the bins from
    (vaf * 100).toInt - percent % (100 / bins) in Float."""
    batch, ref, _ = fixtures.load("gatk_mini")
    c20 = batch.contig_names.index("20")
    hist, tot, recs = oracle.vaf_histogram(batch, ref, [(c20, 10007174, 10007176)], bins=20)
    assert [(x, float(v)) for _, _, x, v in recs] == [(10007174, float(np.float32(5) / np.float32(8)))]  # 10007175 is all reference
    assert tot.tolist() == [1] and hist[0, 60] == 1 and hist.sum() == 1  # 62 % -> bin 60
    hist, tot, recs = oracle.vaf_histogram(batch, ref, [(c20, 10007174, 10007176)], bins=20, min_vaf_percent=63)
    assert tot.tolist() == [0] and not recs
    hist, tot, recs = oracle.vaf_histogram(batch, ref, [(c20, 10007174, 10007176)], bins=20, min_read_depth=9)
    assert tot.tolist() == [0]
    with pytest.raises(AssertionError):
        oracle.vaf_histogram(batch, ref, [(c20, 10007174, 10007176)], bins=0)
