"""The oracle's restatement of the coverage-aware partitioner against the reference's own vectors:
TakeLociIteratorSuite.scala:21-101 and ContigCoverageIteratorSuite.scala:39-190."""
import pytest

from oracle import oracle

CONTIGS = {"chr1": 0, "chr2": 1, "chr3": 2, "chr4": 3}

# (name, trim_ranges, [((contig, locus), (depth, starts))], expected loci sets as the suite writes them)
TAKE_LOCI_CASES = [
    ("simple", False,
     [(("chr1", 10), (1, 1)), (("chr1", 11), (6, 5)), (("chr1", 12), (15, 10)), (("chr1", 13), (10, 10)), (("chr1", 14), (10, 5)),
      (("chr1", 15), (11, 1)), (("chr2", 20), (1, 1)), (("chr2", 21), (4, 3)), (("chr2", 22), (11, 7)), (("chr2", 23), (14, 3)),
      (("chr2", 24), (15, 1))],
     ["chr1:10-12", "chr1:12-13", "chr1:13-15", "chr1:15-16,chr2:20-22", "chr2:22-25"]),
    ("skip extreme depth", False,
     [(("chr1", 10), (1, 1)), (("chr1", 11), (6, 5)), (("chr1", 12), (15, 10)), (("chr1", 13), (16, 10)), (("chr1", 14), (11, 5)),
      (("chr1", 15), (12, 1))],
     ["chr1:10-12", "chr1:12-13", "chr1:14-16"]),
    ("skip extreme depth at end", False,
     [(("chr1", 10), (1, 1)), (("chr1", 11), (6, 5)), (("chr1", 12), (15, 10)), (("chr1", 13), (16, 10))],
     ["chr1:10-12", "chr1:12-13"]),
    ("partition spanning several contigs", False,
     [(("chr1", 10), (1, 1)), (("chr1", 11), (6, 5)), (("chr2", 1), (2, 2)), (("chr3", 7), (3, 3)), (("chr3", 9), (4, 3)),
      (("chr4", 20), (2, 2))],
     ["chr1:10-12,chr2:1-2,chr3:7-10", "chr4:20-21"]),
    ("partition spanning several contigs, trim ranges", True,
     [(("chr1", 10), (1, 1)), (("chr1", 11), (6, 5)), (("chr2", 1), (2, 2)), (("chr3", 7), (3, 3)), (("chr3", 9), (4, 3)),
      (("chr4", 20), (2, 2))],
     ["chr1:10-12,chr2:1-2,chr3:7-8,chr3:9-10", "chr4:20-21"]),
]


def as_strings(sets, names):
    return [",".join(f"{names[c]}:{a}-{b}" for c, a, b in s) for s in sets]


def coverages_of(case):
    return [(CONTIGS[c], x, d, s) for (c, x), (d, s) in case]


@pytest.mark.parametrize("name,trim,inputs,expected", TAKE_LOCI_CASES, ids=[c[0] for c in TAKE_LOCI_CASES])
def test_take_loci_suite(name, trim, inputs, expected):
    names = {v: k for k, v in CONTIGS.items()}
    assert as_strings(oracle.take_loci(coverages_of(inputs), 15, trim), names) == expected


READS = [(10, 20), (11, 21), (11, 21), (30, 40)]  # ContigCoverageIteratorSuite.scala:84-92


def test_contig_coverage_suite():
    assert oracle.contig_coverage([(10, 20)]) == [(x, 1, 1 if x == 10 else 0) for x in range(10, 20)]  # "simple"
    assert oracle.contig_coverage([(0, 10), (1, 9), (2, 8), (3, 11)]) == [  # "contained reads"
        (0, 1, 1), (1, 2, 1), (2, 3, 1), (3, 4, 1), (4, 4, 0), (5, 4, 0), (6, 4, 0), (7, 4, 0), (8, 3, 0), (9, 2, 0), (10, 1, 0)]
    # "skips": the iterator is evaluated at 9, 10, then skipped to 15, 16, 21, and from 25 on
    got = dict((x, (d, s)) for x, d, s in oracle.contig_coverage(READS, 1, [9, 10, 15, 16, 21] + list(range(25, 31)) + [39] + [45]))
    assert got[9] == (1, 1) and got[10] == (3, 2) and got[15] == (3, 0) and got[16] == (3, 0) and got[21] == (2, 0)
    assert min(x for x in got if x >= 25) == 29 and got[29] == (1, 1) and got[30] == (1, 0) and got[39] == (1, 0) and 45 not in got
    assert oracle.contig_coverage(READS, 1, [15, 35])[:2] == [(15, 3, 3), (35, 1, 1)]      # "skips+starts"
    assert oracle.contig_coverage(READS, 1, [21])[0] == (21, 2, 2)                         # "skip completely over read"
    assert oracle.contig_coverage(READS, 1, range(25, 50))[0] == (29, 1, 1)                # "skip completely over reads"
    assert oracle.contig_coverage(READS, 1, range(12, 15)) == [(12, 3, 3), (13, 3, 0), (14, 3, 0)]  # "intersection"
    with pytest.raises(ValueError):                                                        # "throw on unsorted regions"
        oracle.contig_coverage([(0, 10), (1, 9), (2, 21), (1, 31)])
