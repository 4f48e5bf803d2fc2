"""Pins the oracle's germline-assembly components against the reference's vectors (no GPU needed):
  DeBruijnGraphSuite.scala:15-457   k-mer counts, children/parents order, unipath merge, mergeIndex, DFS, real reads
  ReadAlignmentSuite.scala:78-158   scores and CIGAR strings (left-aligned indels)
The end-to-end GermlineAssemblyCallerSuite needs fixtures that are absent from the checkout; the caller
closure is exercised on seeded synthetic reads with planted indels instead (unpinned).
"""
import os

import numpy as np
import pytest

from oracle import oracle

REF52 = "GAGGATCTGCCATGGCCGGGCGAGCTGGAGGAGCGAGGAGGAGGCAGGAGGA"


def test_merge_overlapping_and_counts():  # DeBruijnGraphSuite.scala:15-60
    assert oracle.debruijn(["TCATCTCAAAAGAGATCGA"], 8, "counts")["TCATCTCA"] == 1
    c = oracle.debruijn(["TCATCTTAAAAGACATAAA"], 3, "counts")
    assert (c["TCA"], c["CAT"], c["AAA"]) == (1, 2, 3)
    assert len(oracle.debruijn(["AAATCCCTTTTA"], 4, "counts")) == 9  # all unique k-mers (:134-142)


def test_kmer_quality_gate():  # DeBruijnGraphSuite.scala:62-104
    seq = "TCATCTTAAAAGACATAAA"
    c = oracle.debruijn([seq, seq], 3, "counts", min_mean_kmer_quality=15, qualities=[[30] * 19, [10] * 19])
    assert (c["TCA"], c["CAT"], c["AAA"]) == (1, 2, 3)
    # The suite's listed maps follow the qualities its comments describe (9 bases at 30, 6 at 5, 4 at 30); its
    # qualityScores expression puts 10 bases at 30, for which AAA is one higher. The suite compares with a bare
    # `===`, which asserts nothing, so both layouts are checked here against the rule of DeBruijnGraph.scala:357-362.
    for mixed, aaa_low, aaa_high in (([30] * 9 + [5] * 6 + [30] * 4, 5, 4), ([30] * 10 + [5] * 5 + [30] * 4, 6, 5)):
        low = oracle.debruijn([seq, seq], 3, "counts", min_mean_kmer_quality=15, qualities=[[30] * 19, mixed])
        assert (low["TCA"], low["CAT"], low["AAA"]) == (2, 3, aaa_low)
        high = oracle.debruijn([seq, seq], 3, "counts", min_mean_kmer_quality=25, qualities=[[30] * 19, mixed])
        assert (high["TCA"], high["CAT"], high["AAA"]) == (2, 3, aaa_high)


def test_unique_paths_and_bubbles():  # DeBruijnGraphSuite.scala:144-232
    seq = "AAATCCCTGGGT"
    fwd = oracle.debruijn([seq], 4, "forward", source="AAAT")
    assert len(fwd) == 9
    bwd = oracle.debruijn([seq], 4, "backward", source="GGGT")
    assert len(bwd) == 9 and bwd[0] == "AAAT" and bwd[-1] == "GGGT"
    assert len(oracle.debruijn([seq, "AAATCCCTGGAT"], 4, "forward", source="AAAT")) == 7     # bubble at end
    assert oracle.debruijn([seq, "AAATCGCTGGGT"], 4, "forward", source="AAAT") == ["AAAT", "AATC"]  # bubble in middle
    assert len(oracle.debruijn([seq, "ACATCCCTGGGT"], 4, "forward", source="AAAT")) == 2     # bubble in first k-mer
    assert oracle.debruijn([seq, "AAATCCCTGGAT"], 4, "backward", source="GGGT") == ["TGGG", "GGGT"]
    assert oracle.debruijn([seq, "AAATCCCTGGAT"], 4, "backward", source="GGAT") == ["TGGA", "GGAT"]
    assert oracle.debruijn([seq, "AAATCGCTGGGT"], 4, "backward", source="GGGT") == ["CTGG", "TGGG", "GGGT"]


def test_merge_nodes():  # DeBruijnGraphSuite.scala:234-257
    assert oracle.debruijn(["AAATCCCTGGGT"], 4, "counts", merge_nodes=True) == {"AAATCCCTGGGT": 1}
    assert oracle.debruijn(["AAATCCCTGGGT", "AAATCCCTGGAT"], 4, "counts", merge_nodes=True) == {"AAATCCCTGG": 1, "TGGGT": 1, "TGGAT": 1}


def test_merge_index():  # DeBruijnGraphSuite.scala:259-282
    long_read = ("TGCATGGTGCTGTGAGATCAGCGTGTGTGTGTGTGCAGTGCATGGTGCTGTGTGAGATCAGCATGTGTGTGTGTGCAGTGCATGGTGCTGTGAGATCAGCGTGTGTGTGCAGCGCATGGTGCTGTGTGAGA")
    idx = oracle.debruijn([long_read], 45, "merge_index", merge_nodes=True, source=long_read)
    assert len(idx) == len(long_read) - 44
    for i, (kmer, where) in enumerate(idx):
        assert where == f"{long_read}:{i}"


def _reads(extra=()):
    r = REF52
    return [r[0:25], r[5:30], r[7:32], r[10:35], r[19:41]] + list(extra) + [r[22:44], r[25:47], r[31:52] + "TTT"]


@pytest.mark.parametrize("extra", [(), (REF52[19:41] + "TCGAA",), (REF52[19:41] + "TCGAA", REF52[19:41] + "TCGTA")])
@pytest.mark.parametrize("merge", [False, True])
def test_depth_first_search_single_path(extra, merge):  # DeBruijnGraphSuite.scala:284-425
    paths = oracle.debruijn(_reads(extra), 15, "dfs", merge_nodes=merge, source=REF52[:15], sink=REF52[-15:])
    assert paths == [REF52]


def test_depth_first_search_finds_both_alleles():
    alt = REF52[:26] + "T" + REF52[27:]
    reads = [REF52[i:i + 30] for i in range(0, 23, 2)] + [alt[i:i + 30] for i in range(0, 23, 2)]
    paths = oracle.debruijn(reads, 15, "dfs", merge_nodes=True, source=REF52[:15], sink=REF52[-15:], max_paths=17)
    assert sorted(paths) == sorted([REF52, alt])


def test_alignment_scores():  # ReadAlignmentSuite.scala:78-102
    assert oracle.read_alignment("TCGA", "TCGA", 1e-2, 1e-3, 1e-2)[3] == 0
    assert round(oracle.read_alignment("TCGA", "TCCA", 1e-2, 1e-3, 1e-2)[4]) == 5


@pytest.mark.parametrize("seq,ref,cigar", [
    ("TCGA", "TCGA", "4="), ("TCGA", "TCCA", "2=1X1="), ("TCGATGATCTGAGA", "TCGATGATCTGAGA", "14="),
    ("TCCGA", "TCGA", "1=1I3="), ("TCGACCCTCTGA", "TCGATCTGA", "4=3I5="), ("TCGATCTGA", "TCGACCCTCTGA", "4=3D5="),
    ("TCGACCCTCTTA", "TCGATCTGA", "4=3I3=1X1="),
    ("ATTCTCAAGTTTTAAGTGGTATTCTAATTATGGCAGTAATTAACTGAATAAAGAGATTCATCATGTGCAAAAACTAATCTTGTTTACTTAAAATTGAGAGT",
     "ATTCTCAAGTTTTAAGTGGTTTTCTAATTATGGCAGTAATAAACTGAATAAAGAGATTCATCATGTGCAAAAACTAATCTTGTTTACTTAAAATTGAGAGT", "20=1X19=1X60="),
    ("ATTCTCAAGTTTTAAGTGGTATTCTAATTATGGCAGTAATTAACTGAATAAAGAGATTCATCATGTGCAAAAACTAATCTT" + "GTTTACTTAAAATTGAGAGT",
     "ATTCTCAAGTTTTAAGTGGTTTTCTAATTATGGCAGTAATAAACTGAATAAAGAGATTCATCATGTGCAAAAACTAATCTTCCCGTTTACTTAAAATTGAGAGT", "20=1X19=1X40=3D20="),
    ("AGACACGGAGACACACAGAGATACACGGAAACACAG" + "ACATGCACACACGCGAAGACACAGACACATACACATGCAT",
     "AGACACGGAGACACACAGAGATACACGGAAACACAGAC" + "ACATGCACACACGCGAAGACACAGACACATACACATGCAT", "36=2D40="),
])
def test_alignment_cigars(seq, ref, cigar):  # ReadAlignmentSuite.scala:104-158
    assert oracle.read_alignment(seq, ref)[0] == cigar


def test_real_reads_single_path():  # DeBruijnGraphSuite.scala:427-457
    from guacamole_b200.io import read_sam
    path = "/root/reference/src/test/resources/assemble-reads-set3-chr2-73613071.sam"
    fixture = os.path.join(os.path.dirname(__file__), "golden", "assemble_reads_set3.npz")
    if os.path.exists(path) and not os.path.exists(fixture):
        reads, _ = read_sam(path, 0)
        reads.sort(key=lambda r: r.start)
        np.savez_compressed(fixture, seqs=np.array([r.sequence.decode() for r in reads]),
                            quals=np.array([bytes(r.base_qualities).hex() for r in reads]))
    z = np.load(fixture)
    seqs = [str(s) for s in z["seqs"]]
    reference = ("GAGGATCTGCCATGGCCGGGCGAGCTGGAGGAGGAGGAGGAGGAGGAGGAGGAGGAGGAGGAGGAAGAGGAGGAGGCTGCAGCGGCGGCGGCGGCGAACGTGGACGACGTAGTGGTCGTGGAGGAGGTGGAGGAAGAGGCGGG"
                 "GCG")
    for merge in (False, True):
        paths = oracle.debruijn(seqs, 55, "dfs", min_occurrence=3, merge_nodes=merge, source=reference[:55], sink=reference[-55:])
        assert len(paths) == 1 and paths[0].startswith(reference[:55]) and paths[0].endswith(reference[-55:])
