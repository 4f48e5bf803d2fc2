"""GPU parity for the germline-assembly caller (De Bruijn graph + DFS + affine-gap alignment per active window),
through the C ABI, against the oracle's restatement of GermlineAssemblyCaller.discoverGermlineVariants.

The reference's own end-to-end suite (GermlineAssemblyCallerSuite) needs a BAM and a FASTA that are not in the
checkout, so the closure is compared on seeded synthetic reads with planted SNVs and indels; the graph and
alignment pieces are pinned to the reference's unit vectors in tests/test_oracle_assembly.py.
"""
import numpy as np
import pytest

from guacamole_b200 import synth
from guacamole_b200.native import ga_params
from oracle import oracle

pytestmark = pytest.mark.gpu

# GermlineAssemblyCallerSuite.verifyVariantsAtLocus parameters (kmerSize 47, assemblyWindowRange 120, minOccurrence 5,
# minMeanKmerQuality 20, minAreaVaf 0.1)
SUITE = dict(kmer_size=47, assembly_window_range=120, min_occurrence=5, min_mean_kmer_quality=20, min_area_vaf_percent=10)


def decode(result):
    arena = result["arena"]
    rows = []
    for c in result["calls"]:
        ref = arena[c["ref_off"]:c["ref_off"] + c["ref_len"]].decode()
        alt = arena[c["alt_off"]:c["alt_off"] + c["alt_len"]].decode()
        rows.append((int(c["n_genotypes"]) == 0, int(c["contig_id"]), int(c["start"]), ref, alt, c))
    return rows


def assert_assembly_equal(got, want):
    want_windows = [(w[1], w[2], int(w[3])) for w in want if w[0] == "W"]
    got_windows = [(int(w["contig_id"]), int(w["locus"]), int(w["n_paths"])) for w in got["windows"] if w["consumed"]]
    assert got_windows == want_windows
    rows = decode(got)
    want_a = [w for w in want if w[0] == "A"]
    got_a = [r for r in rows if r[0]]
    assert [(r[1], r[2], r[3], r[4]) for r in got_a] == [(w[1], w[2], w[3], w[4]) for w in want_a]
    for r, w in zip(got_a, want_a):
        ev = r[5]["evidence"]
        assert ev["probability"] == 1.0 and ev["median_mismatches_per_read"] == 0.0
        assert ev["read_depth"] == ev["allele_read_depth"] == ev["forward_depth"] == ev["allele_forward_depth"] == int(w[5])
        # breeze's Float running means depend on the order of the window queue (a heap in the reference, batch order
        # here): equal to Float rounding; the medians are order-free and exact
        np.testing.assert_allclose(ev["mean_mapping_quality"], float(w[6]), rtol=2e-5)
        assert ev["median_mapping_quality"] == float(w[7])
        np.testing.assert_allclose(ev["mean_base_quality"], float(w[8]), rtol=2e-5)
        assert ev["median_base_quality"] == float(w[9])
    want_p = [w for w in want if w[0] == "P"]
    got_p = [r for r in rows if not r[0]]
    assert [(r[1], r[2], r[3], r[4]) for r in got_p] == [(w[1], w[2], w[3], w[4]) for w in want_p]
    for r, w in zip(got_p, want_p):
        np.testing.assert_allclose(r[5]["evidence"]["probability"], float(w[5]), rtol=1e-9)
        assert (r[5]["evidence"]["read_depth"], r[5]["evidence"]["allele_read_depth"]) == (int(w[6]), int(w[7]))


@pytest.mark.parametrize("shortcut", [False, True])
@pytest.mark.parametrize("seed,coverage", [(2000, 30), (2001, 60)])
def test_synthetic_single_sample(gpu_ctx, seed, coverage, shortcut):
    contig_len = 20000
    ref = synth.make_reference(contig_len, seed)
    batch = synth.generate_reads(contig_len, (coverage,), (False,), seed)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    p = ga_params(shortcut_assembly=shortcut, **SUITE)
    ranges = [(0, 200, contig_len - 200)]
    want = oracle.germline_assembly(batch, ref, ranges, p)
    assert sum(1 for w in want if w[0] == "A") >= 5 and sum(1 for w in want if w[0] == "W") >= 20
    got = gpu_ctx.germline_assembly(ranges, p)
    assert_assembly_equal(got, want)
    # speculative windows: everything the reference's walk consumed was assembled, plus the ones its jumps skip
    assert len(got["windows"]) >= sum(1 for w in want if w[0] == "W")


def test_small_kmer_many_paths(gpu_ctx):
    """A short k-mer and minOccurrence 1 keep sequencing errors in the graph: several paths per window, so the
    read-vs-path scoring of AssemblyUtils.scala:92-104 and the too-many-paths cut-off are exercised."""
    contig_len, seed = 6000, 2100
    ref = synth.make_reference(contig_len, seed)
    batch = synth.generate_reads(contig_len, (30,), (False,), seed)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    p = ga_params(kmer_size=21, assembly_window_range=60, min_occurrence=1, min_mean_kmer_quality=0, min_area_vaf_percent=10)
    ranges = [(0, 100, contig_len - 100)]
    want = oracle.germline_assembly(batch, ref, ranges, p)
    n_paths = [int(w[3]) for w in want if w[0] == "W"]
    assert any(n == 2 for n in n_paths)
    got = gpu_ctx.germline_assembly(ranges, p)
    assert_assembly_equal(got, want)


def test_rejects_backward_jump(gpu_ctx):
    contig_len, seed = 2000, 2000
    ref = synth.make_reference(contig_len, seed)
    batch = synth.generate_reads(contig_len, (30,), (False,), seed)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    with pytest.raises(Exception, match="can only move forward"):
        gpu_ctx.germline_assembly([(0, 200, 1800)], ga_params())  # the reference's defaults: range 20 < k-mer 45


# ---- the reference's unit vectors through the device kernels ---------------------------------------------------------
from guacamole_b200.reads import ReadBatch, make_read  # noqa: E402
from tests import test_oracle_assembly as vectors  # noqa: E402


def upload_sequences(ctx, sequences, qualities=None):
    reads = [make_read(s, f"{len(s)}M", start=i, quality_scores=None if qualities is None else qualities[i])
             for i, s in enumerate(sequences)]
    ctx.upload_reads(ReadBatch.from_reads(reads))


@pytest.mark.parametrize("extra", [(), (vectors.REF52[19:41] + "TCGAA",), (vectors.REF52[19:41] + "TCGAA", vectors.REF52[19:41] + "TCGTA")])
def test_depth_first_search_vectors(gpu_ctx, extra):  # DeBruijnGraphSuite.scala:284-425 (merged graph)
    ref = vectors.REF52
    upload_sequences(gpu_ctx, vectors._reads(extra))
    assert gpu_ctx.discover_paths(ref, 15) == [ref]
    # only the first and last k-mer of the reference matter
    assert gpu_ctx.discover_paths(ref[:15] + "A" * 7 + ref[-15:], 15) == [ref]


def test_two_alleles_and_counts(gpu_ctx):
    ref = vectors.REF52
    alt = ref[:26] + "T" + ref[27:]
    reads = [ref[i:i + 30] for i in range(0, 23, 2)] + [alt[i:i + 30] for i in range(0, 23, 2)]
    upload_sequences(gpu_ctx, reads)
    want = oracle.debruijn(reads, 15, "dfs", merge_nodes=True, source=ref[:15], sink=ref[-15:], max_paths=17)
    assert gpu_ctx.discover_paths(ref, 15, max_paths=17) == want and sorted(want) == sorted([ref, alt])
    # minOccurrence prunes the alternate allele's k-mers when only two reads carry it
    few = reads[:12] + reads[12:14]
    upload_sequences(gpu_ctx, few)
    assert gpu_ctx.discover_paths(ref, 15, min_occurrence=3) == oracle.debruijn(few, 15, "dfs", min_occurrence=3, merge_nodes=True,
                                                                             source=ref[:15], sink=ref[-15:])
    # maxPaths stops the search
    upload_sequences(gpu_ctx, reads)
    assert gpu_ctx.discover_paths(ref, 15, max_paths=1) == oracle.debruijn(reads, 15, "dfs", merge_nodes=True, source=ref[:15],
                                                                         sink=ref[-15:], max_paths=1)


def test_kmer_quality_gate_vectors(gpu_ctx):  # DeBruijnGraphSuite.scala:62-104: low-quality k-mers do not reach minOccurrence
    seq = "TCATCTTAAAAGACATAAACCGGTTGCA"
    mixed = [30] * 9 + [5] * 6 + [30] * (len(seq) - 15)
    for threshold in (0, 15, 25):
        quals = [[30] * len(seq), mixed]
        upload_sequences(gpu_ctx, [seq, seq], quals)
        want = oracle.debruijn([seq, seq], 5, "dfs", min_occurrence=2, min_mean_kmer_quality=threshold, merge_nodes=True, qualities=quals,
                               source=seq[:5], sink=seq[-5:])
        assert gpu_ctx.discover_paths(seq, 5, min_occurrence=2, min_mean_kmer_quality=threshold) == want


def test_real_reads_vector(gpu_ctx):  # DeBruijnGraphSuite.scala:427-457
    z = np.load(vectors.os.path.join(vectors.os.path.dirname(vectors.__file__), "golden", "assemble_reads_set3.npz"))
    seqs = [str(s) for s in z["seqs"]]
    reference = ("GAGGATCTGCCATGGCCGGGCGAGCTGGAGGAGGAGGAGGAGGAGGAGGAGGAGGAGGAGGAGGAAGAGGAGGAGGCTGCAGCGGCGGCGGCGGCGAACGTGGACGACGTAGTGGTCGTGGAGGAGGTGGAGGAAGAGGCGGG"
                 "GCG")
    upload_sequences(gpu_ctx, seqs)
    got = gpu_ctx.discover_paths(reference, 55, min_occurrence=3)
    assert len(got) == 1
    assert got == oracle.debruijn(seqs, 55, "dfs", min_occurrence=3, merge_nodes=True, source=reference[:55], sink=reference[-55:])


def test_alignment_vectors(gpu_ctx):  # ReadAlignmentSuite.scala:78-158
    assert gpu_ctx.read_alignment("TCGA", "TCGA", 1e-2, 1e-3, 1e-2)[3] == 0
    assert round(gpu_ctx.read_alignment("TCGA", "TCCA", 1e-2, 1e-3, 1e-2)[4]) == 5
    for seq, ref, cigar in vectors.test_alignment_cigars.pytestmark[0].args[1]:  # the suite's (sequence, reference, cigar) table
        got = gpu_ctx.read_alignment(seq, ref)
        want = oracle.read_alignment(seq, ref)
        assert got[0] == cigar
        assert got[:4] == tuple(want[:4]) and got[4] == want[4]  # the raw FP64 score bit for bit


def test_alignment_random_pairs(gpu_ctx):
    rng = np.random.default_rng(7)
    for _ in range(40):
        n = int(rng.integers(5, 120))
        ref = "".join(rng.choice(list("ACGT"), n + int(rng.integers(0, 40))))
        at = int(rng.integers(0, len(ref) - n + 1))
        seq = list(ref[at:at + n])
        for _ in range(int(rng.integers(0, 4))):  # a few edits
            i = int(rng.integers(0, len(seq)))
            kind = int(rng.integers(0, 3))
            if kind == 0:
                seq[i] = "ACGT"[int(rng.integers(0, 4))]
            elif kind == 1:
                seq[i:i] = list("ACGT"[int(rng.integers(0, 4))] * int(rng.integers(1, 4)))
            else:
                del seq[i:i + int(rng.integers(1, 4))]
        seq = "".join(seq)
        if not seq:
            continue
        got, want = gpu_ctx.read_alignment(seq, ref), oracle.read_alignment(seq, ref)
        assert got[:4] == tuple(want[:4]) and got[4] == want[4]


# ---- planted indels (BASELINE configs[4]: germline-assembly on reads with planted indels, k = 31) -----------------------
def simulate_haplotype_reads(reference: str, edits, n_reads, read_len, rng, alt_fraction=1.0, first=0):
    """Reads of `read_len` bases drawn from the reference or (with probability alt_fraction) from the haplotype that
    carries `edits` = [(ref_pos, deleted_len, inserted_bases)], aligned with the CIGAR the edits imply."""
    reads = []
    span = len(reference) - 2 * read_len
    for i in range(n_reads):
        start = first + int(rng.integers(0, span))
        use_alt = rng.random() < alt_fraction
        seq, ops, pos = [], [], start
        todo = [e for e in edits if use_alt and e[0] > start]  # an edit right at the read's first base is not representable

        def push(op, n):
            if n <= 0:
                return
            if ops and ops[-1][0] == op:
                ops[-1][1] += n
            else:
                ops.append([op, n])

        while len(seq) < read_len:
            nxt = todo[0] if todo else None
            if nxt is not None and nxt[0] == pos:
                _, dlen, ins = todo.pop(0)
                if len(seq) + len(ins) >= read_len:  # the read would end inside the insertion: stop before it
                    break
                if ins:
                    seq.extend(ins)
                    push("I", len(ins))
                if dlen:
                    push("D", dlen)
                    pos += dlen
                continue
            stop = min(nxt[0] if nxt is not None else len(reference), pos + read_len - len(seq))
            seq.extend(reference[pos:stop])
            push("M", stop - pos)
            pos = stop
        while ops and ops[-1][0] != "M":  # no trailing indel operators
            op, n = ops.pop()
            if op == "I":
                del seq[-n:]
        cigar = "".join(f"{n}{op}" for op, n in ops)
        reads.append(make_read("".join(seq), cigar, start=start, chr="chr1", quality_scores=[30] * len(seq),
                               is_positive_strand=bool(i & 1)))
    return reads


@pytest.mark.parametrize("shortcut", [False, True])
def test_planted_indels_k31(gpu_ctx, shortcut):
    from guacamole_b200.io import make_reference
    rng = np.random.default_rng(31)
    n = 6000
    reference = "".join(rng.choice(list("ACGT"), n))
    # a 3-base deletion, a 2-base insertion, a 1-base deletion next to an SNV-free stretch; anchors are the base before
    edits = [(1500, 3, ""), (3000, 0, "GT"), (4500, 1, "")]
    reads = simulate_haplotype_reads(reference, edits, n_reads=1800, read_len=100, rng=rng, alt_fraction=1.0)
    reads += simulate_haplotype_reads(reference, [(2200, 0, "ACA")], n_reads=900, read_len=100, rng=rng, alt_fraction=0.5)  # het insertion
    batch = ReadBatch.from_reads(reads, ["chr1"], 1)
    ref = make_reference(("chr1", 0, reference), contig_length=n)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    p = ga_params(kmer_size=31, assembly_window_range=100, min_occurrence=3, min_mean_kmer_quality=0, min_area_vaf_percent=10,
                  shortcut_assembly=shortcut)
    ranges = [(0, 150, n - 150)]
    want = oracle.germline_assembly(batch, ref, ranges, p)
    got = gpu_ctx.germline_assembly(ranges, p)
    assert_assembly_equal(got, want)
    # every planted indel is called: same length change, anchored at the base before it (an insertion inside a short
    # repeat is reported at its left-most equivalent position, a base or two earlier)
    called = [(r[2], len(r[3]) - len(r[4])) for r in decode(got) if r[0]]
    for pos, change in ((1499, 3), (2199, -3), (2999, -2), (4499, 1)):
        assert any(c == change and pos - 3 <= q <= pos for q, c in called), (pos, change, called)
