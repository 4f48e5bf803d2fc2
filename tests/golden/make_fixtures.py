#!/usr/bin/env python
"""Generates tests/golden/*.npz from the reference checkout's own test resources.

Run in the build container (where /root/reference is mounted):
    python tests/golden/make_fixtures.py
The GPU box has no /root/reference, so tests read only the committed .npz files.  Nothing of the
reference's *source* is copied: the fixtures are decoded test DATA (reads, partial FASTA windows,
golden VCF rows) in this repository's columnar layout.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from guacamole_b200.io import Reference, read_bam, read_sam  # noqa: E402
from guacamole_b200.reads import ReadBatch, filter_reads  # noqa: E402

RES = "/root/reference/src/test/resources/"
OUT = os.path.dirname(os.path.abspath(__file__))

BATCH_FIELDS = ["contig_id", "start", "cigar_off", "cigar", "seq_off", "bases_ascii", "qual", "mapq", "flags", "sample_id"]


def pack(batch: ReadBatch, reference: Reference = None, **extra):
    d = {f: getattr(batch, f) for f in BATCH_FIELDS}
    d["n_samples"] = np.int64(batch.n_samples)
    d["contig_names"] = np.array(batch.contig_names)
    if reference is not None:
        names = [n for n in batch.contig_names if n in reference.contigs]
        d["ref_names"] = np.array(names)
        for i, n in enumerate(names):
            w = reference.contigs[n]
            # store only the present runs (partial FASTA windows), not the zero gaps
            present = w.bases != 0
            edges = np.flatnonzero(np.diff(np.concatenate([[0], present.view(np.int8), [0]])))
            d[f"ref_{i}_starts"] = (edges[0::2] + w.offset).astype(np.int64)
            d[f"ref_{i}_lens"] = (edges[1::2] - edges[0::2]).astype(np.int64)
            d[f"ref_{i}_bases"] = w.bases[present]
            d[f"ref_{i}_length"] = np.int64(w.contig_length)
    for k, v in extra.items():
        d[k] = np.array(v)
    return d


def main():
    # 1. somatic-standard end to end (SomaticStandardEndToEndSuite.scala:13-37)
    normal, contigs = read_sam(RES + "normal.chr20.tough.sam", 0)
    tumor, _ = read_sam(RES + "tumor.chr20.tough.sam", 1)
    reads = filter_reads(normal + tumor, min_alignment_quality=1)  # --min-alignment-quality 1, default filters
    names = [c for c, _ in contigs]
    batch = ReadBatch.from_reads(reads, names, 2)
    ref = Reference.from_partial_fasta(RES + "grch37.partial.fasta")
    golden = [l.rstrip("\n") for l in open(RES + "tough.golden.vcf") if not l.startswith("##")]
    np.savez_compressed(os.path.join(OUT, "tough.npz"), **pack(batch, ref, golden_vcf=golden))

    # 2. PileupSuite SAM files
    for name in ["same_start_reads", "different_start_reads", "testrna"]:
        rs, cs = read_sam(RES + name + ".sam", 0)
        b = ReadBatch.from_reads(rs, [c for c, _ in cs] or sorted({r.contig for r in rs}), 1)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **pack(b))

    # 3. somatic-joint end to end inputs (SomaticJointCallerEndToEndSuite.scala:20-38)
    all_reads, contigs = [], None
    for i, n in enumerate(["normal", "primary", "recurrence"]):
        rs, cs = read_bam(RES + f"cancer-wgs1/{n}.tiny.bam", i)
        contigs = contigs or cs
        all_reads += rs
    all_reads = filter_reads(all_reads, non_duplicate=False, passed_vendor_quality_checks=False)  # --include-duplicates --include-failed-quality-checks
    names = [c for c, _ in contigs]
    batch = ReadBatch.from_reads(all_reads, names, 3)
    ref = Reference.from_partial_fasta(RES + "hg19.partial.fasta")
    golden = [l.rstrip("\n") for l in open(RES + "tiny-sjc-output/all.vcf") if not l.startswith("##")]
    force = [l.rstrip("\n") for l in open(RES + "tiny.vcf") if not l.startswith("#")]
    np.savez_compressed(os.path.join(OUT, "tiny_sjc.npz"), **pack(batch, ref, golden_vcf=golden, force_call_vcf=force))
    # 4. SomaticStandardCallerRealDataSuite.scala:79-150: tumor / normal SAM pairs and the positions they are probed at
    #    (loadTumorNormalReads: nonDuplicate, passedVendorQualityChecks; only reads over a probed position are kept)
    real = {
        "syn1fp": ("synthetic.challenge.set1.tumor.v2.withMDTags.chr2.syn1fp.sam", "synthetic.challenge.set1.normal.v2.withMDTags.chr2.syn1fp.sam",
                   [216094721, 3529313, 8789794, 104043280, 104175801, 126651101, 241901237, 57270796, 120757852]),
        "complexvar": ("synthetic.challenge.set1.tumor.v2.withMDTags.chr2.complexvar.sam",
                       "synthetic.challenge.set1.normal.v2.withMDTags.chr2.complexvar.sam",
                       [134307261, 90376213, 3638733, 109347468, 82949713, 130919744]),
        "simplefp": ("tumor.chr20.simplefp.sam", "normal.chr20.simplefp.sam", [13046318, 25939088, 26211835, 29652479, 54495768]),
    }
    ref = Reference.from_partial_fasta(RES + "grch37.partial.fasta")
    for name, (tumor_sam, normal_sam, positions) in real.items():
        normal, contigs = read_sam(RES + normal_sam, 0)
        tumor, _ = read_sam(RES + tumor_sam, 1)
        reads = [r for r in filter_reads(normal + tumor) if any(r.start <= p < r.end for p in positions)]
        batch = ReadBatch.from_reads(reads, [c for c, _ in contigs], 2)
        np.savez_compressed(os.path.join(OUT, f"real_{name}.npz"), **pack(batch, ref, positions=np.array(positions, np.int64)))
    # 5. VariantSupportSuite.scala:58-122: gatk_mini_bundle_extract.bam, every read over the loci the suite probes
    #    (no duplicate filter here: the suite applies it in one of its tests, and so do ours, from the flags)
    rs, cs = read_bam(RES + "gatk_mini_bundle_extract.bam", 0)
    sel = [r for r in rs if r.contig == "20" and r.start < 10009000 and r.end > 9999900]
    batch = ReadBatch.from_reads(sel, [c for c, _ in cs], 1)
    np.savez_compressed(os.path.join(OUT, "gatk_mini.npz"), **pack(batch, ref))
    for f in sorted(os.listdir(OUT)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
