"""Host side of the somatic-standard caller around the fused device path.

Mirrors `SomaticStandard.Caller.computeVariants`
(reference: src/main/scala/org/hammerlab/guacamole/commands/SomaticStandardCaller.scala:75-153):
reads of the normal (sample 0) and tumor (sample 1) are partitioned over the loci, the per-locus
closure `findPotentialVariantAtLocus` runs fused on the device (guac_somatic_standard), and the
`SomaticGenotypeFilter` chain plus VCF rendering stay on the host, as in the reference where they
are RDD filters and an htsjdk/ADAM writer.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import numpy as np

from ..reads import ReadBatch


def covered_ranges(batch: ReadBatch, loci: Optional[Sequence[Tuple[int, int, int]]] = None) -> List[Tuple[int, int, int]]:
    """Loci covered by at least one read, as sorted disjoint (contig_id, start, end) ranges, intersected
    with `loci` when given.  With skipEmpty = true the operator only ever visits covered loci
    (windowing/SlidingWindow.scala:127-152), so handing the device these ranges instead of whole
    contigs changes nothing but the amount of empty space scanned."""
    ends = batch.ends()
    out: List[Tuple[int, int, int]] = []
    for cid in np.unique(batch.contig_id):
        m = batch.contig_id == cid
        s, e = batch.start[m], ends[m]
        cur_s, cur_e = int(s[0]), int(e[0])
        for a, b in zip(s[1:], e[1:]):
            if a > cur_e:
                out.append((int(cid), cur_s, cur_e))
                cur_s, cur_e = int(a), int(b)
            else:
                cur_e = max(cur_e, int(b))
        out.append((int(cid), cur_s, cur_e))
    if loci is None:
        return out
    clipped = []
    for c, s, e in out:
        for lc, ls, le in loci:
            if lc == c and ls < e and s < le:
                clipped.append((c, max(s, ls), min(e, le)))
    return sorted(clipped)


def success_probability_to_phred(probability: float) -> int:
    """util/PhredUtils.scala:7-14 over ADAM's successProbabilityToPhred: round(-10 log10(1 - min(1 - 1e-16, p)))."""
    p = min(1 - 1e-16, probability)
    return int(math.floor(-10.0 * math.log10(1.0 - p) + 0.5))


@dataclass
class SomaticGenotypeFilterArguments:
    """filters/somatic/SomaticGenotypeFilter.scala:34-80 defaults."""

    min_likelihood: Optional[int] = None
    min_vaf: int = 0
    min_average_mapping_quality: int = 20
    min_average_base_quality: int = 20
    min_tumor_read_depth: int = 5
    min_normal_read_depth: int = 5
    max_tumor_read_depth: int = 2**31 - 1
    min_tumor_alternate_read_depth: Optional[int] = None
    maximum_median_mismatches: Optional[int] = None


def phred_scaled_somatic_likelihood(call) -> int:
    """variants/CalledSomaticAllele.scala:38-43."""
    return success_probability_to_phred(float(call["tumor_variant_evidence"]["probability"]) *
                                        float(call["normal_reference_evidence"]["probability"]))


def somatic_genotype_filter(calls: np.ndarray, args: SomaticGenotypeFilterArguments = SomaticGenotypeFilterArguments()) -> np.ndarray:
    """SomaticGenotypeFilter.apply (filters/somatic/SomaticGenotypeFilter.scala:82-130), including the
    reference's quirk that the "average base quality" filter tests meanMappingQuality
    (filters/somatic/SomaticAverageBaseQualityFilter.scala:11-12)."""
    t, n = calls["tumor_variant_evidence"], calls["normal_reference_evidence"]
    keep = (t["read_depth"] >= args.min_tumor_read_depth) & (t["read_depth"] < args.max_tumor_read_depth)
    keep &= n["read_depth"] >= args.min_normal_read_depth
    if args.min_tumor_alternate_read_depth is not None:
        keep &= t["allele_read_depth"] >= args.min_tumor_alternate_read_depth
    if args.min_likelihood is not None:
        keep &= np.array([phred_scaled_somatic_likelihood(c) >= args.min_likelihood for c in calls], bool)
    vaf = t["allele_read_depth"].astype(np.float32) / t["read_depth"].astype(np.float32)  # Float division (AlleleEvidence.scala:37)
    keep &= vaf.astype(np.float64) * 100.0 > args.min_vaf
    for thr in (args.min_average_mapping_quality, args.min_average_base_quality):
        keep &= (t["mean_mapping_quality"] >= thr) & (n["mean_mapping_quality"] >= thr)
    if args.maximum_median_mismatches is not None:
        keep &= t["median_mismatches_per_read"] <= args.maximum_median_mismatches
    return calls[keep]


def alleles_of(result) -> List[Tuple[bytes, bytes]]:
    a = result["arena"]
    return [(a[c["ref_off"]:c["ref_off"] + c["ref_len"]], a[c["alt_off"]:c["alt_off"] + c["alt_len"]]) for c in result["calls"]]


def vcf_rows(result, contig_names: Sequence[str], calls: Optional[np.ndarray] = None, sample_name: str = "tumor") -> List[str]:
    """Data rows as the reference's ADAM VCF writer renders a CalledSomaticAllele.toBDGGenotype
    (variants/CalledSomaticAllele.scala:45-64, variants/GenotypeOutputArgs.scala:86-113): sorted by
    (start, end); GT 0/1; AD ref,alt; DP; GQ = phredScaledSomaticLikelihood, saturated at 99 as in
    tough.golden.vcf (the only golden that pins it)."""
    calls = result["calls"] if calls is None else calls
    a = result["arena"]
    rows = []
    for c in calls:
        ref = a[c["ref_off"]:c["ref_off"] + c["ref_len"]].decode()
        alt = a[c["alt_off"]:c["alt_off"] + c["alt_len"]].decode()
        t = c["tumor_variant_evidence"]
        dp, ad = int(t["read_depth"]), int(t["allele_read_depth"])
        gq = min(99, phred_scaled_somatic_likelihood(c))
        rows.append((int(c["start"]), f"{contig_names[int(c['contig_id'])]}\t{int(c['start']) + 1}\t.\t{ref}\t{alt}\t.\t.\t.\t"
                                      f"GT:AD:DP:GQ\t0/1:{dp - ad},{ad}:{dp}:{gq}"))
    rows.sort(key=lambda r: r[0])
    return [f"#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\t{sample_name}"] + [r[1] for r in rows]
