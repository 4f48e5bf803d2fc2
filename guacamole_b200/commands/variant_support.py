"""Host mirror of the reference's two small consumers of the pileup operator, over the C ABI.

Names follow the Scala: ``pileups_to_allele_counts`` is VariantSupport.Caller.pileupsToAlleleCounts through
pileupFlatMapMultipleSamples (commands/VariantSupport.scala:83-112); ``variant_loci_from_reads`` and
``generate_vaf_histograms`` are VAFHistogram's (commands/VAFHistogram.scala:194-245).  The per-element work runs on the
device (guac_variant_support, guac_vaf_histogram)."""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

from ..native import Context


def pileups_to_allele_counts(ctx: Context, loci: Sequence[Tuple[int, int, int]], sample_names: Sequence[str],
                             contig_names: Sequence[str]) -> List[str]:
    """AlleleCount.toString rows ("sample, contig, locus, ref, alt, count"), one per distinct allele of every sample's pileup
    at every visited locus (VariantSupport.scala:39-46,100-112)."""
    return [f"{sample_names[s]}, {contig_names[c]}, {x}, {ref}, {alt}, {n}" for s, c, x, ref, alt, n in ctx.variant_support(loci)]


def variant_loci_from_reads(ctx: Context, loci: Sequence[Tuple[int, int, int]], min_read_depth: int = 0,
                            min_variant_allele_frequency: int = 0):
    """(VariantLocus records, {sample id: number of variant loci}) (VAFHistogram.scala:214-245)."""
    _, totals, recs = ctx.vaf_histogram(loci, 20, min_read_depth, min_variant_allele_frequency)
    return recs, {s: int(n) for s, n in enumerate(totals)}


def generate_vaf_histograms(ctx: Context, loci: Sequence[Tuple[int, int, int]], bins: int, min_read_depth: int = 0,
                            min_variant_allele_frequency: int = 0) -> Dict[int, List[Tuple[int, int]]]:
    """{sample id: [(rounded variant allele percentage, number of loci)] sorted by bin} (VAFHistogram.scala:194-212)."""
    hist, _, _ = ctx.vaf_histogram(loci, bins, min_read_depth, min_variant_allele_frequency, want_loci=False)
    return {s: [(b, int(n)) for b, n in enumerate(row) if n] for s, row in enumerate(hist) if row.any()}
