"""Host side of the somatic-joint caller: parameters, record decoding and VCF rendering.

Mirrors `SomaticJoint.Caller` (reference: src/main/scala/org/hammerlab/guacamole/commands/SomaticJointCaller.scala:58-203),
`Parameters` (jointcaller/Parameters.scala:10-25), `InputCollection` defaults (jointcaller/InputCollection.scala:55-101:
first input normal, the rest tumor, all DNA) and `VCFOutput.makeHtsjdVariantContext` (jointcaller/VCFOutput.scala:127-413).
The per-locus model runs on the device (guac_somatic_joint); VCF writing stays on the host as in the
reference (htsjdk there, plain text rendering here).
"""
from __future__ import annotations

import ctypes as C
import math
from decimal import ROUND_HALF_UP, Decimal
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

NORMAL_DNA, TUMOR_DNA, TUMOR_RNA = 0, 1, 2
MAX_SAMPLES, MAX_DEPTH_ENTRIES = 16, 8


class SJParamsC(C.Structure):
    _fields_ = [("n_samples", C.c_int32), ("sample_kind", C.c_int32 * MAX_SAMPLES), ("max_alleles_per_site", C.c_int32),
                ("any_allele_min_supporting_reads", C.c_int32), ("any_allele_min_supporting_percent", C.c_double),
                ("germline_neg_log10_het_prior", C.c_double), ("germline_neg_log10_hom_alt_prior", C.c_double),
                ("somatic_neg_log10_variant_prior", C.c_double), ("somatic_neg_log10_variant_prior_rna", C.c_double),
                ("somatic_neg_log10_variant_prior_with_rna_evidence", C.c_double), ("somatic_vaf_floor", C.c_double),
                ("filter_strand_bias_phred", C.c_double), ("filter_somatic_normal_nonreference_percent", C.c_double),
                ("filter_somatic_normal_depth", C.c_int32), ("only_somatic", C.c_int32), ("include_filtered", C.c_int32),
                ("_pad", C.c_int32)]


_DEPTH = np.dtype([("seq_off", np.uint32), ("seq_len", np.uint32), ("count", np.int32), ("_pad", np.int32)])
SAMPLE_EVIDENCE_DTYPE = np.dtype([
    ("kind", np.int32), ("total_depth", np.int32), ("n_depths", np.int32), ("n_likelihoods", np.int32),
    ("depths", _DEPTH, (MAX_DEPTH_ENTRIES,)), ("log_likelihoods", np.float64, (3,)), ("posteriors", np.float64, (3,)),
    ("alt_vaf", np.float64), ("sb_variant_forward", np.int32), ("sb_variant_reverse", np.int32), ("sb_total_forward", np.int32),
    ("sb_total_reverse", np.int32), ("sb_phred", np.float64), ("sb_filtered", np.int32), ("triggered", np.int32)])
ALLELE_EVIDENCE_DTYPE = np.dtype([
    ("contig_id", np.int32), ("site_index", np.int32), ("start", np.int64), ("ref_off", np.uint32), ("ref_len", np.uint32),
    ("alt_off", np.uint32), ("alt_len", np.uint32), ("is_germline_call", np.int32), ("is_somatic_call", np.int32),
    ("germline_genotype", np.int32), ("tumor_rna_expressed", np.int32), ("has_insufficient_normal", np.int32),
    ("in_reference_reads", np.int32), ("in_total_reads", np.int32), ("in_filtered", np.int32), ("failing_filters", np.uint32),
    ("forced", np.int32), ("first_sample_evidence", np.int64)])


def sj_params(sample_kinds: Sequence[int], **kw) -> SJParamsC:
    """Parameters.defaults (jointcaller/Parameters.scala:47-109) for the given per-sample kinds."""
    p = SJParamsC()
    p.n_samples = len(sample_kinds)
    for i, k in enumerate(sample_kinds):
        p.sample_kind[i] = k
    p.max_alleles_per_site = 4
    p.any_allele_min_supporting_reads = 2
    p.any_allele_min_supporting_percent = 2.0
    p.germline_neg_log10_het_prior = 4
    p.germline_neg_log10_hom_alt_prior = 5
    p.somatic_neg_log10_variant_prior = 6
    p.somatic_neg_log10_variant_prior_rna = 1
    p.somatic_neg_log10_variant_prior_with_rna_evidence = 1
    p.somatic_vaf_floor = 0.05
    p.filter_strand_bias_phred = 60
    p.filter_somatic_normal_nonreference_percent = 3.0
    p.filter_somatic_normal_depth = 30
    for k, v in kw.items():
        if not hasattr(p, k):
            raise AttributeError(k)
        setattr(p, k, v)
    return p


def default_sample_kinds(n: int) -> List[int]:
    """InputCollection defaults: first sample normal DNA, the others tumor DNA (InputCollection.scala:73-81)."""
    return [NORMAL_DNA] + [TUMOR_DNA] * (n - 1)


def decode(alleles: np.ndarray, samples: np.ndarray, arena: bytes, n_samples: int) -> List[dict]:
    """Flat records -> one dict per MultiSampleSingleAlleleEvidence."""
    def s(off, ln):
        return arena[int(off):int(off) + int(ln)].decode()
    out = []
    for a in alleles:
        ev = dict(contig_id=int(a["contig_id"]), site_index=int(a["site_index"]), start=int(a["start"]),
                  ref=s(a["ref_off"], a["ref_len"]), alt=s(a["alt_off"], a["alt_len"]),
                  is_germline_call=bool(a["is_germline_call"]), is_somatic_call=bool(a["is_somatic_call"]),
                  germline_genotype=int(a["germline_genotype"]), tumor_rna_expressed=bool(a["tumor_rna_expressed"]),
                  insufficient_normal=(dict(reference_reads=int(a["in_reference_reads"]), total_reads=int(a["in_total_reads"]),
                                            filtered=bool(a["in_filtered"])) if a["has_insufficient_normal"] else None),
                  failing_filters=int(a["failing_filters"]), forced=bool(a["forced"]), samples=[])
        f = int(a["first_sample_evidence"])
        for se in samples[f:f + n_samples + 2]:
            nl = int(se["n_likelihoods"])
            ev["samples"].append(dict(
                kind=int(se["kind"]), total_depth=int(se["total_depth"]),
                allelic_depths=[(s(d["seq_off"], d["seq_len"]), int(d["count"])) for d in se["depths"][:int(se["n_depths"])]],
                log_likelihoods=[float(x) for x in se["log_likelihoods"][:nl]], posteriors=[float(x) for x in se["posteriors"][:nl]],
                alt_vaf=float(se["alt_vaf"]),
                strand_bias=(int(se["sb_variant_forward"]), int(se["sb_variant_reverse"]), int(se["sb_total_forward"]),
                             int(se["sb_total_reverse"])), sb_phred=float(se["sb_phred"]), sb_filtered=bool(se["sb_filtered"]),
                triggered=bool(se["triggered"])))
        out.append(ev)
    return out


# ---- VCF rendering (jointcaller/VCFOutput.scala:127-413 through htsjdk 2.6.1) -----------------------------

def java_g8(x: float) -> str:
    """Java's String.format("%.8g", x): 8 significant digits, HALF_UP, fixed notation for 1e-4 <= |x| < 1e8,
    trailing zeros kept."""
    if math.isnan(x):
        return "NaN"
    if math.isinf(x):
        return "Infinity" if x > 0 else "-Infinity"
    if x == 0:
        return "0.0000000"
    d = Decimal(x)  # the exact value of the double, as the JVM's Formatter takes it
    exp = d.adjusted()
    rounded = d.quantize(Decimal(1).scaleb(exp - 7), rounding=ROUND_HALF_UP)
    exp = rounded.adjusted()
    if -4 <= exp < 8:
        return f"{rounded:.{max(0, 7 - exp)}f}"
    mant = rounded.scaleb(-exp)
    return f"{mant:.7f}e{'+' if exp >= 0 else '-'}{abs(exp):02d}"


def java_fixed(x: float, frac: int) -> str:
    """Java's String.format("%.<frac>f", x): HALF_UP on the exact value of the double (Python's own formatting rounds
    half-even: 0.625 -> "0.62", the JVM prints "0.63")."""
    if math.isnan(x):
        return "NaN"
    if math.isinf(x):
        return "Infinity" if x > 0 else "-Infinity"
    q = Decimal(x).quantize(Decimal(1).scaleb(-frac), rounding=ROUND_HALF_UP)
    s = f"{q:.{frac}f}"
    return s[1:] if s.startswith("-") and float(q) == 0 else s


def format_vcf_double(d: float) -> str:
    """htsjdk VCFEncoder.formatVCFDouble."""
    if d < 1:
        if d < 0.01:
            if abs(d) >= 1e-20:
                dd = Decimal(d)
                exp = dd.adjusted()
                r = dd.quantize(Decimal(1).scaleb(exp - 3), rounding=ROUND_HALF_UP)
                exp = r.adjusted()
                return f"{r.scaleb(-exp):.3f}e{'+' if exp >= 0 else '-'}{abs(exp):02d}"
            return "0.00"
        return java_fixed(d, 3)
    return java_fixed(d, 2)


def _allele_str(x: str) -> str:
    return "." if x == "" else x


def _depth(ev_sample, allele):
    return dict(ev_sample["allelic_depths"]).get(allele, 0)


def _sample_depth(ev_sample):  # NormalDNA...Evidence.depth = allelicDepths.values.sum (of the truncated map)
    return sum(c for _, c in ev_sample["allelic_depths"])


def _sample_vaf(ev_sample, alt):
    d = _sample_depth(ev_sample)
    return _depth(ev_sample, alt) / d if d else float("nan")


def vcf_row(ev: dict, contig_names: Sequence[str], sample_names: Sequence[str], sample_kinds: Sequence[int],
            include_pooled_normal: Optional[bool] = None, include_pooled_tumor: Optional[bool] = None,
            somatic_genotype_policy: str = "presence") -> str:
    """One data row of VCFOutput.writeVcf for one MultiSampleSingleAlleleEvidence."""
    S = len(sample_names)
    ref, alt = ev["ref"], ev["alt"]
    if include_pooled_normal is None:
        include_pooled_normal = sum(k == NORMAL_DNA for k in sample_kinds) > 1  # SomaticJointCaller.scala:223
    if include_pooled_tumor is None:
        include_pooled_tumor = sum(k == TUMOR_DNA for k in sample_kinds) > 1
    columns = [(i, sample_names[i], sample_kinds[i]) for i in range(S)]
    if include_pooled_normal:
        columns.append((S, "pooled_normal", NORMAL_DNA))
    if include_pooled_tumor:
        columns.append((S + 1, "pooled_tumor", TUMOR_DNA))

    def gt_index(a):
        return "0" if a == ref else "1"

    def adp(se):
        alleles = list(dict.fromkeys([ref, alt] + [a for a, _ in se["allelic_depths"]]))
        return ",".join(f"{_allele_str(a)}={_depth(se, a)}" for a in alleles)

    any_ff = any(ev["samples"][i]["sb_filtered"] for i, _, _ in columns)
    keys = ["GT", "AD", "ADP", "DP"] + (["FF"] if any_ff else []) + ["FS", "RL", "TRIGGERED", "VAF"]
    cells = []
    for idx, _, kind in columns:
        se = ev["samples"][idx]
        vals: Dict[str, str] = {}
        if kind == NORMAL_DNA:
            names = [(ref, ref), (ref, alt), (alt, alt)]
            best = 0  # posteriors.maxBy(_._2): the first strictly greater entry wins
            for g in (1, 2):
                if se["posteriors"][g] > se["posteriors"][best]:
                    best = g
            a1, a2 = names[best]
            vals["RL"] = ",".join(f"[{x}/{y}]={java_g8(p)}" for (x, y), p in zip(names, se["posteriors"]))
            vals["TRIGGERED"] = "."
        else:
            mixtures = [[(ref, 1.0)]] + ([[(alt, se["alt_vaf"]), (ref, 1.0 - se["alt_vaf"])]] if ref != alt else [])
            if somatic_genotype_policy == "presence":
                present = sorted([(a, c) for a, c in se["allelic_depths"] if c > 0], key=lambda t: -t[1])[:2]
                top = {a for a, _ in present}
                if top == {ref, alt}:
                    a1, a2 = ref, alt
                elif top == {ref} or top == {alt}:
                    a1 = a2 = next(iter(top))
                else:
                    a1 = a2 = ref
            else:
                a1, a2 = (ref, alt) if se["triggered"] else (ref, ref)
            vals["RL"] = ",".join("[" + "|".join(f"{_allele_str(a)}->{java_fixed(f, 2)}" for a, f in m) + f"]={java_g8(p)}"
                                  for m, p in zip(mixtures, se["posteriors"]))
            vals["TRIGGERED"] = ("EXPRESSED" if kind == TUMOR_RNA else "YES") if se["triggered"] else "NO"
        vals["GT"] = "/".join(sorted([gt_index(a1), gt_index(a2)]))
        vals["AD"] = f"{_depth(se, ref)},{_depth(se, alt)}"
        vals["ADP"] = adp(se)
        vals["DP"] = str(_sample_depth(se))
        vals["FS"] = str(int(math.floor(se["sb_phred"] + 0.5)))
        vals["VAF"] = format_vcf_double(_sample_vaf(se, alt))
        if any_ff:
            vals["FF"] = "STRAND_BIAS" if se["sb_filtered"] else "."
        cells.append(":".join(vals[k] for k in keys))
    triggers = []
    if ev["is_germline_call"]:
        triggers.append("GERMLINE_POOLED")
    else:
        tumor_idx = [i for i in range(S) if sample_kinds[i] == TUMOR_DNA]
        if ev["samples"][S + 1]["triggered"]:
            triggers.append("SOMATIC_POOLED")
        if any(ev["samples"][i]["triggered"] for i in tumor_idx):
            triggers.append("SOMATIC_INDIVIDUAL")
    names = [n for bit, n in ((2, "INSUFFICIENT_NORMAL"), (1, "STRAND_BIAS")) if ev["failing_filters"] & bit]
    if names:
        flt = ";".join(sorted(names))
    else:
        flt = "PASS" if (ev["is_germline_call"] or ev["is_somatic_call"]) else "."
    info = f"TRIGGER={','.join(triggers) if triggers else 'NONE'};TUMOR_EXPRESSION={'YES' if ev['tumor_rna_expressed'] else 'NO'}"
    return "\t".join([contig_names[ev["contig_id"]], str(ev["start"] + 1), ".", ref, alt, ".", flt, info, ":".join(keys)] + cells)


def vcf_rows(evidences: List[dict], contig_names, sample_names, sample_kinds, **kw) -> List[str]:
    S = len(sample_names)
    ipn = kw.get("include_pooled_normal")
    ipt = kw.get("include_pooled_tumor")
    if ipn is None:
        ipn = sum(k == NORMAL_DNA for k in sample_kinds) > 1
    if ipt is None:
        ipt = sum(k == TUMOR_DNA for k in sample_kinds) > 1
    header = "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\t" + "\t".join(
        list(sample_names) + (["pooled_normal"] if ipn else []) + (["pooled_tumor"] if ipt else []))
    ordered = sorted(evidences, key=lambda e: (contig_names[e["contig_id"]], e["start"]))  # VCFOutput.scala:97-100
    return [header] + [vcf_row(e, contig_names, sample_names, sample_kinds, **kw) for e in ordered]


def loci_from_vcf_lines(lines: Sequence[str], contig_names: Sequence[str]) -> List[Tuple[int, int, int]]:
    """--loci-file / --force-call-loci-file given a VCF: one range [pos-1, pos-1+len(ref)) per record."""
    out = []
    for l in lines:
        f = l.split("\t")
        out.append((contig_names.index(f[0]), int(f[1]) - 1, int(f[1]) - 1 + len(f[3])))
    return sorted(out)
