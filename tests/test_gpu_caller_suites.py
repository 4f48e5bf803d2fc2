"""The reference's caller unit suites through the C ABI on the device (vectors in tests/caller_suite_cases.py)."""
import pytest

from tests import caller_suite_cases as cases
from tests.golden import fixtures

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("impl", ["tile", "gather", "stream"])
@pytest.mark.parametrize("case", cases.SS_CASES, ids=[c[0] for c in cases.SS_CASES])
def test_somatic_standard_caller_suite(gpu_ctx, case, impl):
    def run(batch, ref, ranges, params):
        gpu_ctx.upload_reference(ref, batch.contig_names)
        gpu_ctx.upload_reads(batch)
        return gpu_ctx.somatic_standard(ranges, params)

    gpu_ctx.set_option("ss_impl", {"tile": 0, "gather": 1, "stream": 2}[impl])
    try:
        cases.check_ss_case(case, run)
    finally:
        gpu_ctx.set_option("ss_impl", 2)


def joint_run(gpu_ctx):
    def run(batch, ref, loci, params, force_loci):
        gpu_ctx.upload_reference(ref, batch.contig_names)
        gpu_ctx.upload_reads(batch)
        return gpu_ctx.somatic_joint(loci, params, force_loci=force_loci)

    return run


@pytest.mark.parametrize("fallback_mod", [0, 1, 3])
def test_somatic_joint_caller_suite(gpu_ctx, fallback_mod):
    """fallback_mod 1 / 3: every / every third tile goes through the dense per-locus discovery instead of the prefilter."""
    batch, ref, _ = fixtures.load("tiny_sjc")
    gpu_ctx.set_option("ss_force_fallback_mod", fallback_mod)
    try:
        cases.check_joint_cases(batch, ref, joint_run(gpu_ctx))
    finally:
        gpu_ctx.set_option("ss_force_fallback_mod", 0)


@pytest.mark.parametrize("impl", ["tile", "gather", "stream"])
@pytest.mark.parametrize("name", ["syn1fp", "complexvar", "simplefp"])
def test_somatic_standard_caller_real_data_suite(gpu_ctx, name, impl):
    batch, ref, extra = fixtures.load("real_" + name)
    gpu_ctx.upload_reference(ref, batch.contig_names)
    gpu_ctx.upload_reads(batch)
    gpu_ctx.set_option("ss_impl", {"tile": 0, "gather": 1, "stream": 2}[impl])
    try:
        cases.check_real_data_case(name, batch, ref, extra["positions"], lambda b, r, ranges, p: gpu_ctx.somatic_standard(ranges, p))
    finally:
        gpu_ctx.set_option("ss_impl", 2)


@pytest.mark.parametrize("case", cases.ALLELE_EVIDENCE_CASES, ids=["all reads support", "one read supports"])
def test_allele_evidence_suite(gpu_ctx, case):
    def run(batch, ref, ranges, params):
        gpu_ctx.upload_reference(ref, batch.contig_names)
        gpu_ctx.upload_reads(batch)
        return gpu_ctx.germline_standard(ranges, params)

    cases.check_allele_evidence_case(case, run)
