"""The reference's caller unit suites against the oracle (no GPU needed); the same vectors run through the device in
tests/test_gpu_caller_suites.py."""
import pytest

from oracle import oracle
from tests import caller_suite_cases as cases
from tests.golden import fixtures


@pytest.mark.parametrize("case", cases.SS_CASES, ids=[c[0] for c in cases.SS_CASES])
def test_somatic_standard_caller_suite(case):
    cases.check_ss_case(case, oracle.somatic_standard)


def test_somatic_joint_caller_suite():
    batch, ref, _ = fixtures.load("tiny_sjc")
    cases.check_joint_cases(batch, ref, oracle.somatic_joint)


@pytest.mark.parametrize("name", ["syn1fp", "complexvar", "simplefp"])
def test_somatic_standard_caller_real_data_suite(name):
    batch, ref, extra = fixtures.load("real_" + name)
    cases.check_real_data_case(name, batch, ref, extra["positions"], oracle.somatic_standard)


@pytest.mark.parametrize("case", cases.ALLELE_EVIDENCE_CASES, ids=["all reads support", "one read supports"])
def test_allele_evidence_suite(case):
    cases.check_allele_evidence_case(case, oracle.germline_standard)
