"""The C++ host mirror (include/guac_b200.hpp: pileupFlatMapOneSample / TwoSamples / MultipleSamples over the C ABI) against
the vectors of the reference's PileupFlatMapUtilsSuite — tests/cpp/test_operators.cpp, compiled here with g++ and run on
the device."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "guacamole_b200", "csrc")


def compile_operators_test(out):
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "cpp", "test_operators.cpp"),
           "-L", LIBDIR, "-lguac_b200", f"-Wl,-rpath,{LIBDIR}", "-o", out]
    return subprocess.run(cmd, capture_output=True, text=True)


def test_header_compiles_and_links(built, tmp_path):
    """No GPU needed: the mirror is plain C++17 over the C ABI."""
    r = compile_operators_test(str(tmp_path / "test_operators"))
    assert r.returncode == 0, r.stderr


@pytest.mark.gpu
def test_pileup_flat_map_utils_suite_in_cpp(built, tmp_path):
    exe = str(tmp_path / "test_operators")
    r = compile_operators_test(exe)
    assert r.returncode == 0, r.stderr
    run = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert run.returncode == 0 and run.stdout.strip().endswith("ok"), run.stdout + run.stderr
