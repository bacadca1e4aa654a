"""Host-side proof that magritte_b200/csrc/mgb_libm.h reproduces the box's libm bit for bit.

The same header is what the CUDA kernels compile (with __fma_rn in place of fma()); here it is built
for the host with explicit fma() and -ffp-contract=off and compared with glibc over ALL 2^32 float
inputs for expf / logf / erff / expm1f and over a lattice of 2^32/257 x-values times 73 exponents for
powf (the exponents the reference can produce: 1 - i/n, solver.tpp:1083, plus edge cases).
reference call sites: solver.tpp:1109,1120,2450; solver_interp_helper.tpp:37,39,67,68,98.
"""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def checker(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("libm") / "libm_host_check")
    subprocess.check_call(["gcc", "-O2", "-mfma", "-ffp-contract=off", "-fopenmp",
                           "-I", os.path.join(ROOT, "magritte_b200", "csrc"),
                           os.path.join(ROOT, "tests", "csrc", "libm_host_check.c"), "-o", exe, "-lm"])
    return exe


@pytest.mark.parametrize("fn,stride", [("expf", 1), ("logf", 1), ("erff", 1), ("erff_u", 1), ("expm1f", 1), ("powf", 257)])
def test_bit_exact_against_host_libm(checker, fn, stride):
    out = subprocess.run([checker, fn, str(stride)], capture_output=True, text=True)
    assert out.returncode == 0 and "mismatches=0" in out.stdout, out.stdout + out.stderr
