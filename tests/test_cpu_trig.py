"""gdtrig::sinf_glibc / cosf_glibc (gpudrive_lab_b200/csrc/glibc_trig.h), the restatement of glibc 2.39's
FMA sinf / cosf that the kernels use for per-agent state, is bit-identical to this machine's libm on a
strided sweep of |x| < 120 (the full 748M-input sweep is `check_trig 1`). The reference's trig is libm's
(src/dynamics.hpp:62-76, external/madrona/include/madrona/math.inl:680-691)."""
import os
import subprocess

import pytest

from helpers import ROOT


def test_glibc_exact_trig(tmp_path):
    exe = str(tmp_path / "check_trig")
    src = os.path.join(ROOT, "tests", "csrc", "check_trig.cpp")
    subprocess.check_call(["g++", "-O2", "-mfma", "-ffp-contract=off", "-o", exe, src, "-lm"])
    flags = open("/proc/cpuinfo").read()
    if " fma " not in flags and " fma\n" not in flags:
        pytest.skip("host CPU has no FMA: its libm selects a different sinf/cosf variant")
    r = subprocess.run([exe, "997"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:]
    assert "sin mismatches 0, cos mismatches 0" in r.stdout
