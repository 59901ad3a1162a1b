"""The line-FFT core of the pruned w-plane FFT kernels (csrc/es_fft_core.h),
compiled for the host and run against numpy.fft: radix passes, twiddle tables
and the digit-reversal map, for both flow graphs (DIF: gridding, DIT:
degridding), both directions, both precisions, on the reference's grid sizes
(2 * good_size, 11-smooth; es_fft_utils.cpp:181-218)."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "ska-sdp-func_b200", "csrc")
EXE = os.path.join(ROOT, "build", "fft_core_host")


@pytest.fixture(scope="module")
def exe():
    os.makedirs(os.path.dirname(EXE), exist_ok=True)
    obj = EXE + "_plan.o"
    subprocess.run(["gcc", "-std=c11", "-O2", "-I", CSRC, "-c",
                    os.path.join(CSRC, "es_fft_plan.c"), "-o", obj], check=True)
    subprocess.run(["g++", "-std=c++17", "-O2", "-I", CSRC,
                    os.path.join(ROOT, "tests", "fft_core_host.cpp"), obj,
                    "-lm", "-o", EXE], check=True)
    return EXE


# grid sizes of BASELINE configs C1..C5, the reference test (1540, 1920) and odd cases
SIZES = [392, 1540, 1920, 6160, 11520, 14400, 24640, 98, 120, 2 * 3 * 5 * 7 * 11]


@pytest.mark.parametrize("n", SIZES)
@pytest.mark.parametrize("double", [0, 1])
def test_line_fft_matches_numpy(exe, n, double, tmp_path):
    out = str(tmp_path / "fft.bin")
    for sign in (1, -1):
        for dit in (0, 1):
            subprocess.run([exe, str(n), str(double), str(sign), str(dit), out],
                           check=True, capture_output=True)
            d = np.fromfile(out, dtype=np.complex128)
            x, y = d[:n], d[n:]
            ref = np.fft.ifft(x) * n if sign > 0 else np.fft.fft(x)
            err = np.max(np.abs(y - ref)) / np.max(np.abs(ref))
            assert err < (1e-14 if double else 2e-6), (n, double, sign, dit, err)


def test_unsupported_sizes_are_reported(exe, tmp_path):
    out = str(tmp_path / "fft.bin")
    r = subprocess.run([exe, str(2 * 13), "0", "1", "0", out], capture_output=True,
                       text=True)
    assert "unsupported" in r.stdout
