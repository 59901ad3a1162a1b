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
SIZES = [392, 1540, 1920, 6160, 11520, 14400, 24640, 98, 120, 28, 44, 2 * 3 * 5 * 7 * 11]


def _check(exe, n, double, lshift, out, radices=None):
    lines = 1 << lshift
    for sign in (1, -1):
        for dit in (0, 1):
            cmd = [exe, str(n), str(double), str(sign), str(dit), str(lshift), out]
            if radices:
                cmd.append(radices)
            r = subprocess.run(cmd, check=True, capture_output=True, text=True)
            assert "unsupported" not in r.stdout, r.stdout
            d = np.fromfile(out, dtype=np.complex128).reshape(2, lines, n)
            x, y = d[0], d[1]
            ref = np.fft.ifft(x, axis=1) * n if sign > 0 else np.fft.fft(x, axis=1)
            err = np.max(np.abs(y - ref)) / np.max(np.abs(ref))
            assert err < (1e-14 if double else 2e-6), (n, double, sign, dit, err)


@pytest.mark.parametrize("n", SIZES)
@pytest.mark.parametrize("double", [0, 1])
def test_line_fft_matches_numpy(exe, n, double, tmp_path):
    _check(exe, n, double, 0, str(tmp_path / "fft.bin"))


@pytest.mark.parametrize("lshift", [1, 2])
def test_interleaved_lines(exe, lshift, tmp_path):
    """bundles of 2 / 4 lines share one pass (column kernels)"""
    _check(exe, 1540, 0, lshift, str(tmp_path / "fft.bin"))
    _check(exe, 392, 1, lshift, str(tmp_path / "fft.bin"))


# every butterfly radix at least once, as first, middle and last pass
@pytest.mark.parametrize("radices", [
    "32,28,3", "3,28,32", "24,22,5", "5,24,22", "20,15,14", "14,20,15", "12,10,9",
    "9,12,10", "16,6,11", "11,16,6", "8,7,4", "4,2,7", "2,3,5,7,11", "6,6", "30"])
def test_forced_radices(exe, radices, tmp_path):
    vals = [int(v) for v in radices.split(",")]
    n = int(np.prod(vals))
    out = str(tmp_path / "fft.bin")
    if 30 in vals:  # no butterfly for 30
        r = subprocess.run([exe, str(n), "0", "1", "0", "0", out, radices],
                           capture_output=True, text=True)
        assert "unsupported" in r.stdout
        return
    _check(exe, n, 0, 0, out, radices)
    if max(vals) <= 16:
        _check(exe, n, 1, 1, out, radices)


def test_unsupported_sizes_are_reported(exe, tmp_path):
    out = str(tmp_path / "fft.bin")
    r = subprocess.run([exe, str(2 * 13), "0", "1", "0", "0", out], capture_output=True,
                       text=True)
    assert "unsupported" in r.stdout


def test_screen_phasor_matches_libm(tmp_path):
    """csrc/es_screen.h: the fp32 w-screen phasor without conversion / special
    function instructions, against libm (max abs error of cos/sin, max rel
    error of 1/n, max abs error of the sincos polynomial)."""
    exe_path = os.path.join(ROOT, "build", "screen_host")
    os.makedirs(os.path.dirname(exe_path), exist_ok=True)
    subprocess.run(["g++", "-std=c++17", "-O2", "-I", CSRC,
                    os.path.join(ROOT, "tests", "screen_host.cpp"), "-lm", "-o", exe_path],
                   check=True)
    r = subprocess.run([exe_path], check=True, capture_output=True, text=True)
    ph, inv, sc, fx32, fx24 = (float(v) for v in r.stdout.split())
    assert ph < 4e-7 and inv < 2e-7 and sc < 2e-7, r.stdout
    # table form (es_phasor_f32): phase reduced by one fp64 fma into 32 / 24
    # bit fixed point; host build uses libm's cosf / sinf on the reduced phase
    assert fx32 < 4e-7 and fx24 < 5e-7, r.stdout
