#!/usr/bin/env python
"""Where does the parity error at the C4 grid size (G = 24640, N = 16384,
fp32) come from? Error maps of the product against the fp64 oracle for the
case of tests/test_gpu_timed_shapes.py::test_c4_grid_size_24640, with the own
line FFTs, with the cuFFT fallback, and without w-stacking."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ska-sdp-func_b200"))
import numpy as np  # noqa: E402

from oracle import es_oracle as eo  # noqa: E402
from ska_sdp_func.grid_data import GridderUvwEsFft  # noqa: E402
from tests import synth  # noqa: E402


def describe(tag, got, want):
    got = got.astype(np.float64)
    err = np.abs(got - want)
    n = want.shape[0]
    mx = np.max(np.abs(want))
    i, j = np.unravel_index(np.argmax(err), err.shape)
    wi, wj = np.unravel_index(np.argmax(np.abs(want)), want.shape)
    q = n // 4
    c = (slice(q, n - q), slice(q, n - q))
    print(f"{tag}: max|want| {mx:.4e} at {(wi, wj)}; max err {err.max():.3e} "
          f"({err.max() / mx:.3e} of max) at {(i, j)} where want = {want[i, j]:.3e}; "
          f"rms err {np.sqrt(np.mean(err ** 2)):.3e}, rms want "
          f"{np.sqrt(np.mean(want ** 2)):.3e}; centre box: max err "
          f"{err[c].max():.3e} / max want {np.abs(want[c]).max():.3e} = "
          f"{err[c].max() / np.abs(want[c]).max():.3e}", flush=True)
    # error by ring: max err and max want in 8 bands of max(|i|, |j|)
    ii = np.abs(np.arange(n) - n // 2)
    r = np.maximum(ii[:, None], ii[None, :])
    for b in range(8):
        m = (r >= b * n // 16) & (r < (b + 1) * n // 16)
        print(f"   band {b}: max err {err[m].max():.3e}  max want {np.abs(want[m]).max():.3e}"
              f"  ratio {err[m].max() / np.abs(want[m]).max():.3e}", flush=True)


def run(case, eps, wstack, lo_hi, tag, want=None):
    n = case["image_size"]
    real = case["uvw"].dtype
    dirty = np.zeros((n, n), dtype=real)
    g = GridderUvwEsFft(case["uvw"], case["freq_hz"], case["vis"], case["weight"],
                        dirty, case["pixel_size"], case["pixel_size"], eps, wstack)
    info = g.info()
    planes = info["num_total_w_grids"]
    lo, hi = lo_hi if wstack else (0, planes)
    g.set_plane_range(lo, hi, apply_correction=True)
    g.grid_uvw_es_fft(case["uvw"], case["freq_hz"], case["vis"], case["weight"], dirty)
    st = g.stats()
    print(tag, "planes", (lo, hi), "of", planes, "records", st["num_entries"],
          "pruned_fft", st["pruned_fft"], flush=True)
    if want is None:
        o = eo.OracleGridder(case["uvw"], case["freq_hz"], case["vis"], case["weight"],
                             dirty, case["pixel_size"], case["pixel_size"], eps, wstack)
        t0 = time.time()
        want, _ = o.grid_uvw_es_fft(case["uvw"], case["freq_hz"], case["vis"],
                                    case["weight"], planes=(lo, hi))
        print("  oracle", round(time.time() - t0, 1), "s", flush=True)
    describe(tag, dirty, want)
    return dirty, want


def main():
    case = synth.ska_low_snapshot(512, 2, 16384, seed=43, single=True)
    planes = 1488
    mid = (planes // 2 - 1, planes // 2 + 1)
    ours, want = run(case, 1e-5, True, mid, "own FFT, w-stacking, planes mid")
    os.environ["SKA_SDP_FUNC_ES_USE_CUFFT"] = "1"
    cu, _ = run(case, 1e-5, True, mid, "cuFFT fallback, same planes", want)
    del os.environ["SKA_SDP_FUNC_ES_USE_CUFFT"]
    describe("own FFT vs cuFFT fallback", ours, cu.astype(np.float64))
    # planes near w = 0 where most visibilities are
    ours0, want0 = run(case, 1e-5, True, (2, 4), "own FFT, planes 2-3")
    # no w-stacking: one plane, all visibilities
    run(case, 1e-5, False, None, "own FFT, 2-D mode")
    # other radix lists for the same length
    for rad in ("32,22,7,5", "28,22,8,5", "11,10,14,16"):
        os.environ["SKA_SDP_FUNC_ES_FFT_RADICES"] = rad
        run(case, 1e-5, True, mid, f"own FFT radices {rad}", want)
    del os.environ["SKA_SDP_FUNC_ES_FFT_RADICES"]


if __name__ == "__main__":
    main()
