"""The oracle against its pins (runs on CPU, no GPU needed).

Pins, in decreasing authority:
  1. golden vectors generated from oracle/_ref = the reference's own kernel
     source compiled for the host (tests/golden/make_golden.py);
  2. oracle/_ref itself when the prebuilt library is present;
  3. the explicit DFT (signs, axis order, normalisation, 1/n);
  4. the reference's only own numeric criterion for this path: adjointness
     (tests/grid_data/test_gridder_uvw_es_fft.py:500-529,637).
"""
import os

import numpy as np
import pytest

from oracle import dft, es_oracle as eo, ref_oracle as ro
from tests import synth

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "golden_cases.npz")


def load_golden(name):
    z = np.load(GOLDEN)
    return {k.split("/", 1)[1]: z[k] for k in z.files if k.startswith(name + "/")}


def oracle_for(g):
    is_double = int(g["plan_prec"]) == 8
    # the golden w-range was computed in the caller's precision
    uvw = g["uvw"] if is_double else g["uvw"].astype(np.float32)
    freq = g["freq"] if is_double else g["freq"].astype(np.float32)
    return eo.OracleGridder(uvw, freq, g["vis"], g["weight"], g["image"],
                            float(g["pix"]), float(g["pix"]), float(g["eps"]),
                            bool(g["wstack"]), is_double=is_double)


def rel_max(a, b):
    return float(np.max(np.abs(a - b)) / np.max(np.abs(b)))


def test_oracle_fft_matches_numpy():
    rng = np.random.default_rng(0)
    for n in (12, 98, 120, 392, 770):  # 11-smooth sizes incl. factors 7, 11
        a = rng.standard_normal((n, n)) + 1j * rng.standard_normal((n, n))
        fwd = eo.fft2(a, -1)
        bwd = eo.fft2(a, +1)
        assert rel_max(fwd, np.fft.fft2(a)) < 1e-13
        assert rel_max(bwd, np.fft.ifft2(a) * n * n) < 1e-13


@pytest.mark.parametrize("name", ["f64_3d", "f64_2d", "f64_w14", "f32plan_f64"])
def test_oracle_matches_golden(name):
    g = load_golden(name)
    o = oracle_for(g)
    assert o.plan["grid_size"] == int(g["grid_size"])
    assert o.plan["support"] == int(g["support"])
    assert o.plan["num_total_w_grids"] == int(g["planes"])
    dirty, _ = o.grid_uvw_es_fft(g["uvw"], g["freq"], g["vis"], g["weight"])
    pred, img, _ = o.ifft_grid_uvw_es(g["uvw"], g["freq"], g["image"])
    assert rel_max(dirty, g["dirty"]) < 1e-12
    assert rel_max(pred, g["pred"]) < 1e-12
    assert rel_max(img, g["image_corrected"]) < 1e-13


def test_golden_fp32_reference_is_further_from_oracle64_than_tolerance_allows():
    """Documents SURVEY.md 0.4: the reference's literal fp32 arithmetic is
    itself ~1e-4..1e-6 away from the fp64-coordinate evaluation of the same
    algorithm, so fp32 parity is defined against oracle64."""
    g = load_golden("f32plan_f64")
    err_grid = rel_max(g["dirty_ref32"], g["dirty"])
    err_pred = rel_max(g["pred_ref32"], g["pred"])
    assert 1e-8 < err_grid < 1e-2
    assert 1e-8 < err_pred < 1e-2


@pytest.mark.skipif(not ro.available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("wstack,eps", [(True, 1e-5), (False, 1e-5), (True, 1e-12)])
def test_oracle_matches_reference_host_build_c1(wstack, eps):
    """BASELINE config C1: the reference test generator at 256^2, fp64."""
    c = synth.reference_test_case(256)
    mn, mx = ro.get_w_range(c["uvw"], c["freq_hz"])
    ref = ro.RefHostGridder(1000, 10, 256, c["pixel_size"], eps, wstack, mn, mx, 8, 8)
    d_ref, _ = ref.grid(c["uvw"], c["freq_hz"], c["vis"], c["weight"], fast=False)
    v_ref, i_ref, _ = ref.degrid(c["uvw"], c["freq_hz"], c["image"], fast=False)
    o = eo.OracleGridder(c["uvw"], c["freq_hz"], c["vis"], c["weight"],
                         c["image"], c["pixel_size"], c["pixel_size"], eps, wstack)
    d, _ = o.grid_uvw_es_fft(c["uvw"], c["freq_hz"], c["vis"], c["weight"])
    v, i, _ = o.ifft_grid_uvw_es(c["uvw"], c["freq_hz"], c["image"])
    assert rel_max(d, d_ref) < 1e-12
    assert rel_max(v, v_ref) < 1e-12
    assert rel_max(i, i_ref) < 1e-13


@pytest.mark.parametrize("eps,tol", [(1e-5, 2e-5), (1e-7, 2e-7), (1e-12, 1e-10)])
def test_oracle_matches_explicit_dft(eps, tol):
    rng = np.random.default_rng(1)
    n, rows, chans = 48, 40, 2
    pix = 12.0 * np.pi / 180 / n
    freq = 1e9 + np.arange(chans) * 5e7
    uvw = (rng.random((rows, 3)) - 0.5) / (pix * freq[-1] / eo.C_0) * 0.8
    uvw[:, 2] *= 0.5
    vis = rng.random((rows, chans)) - 0.5 + 1j * (rng.random((rows, chans)) - 0.5)
    wgt = rng.random((rows, chans)) + 0.5
    img = rng.random((n, n)) - 0.5
    o = eo.OracleGridder(uvw, freq, vis, wgt, img, pix, pix, eps, True)
    d, _ = o.grid_uvw_es_fft(uvw, freq, vis, wgt)
    v, _, _ = o.ifft_grid_uvw_es(uvw, freq, img)
    d_dft = dft.dft_grid(uvw, freq, vis, wgt, n, pix)
    v_dft = dft.dft_degrid(uvw, freq, img, pix)
    assert np.linalg.norm(d - d_dft) / np.linalg.norm(d_dft) < tol
    assert np.linalg.norm(v - v_dft) / np.linalg.norm(v_dft) < tol


@pytest.mark.parametrize("wstack", [False, True])
def test_oracle_adjointness_reference_criterion(wstack):
    """<grid(vis), img> == Re<vis, degrid(img)> to 1e-12 in fp64 at eps=1e-12
    (reference threshold, test_gridder_uvw_es_fft.py:516-529,637)."""
    c = synth.reference_test_case(256)
    o = eo.OracleGridder(c["uvw"], c["freq_hz"], c["vis"], c["weight"],
                         c["image"], c["pixel_size"], c["pixel_size"], 1e-12, wstack)
    d, _ = o.grid_uvw_es_fft(c["uvw"], c["freq_hz"], c["vis"], c["weight"])
    v, _, _ = o.ifft_grid_uvw_es(c["uvw"], c["freq_hz"], c["image"])
    adj1 = np.vdot(d, c["image"])
    adj2 = np.vdot(v, c["vis"]).real
    assert abs(adj1 - adj2) / max(abs(adj1), abs(adj2)) < 1e-12


def test_oracle_accumulates_and_degrid_adds():
    """Reference side effects (SURVEY.md A.5 #1,#2): gridding accumulates into
    the image before correcting; degridding adds to vis and ignores weight."""
    g = load_golden("f64_3d")
    o = oracle_for(g)
    base = np.full_like(g["dirty"], 0.25)
    d0, _ = o.grid_uvw_es_fft(g["uvw"], g["freq"], g["vis"], g["weight"])
    d1, _ = o.grid_uvw_es_fft(g["uvw"], g["freq"], g["vis"], g["weight"], base)
    # (acc + 0.25) * corr = acc * corr + 0.25 * corr; corr = image_corrected / image
    corr = g["image_corrected"] / g["image"]
    assert rel_max(d1, d0 + 0.25 * corr) < 1e-12
    pred0, _, _ = o.ifft_grid_uvw_es(g["uvw"], g["freq"], g["image"])
    start = np.full(pred0.shape, 1.0 - 2.0j)
    pred1, _, _ = o.ifft_grid_uvw_es(g["uvw"], g["freq"], g["image"], vis=start)
    assert rel_max(pred1, pred0 + start) < 1e-12
