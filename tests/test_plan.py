"""Planning maths of the product library against the reference (CPU only).

The product's C planning code (csrc/es_plan.c) must reproduce the reference's
grid size, support, beta, w-plane lattice and correction tables exactly
(SURVEY.md A.1/A.2, 7.5 #4). Golden values come from oracle/_ref
(tests/golden/make_golden.py).
"""
import json
import os

import numpy as np
import pytest

from oracle import es_oracle as eo
from ska_sdp_func.grid_data import GridderUvwEsFft

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "golden_plans.json")


def test_plan_scalars_match_reference_sweep():
    plans = json.load(open(GOLDEN))
    assert len(plans) > 500
    for rec in plans:
        info, qk, qn, qw, cc = GridderUvwEsFft.plan_info(
            10, 2, rec["image_size"], rec["is_double"], rec["pixel_size"],
            rec["epsilon"], rec["min_abs_w"], rec["max_abs_w"],
            rec["do_wstacking"])
        for key in ("grid_size", "support", "num_total_w_grids"):
            assert info[key] == rec[key], (key, rec)
        for key in ("beta", "uv_scale", "w_scale", "inv_w_scale",
                    "min_plane_w", "max_plane_w", "conv_corr_norm_factor"):
            assert info[key] == pytest.approx(rec[key], rel=4e-15, abs=1e-300), (key, rec)
        sums = [qk.sum(), qn.sum(), qw.sum(), cc.sum(), cc[-1]]
        assert np.allclose(sums, rec["table_sums"], rtol=1e-13, atol=0)


def test_baseline_config_plans():
    """W, G of the BASELINE configs."""
    expect = [  # image, is_double, eps -> support, grid
        (256, 1, 1e-5, 8, 392), (4096, 0, 1e-5, 8, 6160),
        (8192, 1, 1e-10, 12, 14400), (16384, 0, 1e-5, 8, 24640),
        (8192, 0, 1e-4, 7, 11520), (1024, 1, 1e-12, 14, 1920),
        (1024, 0, 1e-5, 8, 1540),
    ]
    for n, dbl, eps, w, g in expect:
        info = GridderUvwEsFft.plan_info(10, 2, n, dbl, 1e-5, eps, 0.0, 100.0, True)[0]
        assert (info["support"], info["grid_size"]) == (w, g)


def test_unattainable_epsilon_falls_back_to_last_kernel():
    """Documented limits are not enforced: an impossible epsilon silently picks
    the last table row, W=16 sigma=2.0 (es_fft_utils.cpp:492-496)."""
    info = GridderUvwEsFft.plan_info(10, 2, 256, 1, 1e-5, 1e-30, 0.0, 10.0, True)[0]
    assert info["support"] == 16 and info["grid_size"] == 528


def test_plan_matches_oracle_plan_maths():
    for n, dbl, eps, pix, mn, mx in [(256, 1, 1e-5, 1.36e-4, 0.3, 812.0),
                                     (1000, 0, 1e-4, 3e-5, 0.0, 4000.0),
                                     (4096, 0, 1e-5, 8.5e-6, 10.0, 5e4)]:
        info, qk, qn, qw, cc = GridderUvwEsFft.plan_info(
            10, 2, n, dbl, pix, eps, mn, mx, True)
        pl = eo.make_plan(n, pix, eps, True, mn, mx, bool(dbl))
        for key in ("grid_size", "support", "num_total_w_grids"):
            assert info[key] == pl[key]
        for key in ("beta", "uv_scale", "w_scale", "min_plane_w",
                    "conv_corr_norm_factor"):
            assert info[key] == pytest.approx(pl[key], rel=4e-15)
        assert np.allclose(qk, pl["quad_kernel"], rtol=0, atol=1e-15)
        assert np.allclose(qn, pl["quad_nodes"], rtol=0, atol=1e-15)
        assert np.allclose(qw, pl["quad_weights"], rtol=0, atol=1e-15)
        assert np.allclose(cc, pl["conv_corr"], rtol=0, atol=1e-14)


def test_degenerate_w_range_is_guarded():
    """max_abs_w == min_abs_w divides by zero in the reference (es_fft.cpp:358);
    the drop-in keeps a finite lattice."""
    info = GridderUvwEsFft.plan_info(10, 2, 256, 1, 1e-4, 1e-5, 7.0, 7.0, True)[0]
    assert np.isfinite(info["w_scale"]) and info["w_scale"] > 0
    assert info["num_total_w_grids"] == 2 + info["support"] - 2
