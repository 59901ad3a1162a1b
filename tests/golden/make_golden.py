#!/usr/bin/env python
"""Generate the golden vectors in tests/golden/ from oracle/_ref, i.e. from the
reference's own kernel source compiled for the host (see
oracle/ref_host/ref_driver.cpp). Run in the authoring container only
(`make -C oracle ref` first); the outputs are committed and are what pins the
oracle and the CUDA path when /root/reference is not available.

  golden_cases.npz : small grid + degrid problems (inputs and reference
                     outputs) for fp64 3-D, fp64 2-D, an fp32-plan problem
                     evaluated in fp64 ("oracle64") and the literal fp32 host
                     run of the same problem
  golden_plans.json: planning scalars over a sweep of (image size, epsilon,
                     precision, w-range)
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import ref_oracle as ro  # noqa: E402


def small_case(seed, n, rows, chans, fov_deg, w_frac):
    rng = np.random.default_rng(seed)
    pix = fov_deg * np.pi / 180 / n
    freq = 1e9 + np.arange(chans) * 5e7
    uvw = (rng.random((rows, 3)) - 0.5) / (pix * freq[-1] / ro.C_0) * 0.9
    uvw[:, 2] *= w_frac
    # a visibility exactly on the grid origin (W+1-tap footprint), one far
    # off the grid (dropped) and one straddling the grid edge (clipped)
    uvw[0] = 0.0
    uvw[1, :2] = 5.0 / (pix * freq[0] / ro.C_0)
    uvw[2, 0] = 0.4985 / (pix * freq[-1] / ro.C_0)
    vis = rng.random((rows, chans)) - 0.5 + 1j * (rng.random((rows, chans)) - 0.5)
    wgt = rng.random((rows, chans)) + 0.5
    img = rng.random((n, n)) - 0.5
    return pix, freq, uvw, vis, wgt, img


def main():
    out = {}
    specs = [
        # name, seed, N, R, C, fov, w_frac, eps, wstack, plan_prec, compute_prec
        ("f64_3d", 11, 64, 80, 3, 14.7, 0.5, 1e-7, True, 8, 8),
        ("f64_2d", 12, 48, 60, 2, 3.0, 0.5, 1e-5, False, 8, 8),
        ("f64_w14", 13, 32, 40, 2, 8.0, 0.3, 1e-12, True, 8, 8),
        ("f32plan_f64", 14, 64, 80, 3, 10.0, 0.5, 1e-5, True, 4, 8),
    ]
    for name, seed, n, rows, chans, fov, wf, eps, wst, pp, cp in specs:
        pix, freq, uvw, vis, wgt, img = small_case(seed, n, rows, chans, fov, wf)
        if pp == 4:
            # fp32 problem: inputs are rounded to fp32 first, as a caller
            # holding fp32 arrays would pass them
            uvw = uvw.astype(np.float32).astype(np.float64)
            freq = freq.astype(np.float32).astype(np.float64)
            vis = vis.astype(np.complex64).astype(np.complex128)
            wgt = wgt.astype(np.float32).astype(np.float64)
            img = img.astype(np.float32).astype(np.float64)
            mn, mx = ro.get_w_range(uvw.astype(np.float32),
                                    freq.astype(np.float32))
        else:
            mn, mx = ro.get_w_range(uvw, freq)
        g = ro.RefHostGridder(rows, chans, n, pix, eps, wst, mn, mx, pp, cp)
        dirty, _ = g.grid(uvw, freq, vis, wgt, fast=False)
        pred, img_corr, _ = g.degrid(uvw, freq, img, fast=False)
        for key, val in dict(pix=pix, eps=eps, wstack=int(wst), plan_prec=pp,
                             min_abs_w=mn, max_abs_w=mx, freq=freq, uvw=uvw,
                             vis=vis, weight=wgt, image=img, dirty=dirty,
                             pred=pred, image_corrected=img_corr,
                             grid_size=g.info["grid_size"],
                             support=g.info["support"],
                             planes=g.info["num_total_w_grids"]).items():
            out[f"{name}/{key}"] = np.asarray(val)
        if pp == 4:
            # the literal fp32 host run of the same problem, for the
            # three-way error table
            g32 = ro.RefHostGridder(rows, chans, n, pix, eps, wst, mn, mx, 4, 4)
            d32, _ = g32.grid(uvw, freq, vis, wgt, fast=False)
            p32, _, _ = g32.degrid(uvw, freq, img, fast=False)
            out[f"{name}/dirty_ref32"] = d32
            out[f"{name}/pred_ref32"] = p32
    np.savez_compressed(os.path.join(HERE, "golden_cases.npz"), **out)

    plans = []
    for n in (64, 256, 500, 777, 1000, 1024, 4096, 8192, 16384):
        for eps in (1e-2, 1e-3, 1e-4, 1e-5, 1e-6, 1e-7, 1e-9, 1e-10, 1e-12,
                    1e-13, 1e-15):
            for dbl in (0, 1):
                for (pix, mn, mx, wst) in ((1e-5, 3.0, 900.0, 1),
                                           (8.522e-6, 0.0, 5.3e4, 1),
                                           (4.26e-5, 12.5, 2.0e4, 1),
                                           (1e-5, 0.0, 0.0, 0)):
                    g = ro.RefHostGridder(10, 2, n, pix, eps, bool(wst), mn, mx,
                                          8 if dbl else 4, 8)
                    qk, qn, qw, cc = g.tables()
                    rec = dict(image_size=n, epsilon=eps, is_double=dbl,
                               pixel_size=pix, min_abs_w=mn, max_abs_w=mx,
                               do_wstacking=wst)
                    rec.update({k: g.info[k] for k in (
                        "grid_size", "support", "num_total_w_grids", "beta",
                        "uv_scale", "w_scale", "inv_w_scale", "min_plane_w",
                        "max_plane_w", "conv_corr_norm_factor")})
                    rec["table_sums"] = [float(qk.sum()), float(qn.sum()),
                                         float(qw.sum()), float(cc.sum()),
                                         float(cc[-1])]
                    plans.append(rec)
    with open(os.path.join(HERE, "golden_plans.json"), "w") as f:
        json.dump(plans, f, indent=0)
    print("cases:", len(specs), "plans:", len(plans))


if __name__ == "__main__":
    main()
