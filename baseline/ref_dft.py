#!/usr/bin/env python
"""Drives the UNMODIFIED reference (baseline/_ref, built for sm_100 by
baseline/build_ref_gpu.sh) through its own Python API of the DFT predict,
``ska_sdp_func.visibility.dft_point_v00 / dft_point_v01``, on device arrays --
the reference's CUDA kernels on this GPU. Harness only; runs in its OWN
process (the reference library and the product library share file name and
symbols).

  python baseline/ref_dft.py --case '{"seed": 8}' [--out x.npz] [--steps 3]
prints one JSON line {"available": true, "ms_v01_c128": .., ...}; with --out
also stores v00 / v01 visibilities in complex128 and complex64.
"""
import argparse
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.path.join(HERE, "_ref")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case", default="{}")
    ap.add_argument("--out", default="")
    ap.add_argument("--steps", type=int, default=1)
    ap.add_argument("--skip-v00", action="store_true")
    args = ap.parse_args()
    if not os.path.exists(os.path.join(REF, "ska_sdp_func", "__init__.py")):
        print(json.dumps({"available": False, "why": "baseline/_ref is not "
                          "installed (baseline/build_ref_gpu.sh)"}))
        return
    for p in (os.path.join(HERE, "cupy_shim"), REF):
        sys.path.insert(0, p)
    os.environ["SKA_SDP_FUNC_LIB_DIR"] = REF
    import numpy as np
    import torch
    import cupy  # the shim
    import ska_sdp_func
    from ska_sdp_func.visibility import dft_point_v00, dft_point_v01
    assert os.path.realpath(ska_sdp_func.__file__).startswith(os.path.realpath(REF))
    sys.path.insert(0, ROOT)
    from tests.synth_dft import dft_case

    c = dft_case(**json.loads(args.case))
    dev = torch.device("cuda", 0)
    keys = ("directions", "fluxes", "uvw") + (() if args.skip_v00 else ("uvw_lambda",))
    t = {k: torch.from_numpy(c[k]).to(dev) for k in keys}
    d = {k: cupy.asarray(v) for k, v in t.items()}

    def timed(fn):
        best = None
        for _ in range(args.steps):
            torch.cuda.synchronize()
            e0, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            best = ms if best is None else min(best, ms)
        return best

    res = {"available": True}
    out = {}
    for dtype, tag in ((torch.complex128, "c128"), (torch.complex64, "c64")):
        vis = torch.zeros(c["vis_shape"], dtype=dtype, device=dev)
        res[f"ms_v01_{tag}"] = timed(lambda: dft_point_v01(
            d["directions"], d["fluxes"], d["uvw"], c["channel_start_hz"],
            c["channel_step_hz"], cupy.asarray(vis)))
        out[f"v01_{tag}"] = vis.cpu().numpy()
        if not args.skip_v00:
            vis = torch.zeros(c["vis_shape"], dtype=dtype, device=dev)
            res[f"ms_v00_{tag}"] = timed(lambda: dft_point_v00(
                d["directions"], d["fluxes"], d["uvw_lambda"], cupy.asarray(vis)))
            out[f"v00_{tag}"] = vis.cpu().numpy()
    if args.out:
        np.savez(args.out, **out)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
