#!/usr/bin/env python
"""Drives the UNMODIFIED reference (baseline/_ref: its libska_sdp_func.so built
for sm_100 by baseline/build_ref_gpu.sh and its own Python package) through
its own public API, ``ska_sdp_func.grid_data.GridderUvwEsFft`` -- the
reference's CUDA implementation of the path on this GPU. Harness only.

Runs in its OWN process: the reference library and the product library have
the same file name and symbols and must not share one.

  python baseline/ref_gpu.py --workload C2 --direction grid --steps 2 [--out x.npz]
prints one JSON line {"available": true, "ms_per_call": ..., "Mvis_per_s": ...};
with --out also stores the dirty image / predicted visibilities (parity tests).
"""
import argparse
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.path.join(HERE, "_ref")


def load_reference():
    """Imports the reference package from baseline/_ref (never the product)."""
    if not os.path.exists(os.path.join(REF, "ska_sdp_func", "__init__.py")):
        return None
    for p in (os.path.join(HERE, "cupy_shim"), REF):
        sys.path.insert(0, p)
    os.environ["SKA_SDP_FUNC_LIB_DIR"] = REF
    import cupy  # the shim
    import ska_sdp_func
    from ska_sdp_func.grid_data import GridderUvwEsFft
    from ska_sdp_func.utility import Lib
    assert os.path.realpath(ska_sdp_func.__file__).startswith(os.path.realpath(REF))
    Lib.handle()
    loaded = Lib.handle()._name  # path of the CDLL the reference package opened
    assert os.path.realpath(loaded).startswith(os.path.realpath(REF)), loaded
    return cupy, GridderUvwEsFft


def run(case, eps, wstack, direction, steps, warmup=1, keep=False):
    """Times `steps` reference calls on device-resident inputs (CUDA events on
    the legacy default stream the reference launches on). Returns a dict."""
    import numpy as np
    import torch
    ref = load_reference()
    if ref is None:
        return {"available": False, "why": "baseline/_ref is not installed "
                "(run baseline/build_ref_gpu.sh where /root/reference exists)"}
    cupy, gridder_class = ref
    dev = torch.device("cuda", torch.cuda.current_device())
    t = {k: torch.from_numpy(case[k]).to(dev)
         for k in ("uvw", "freq_hz", "vis", "weight")}
    n = case["image_size"]
    img0 = torch.zeros((n, n), dtype=t["uvw"].dtype, device=dev)
    if direction == "degrid":
        if case.get("image") is not None:
            img0.copy_(torch.from_numpy(case["image"]))
        else:
            img0.uniform_(-0.5, 0.5, generator=torch.Generator(dev).manual_seed(5))
    img = img0.clone()
    vis = t["vis"].clone()
    wrap = cupy.asarray
    uvw_c, freq_c, wgt_c = wrap(t["uvw"]), wrap(t["freq_hz"]), wrap(t["weight"])
    try:
        g = gridder_class(uvw_c, freq_c, wrap(vis), wgt_c, wrap(img),
                          case["pixel_size"], case["pixel_size"], eps, wstack)
    except Exception as err:  # pylint: disable=broad-except
        return {"available": False, "why": f"plan creation failed: {err}"}

    def call():
        if direction == "grid":
            img.zero_()
            g.grid_uvw_es_fft(uvw_c, freq_c, wrap(vis), wgt_c, wrap(img))
        else:
            img.copy_(img0)
            vis.zero_()
            g.ifft_grid_uvw_es(uvw_c, freq_c, wrap(vis), wgt_c, wrap(img))

    for _ in range(warmup):
        call()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        call()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / max(steps, 1)
    n_vis = case["vis"].size
    out = {"available": True, "ms_per_call": ms,
           "Mvis_per_s": n_vis / (ms * 1e-3) / 1e6, "steps": steps,
           "warmup": warmup, "direction": direction,
           "how": "unmodified reference libska_sdp_func.so (sm_100 cubins) through "
                  "its own ska_sdp_func.grid_data.GridderUvwEsFft, device-resident "
                  "inputs, CUDA events on the legacy default stream"}
    if keep:
        out["dirty"] = img.cpu().numpy()
        out["vis"] = vis.cpu().numpy()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--direction", default="grid", choices=["grid", "degrid"])
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--single", type=int, default=-1,
                    help="C1 only: 1 = fp32 inputs, 0 = fp64")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    sys.path.insert(0, ROOT)
    import numpy as np
    from tests import synth
    import bench  # workload table only; imports neither library
    wl = bench.WORKLOADS[args.workload]
    case = wl["make"](0)
    if args.workload == "C1" and args.single >= 0:
        case = synth.reference_test_case(256, single=bool(args.single))
    res = run(case, wl["eps"], wl["wstack"], args.direction, args.steps,
              args.warmup, keep=bool(args.out))
    if args.out and res.get("available"):
        np.savez(args.out, dirty=res.pop("dirty"), vis=res.pop("vis"))
    res.pop("dirty", None)
    res.pop("vis", None)
    res["workload"] = args.workload
    print(json.dumps(res))


if __name__ == "__main__":
    main()
