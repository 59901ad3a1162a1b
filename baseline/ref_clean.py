#!/usr/bin/env python
"""Drives the UNMODIFIED reference (baseline/_ref, built for sm_100 by
baseline/build_ref_gpu.sh) through its own Python API of the Hogbom CLEAN,
``ska_sdp_func.clean.hogbom_clean``, on device arrays -- the reference's CUDA
path on this GPU. Harness only; runs in its OWN process.

  python baseline/ref_clean.py --case '{"seed": 1, "n": 256}' --cycles 1000 [--out x.npz]
prints one JSON line {"available": true, "ms": ..}.
"""
import argparse
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.path.join(HERE, "_ref")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case", default="{}")
    ap.add_argument("--cycles", type=int, default=1000)
    ap.add_argument("--threshold", type=float, default=None)
    ap.add_argument("--out", default="")
    ap.add_argument("--steps", type=int, default=1)
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
    from ska_sdp_func.clean import hogbom_clean
    assert os.path.realpath(ska_sdp_func.__file__).startswith(os.path.realpath(REF))
    sys.path.insert(0, ROOT)
    from tests.synth_clean import clean_case

    c = clean_case(**json.loads(args.case))
    threshold = c["threshold"] if args.threshold is None else args.threshold
    dev = torch.device("cuda", 0)
    dirty = torch.from_numpy(c["dirty"]).to(dev)
    psf = torch.from_numpy(c["psf"]).to(dev)
    best = None
    for _ in range(args.steps):
        out = [torch.zeros_like(dirty) for _ in range(3)]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        hogbom_clean(cupy.asarray(dirty), cupy.asarray(psf), c["cbeam_details"],
                     c["loop_gain"], threshold, args.cycles,
                     cupy.asarray(out[0]), cupy.asarray(out[1]),
                     cupy.asarray(out[2]))
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3
        best = ms if best is None else min(best, ms)
    if args.out:
        np.savez(args.out, model=out[0].cpu().numpy(),
                 residual=out[1].cpu().numpy(), skymodel=out[2].cpu().numpy())
    print(json.dumps({"available": True, "ms": best,
                      "components": int((out[0] != 0).sum())}))


if __name__ == "__main__":
    main()
