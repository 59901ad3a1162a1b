#!/usr/bin/env python
"""Device times of the point-source DFT predict on device-resident arrays,
ours and -- in its own process -- the unmodified reference CUDA build
(baseline/ref_dft.py), same seeded case.
  python profiles/run_dft.py [--pols 4] [--steps 3]
Prints one JSON line (ms, G (visibility, source) terms per second, parity of
ours against the fp64 oracle on a sample of rows)."""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "ska-sdp-func_b200")):
    sys.path.insert(0, p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pols", type=int, default=4)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--components", type=int, default=1000)
    ap.add_argument("--baselines", type=int, default=8128)
    ap.add_argument("--times", type=int, default=4)
    ap.add_argument("--chan", type=int, default=64)
    ap.add_argument("--no-ref", action="store_true")
    args = ap.parse_args()
    import numpy as np
    import torch
    from oracle import dft_oracle
    from tests.synth_dft import dft_case
    from ska_sdp_func.visibility import dft_point_v00, dft_point_v01
    kw = dict(seed=51, num_components=args.components, num_pols=args.pols,
              num_channels=args.chan, num_baselines=args.baselines,
              num_times=args.times)
    c = dft_case(**kw)
    dev = torch.device("cuda", 0)
    t = {k: torch.from_numpy(c[k]).to(dev)
         for k in ("directions", "fluxes", "uvw", "uvw_lambda")}

    def timed(fn):
        best = None
        for _ in range(args.steps + 1):
            torch.cuda.synchronize()
            e0, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            best = ms if best is None else min(best, ms)
        return best

    terms = float(np.prod(c["vis_shape"][:3])) * args.components
    out = dict(case=kw, num_vis=int(np.prod(c["vis_shape"][:3])), terms=terms,
               ours={}, parity={})
    sample = slice(0, 64)
    cs = dict(c, uvw=c["uvw"][:1, sample], uvw_lambda=c["uvw_lambda"][:1, sample])
    want01 = dft_oracle.dft_point_v01(cs["directions"], cs["fluxes"], cs["uvw"],
                                      c["channel_start_hz"], c["channel_step_hz"],
                                      exact=True)
    want00 = dft_oracle.dft_point_v00(cs["directions"], cs["fluxes"],
                                      cs["uvw_lambda"], exact=True)
    for dtype, tag in ((torch.complex128, "c128"), (torch.complex64, "c64")):
        vis = torch.zeros(c["vis_shape"], dtype=dtype, device=dev)
        ms = timed(lambda: dft_point_v01(t["directions"], t["fluxes"], t["uvw"],
                                         c["channel_start_hz"],
                                         c["channel_step_hz"], vis))
        out["ours"][f"ms_v01_{tag}"] = ms
        out["ours"][f"Gterms_per_s_v01_{tag}"] = terms / ms / 1e6
        got = vis[:1, sample].cpu().numpy().astype(np.complex128)
        out["parity"][f"v01_{tag}"] = float(np.abs(got - want01).max()
                                            / np.abs(want01).max())
        vis.zero_()
        ms = timed(lambda: dft_point_v00(t["directions"], t["fluxes"],
                                         t["uvw_lambda"], vis))
        out["ours"][f"ms_v00_{tag}"] = ms
        out["ours"][f"Gterms_per_s_v00_{tag}"] = terms / ms / 1e6
        got = vis[:1, sample].cpu().numpy().astype(np.complex128)
        out["parity"][f"v00_{tag}"] = float(np.abs(got - want00).max()
                                            / np.abs(want00).max())
    if not args.no_ref:
        r = subprocess.run(
            [sys.executable, os.path.join(ROOT, "baseline", "ref_dft.py"),
             "--case", json.dumps(kw), "--steps", str(args.steps)],
            capture_output=True, text=True, timeout=900)
        lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
        ref = json.loads(lines[-1]) if lines else dict(
            available=False, why=(r.stdout + r.stderr)[-400:])
        out["reference_cuda"] = ref
        if ref.get("available"):
            out["speedup"] = {k[3:]: ref[k] / out["ours"][k]
                              for k in ref if k.startswith("ms_")}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
