#!/usr/bin/env python
"""Wall time of one Hogbom CLEAN call on device-resident images (the call
synchronises: host timer around it), ours and -- in its own process -- the
unmodified reference CUDA build (baseline/ref_clean.py), same seeded case;
parity of ours against the oracle.
  python profiles/run_clean.py [--n 256] [--cycles 2000] [--single 0]
Prints one JSON line."""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "ska-sdp-func_b200")):
    sys.path.insert(0, p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=256)
    ap.add_argument("--cycles", type=int, default=2000)
    ap.add_argument("--single", type=int, default=0)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--no-ref", action="store_true")
    ap.add_argument("--no-oracle", action="store_true")
    args = ap.parse_args()
    import numpy as np
    import torch
    from oracle import clean_oracle
    from tests.synth_clean import clean_case
    from ska_sdp_func.clean import hogbom_clean
    kw = dict(seed=61, n=args.n, num_sources=30, single=bool(args.single),
              beam=(2.0, 2.0, 1.0, 64.0))
    c = clean_case(**kw)
    threshold = -1.0      # every cycle runs: equal work on both sides
    dev = torch.device("cuda", 0)
    dirty = torch.from_numpy(c["dirty"]).to(dev)
    psf = torch.from_numpy(c["psf"]).to(dev)
    best = None
    for _ in range(args.steps + 1):
        out = [torch.zeros_like(dirty) for _ in range(3)]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        hogbom_clean(dirty, psf, c["cbeam_details"], c["loop_gain"], threshold,
                     args.cycles, out[0], out[1], out[2])
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3
        best = ms if best is None else min(best, ms)
    res = dict(case=kw, cycles=args.cycles, ours=dict(
        ms=best, us_per_cycle=best * 1e3 / args.cycles,
        components=int((out[0] != 0).sum())))
    if not args.no_oracle:
        want = clean_oracle.hogbom_clean(c["dirty"], c["psf"], c["cbeam_details"],
                                         c["loop_gain"], threshold, args.cycles)
        sky = out[2].cpu().numpy().astype(np.float64)
        res["parity"] = dict(
            model_identical=bool(np.array_equal(out[0].cpu().numpy(),
                                                want["clean_model"])),
            residual_identical=bool(np.array_equal(out[1].cpu().numpy(),
                                                   want["residual"])),
            skymodel=float(np.abs(sky - want["skymodel"]).max()
                           / np.abs(want["skymodel"]).max()))
    if not args.no_ref:
        r = subprocess.run(
            [sys.executable, os.path.join(ROOT, "baseline", "ref_clean.py"),
             "--case", json.dumps(kw), "--cycles", str(args.cycles),
             "--threshold", str(threshold), "--steps", str(args.steps)],
            capture_output=True, text=True, timeout=1200)
        lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
        ref = json.loads(lines[-1]) if lines else dict(
            available=False, why=(r.stdout + r.stderr)[-400:])
        res["reference_cuda"] = ref
        if ref.get("available"):
            res["speedup"] = ref["ms"] / best
    print(json.dumps(res))


if __name__ == "__main__":
    main()
