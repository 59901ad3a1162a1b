#!/usr/bin/env python
"""Device times of the tiled functions (count + prefix sum, bucket sort,
indexing) on device-resident arrays, ours and -- in its own process -- the
unmodified reference CUDA build (baseline/ref_tiled.py), same seeded case.
  python profiles/run_tiled.py [--single 1] [--steps 5]
Prints one JSON line."""
import argparse
import ctypes
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "ska-sdp-func_b200")):
    sys.path.insert(0, p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--single", type=int, default=0)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--baselines", type=int, default=32768)
    ap.add_argument("--chan", type=int, default=512)
    ap.add_argument("--grid", type=int, default=8192)
    ap.add_argument("--no-ref", action="store_true")
    args = ap.parse_args()
    import numpy as np
    import torch
    from tests.synth_tiled import geometry_args, num_tiles, tiled_case
    from ska_sdp_func.visibility import (bucket_sort, count_and_prefix_sum,
                                         tiled_indexing)
    kw = dict(seed=41, single=bool(args.single), num_times=1,
              num_baselines=args.baselines, num_chan=args.chan,
              grid_size=args.grid, core=0.5, spread_v=0.45)
    c = tiled_case(**kw)
    dev = torch.device("cuda", 0)
    t = {k: torch.from_numpy(c[k]).to(dev) for k in ("uvw", "freqs", "vis", "weights")}
    nt = num_tiles(c)
    real = t["uvw"].dtype
    offsets = torch.zeros(nt + 1, dtype=torch.int32, device=dev)
    counts = torch.zeros(nt, dtype=torch.int32, device=dev)
    skipped = torch.zeros(1, dtype=torch.int32, device=dev)
    n = ctypes.c_int(0)

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

    ms_count = timed(lambda: count_and_prefix_sum(
        t["uvw"], t["freqs"], t["vis"], *geometry_args(c), n, offsets, counts,
        skipped))
    m = n.value
    s = {k: torch.zeros(m, dtype=real, device=dev) for k in ("uu", "vv", "vis", "weight")}
    s["tile"] = torch.zeros(m, dtype=torch.int32, device=dev)
    s["index"] = torch.zeros(m, dtype=torch.int32, device=dev)

    def sort():
        offs = offsets.clone()
        bucket_sort(t["uvw"], t["freqs"], t["vis"], t["weights"], *geometry_args(c),
                    s["uu"], s["vv"], s["weight"], s["tile"], s["vis"], offs)

    def index():
        offs = offsets.clone()
        tiled_indexing(t["uvw"], t["freqs"], *geometry_args(c), args.chan,
                       args.baselines, 1, s["tile"], s["uu"], s["vv"], s["index"],
                       offs)

    ms_sort = timed(sort)
    ms_index = timed(index)
    size = 4 if args.single else 8
    n_vis = args.baselines * args.chan
    out = dict(case=kw, num_vis=n_vis, num_pairs=m, num_tiles=nt,
               skipped=int(skipped.cpu()[0]),
               ours=dict(ms_count=ms_count, ms_sort=ms_sort, ms_index=ms_index,
                         Mvis_per_s_count=n_vis / ms_count / 1e3,
                         Mpairs_per_s_sort=m / ms_sort / 1e3,
                         # outputs written once + vis / weight read once
                         sort_output_GBps=m * (6 * size + 4) / ms_sort / 1e6))
    if not args.no_ref:
        r = subprocess.run(
            [sys.executable, os.path.join(ROOT, "baseline", "ref_tiled.py"),
             "--case", json.dumps(kw), "--steps", str(args.steps)],
            capture_output=True, text=True, timeout=900)
        lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
        out["reference_cuda"] = json.loads(lines[-1]) if lines else dict(
            available=False, why=(r.stdout + r.stderr)[-400:])
        ref = out["reference_cuda"]
        if ref.get("available"):
            out["speedup"] = dict(count=ref["ms_count"] / ms_count,
                                  sort=ref["ms_sort"] / ms_sort,
                                  index=ref["ms_index"] / ms_index)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
