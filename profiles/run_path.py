#!/usr/bin/env python
"""Small driver for profiling: runs N grid (or degrid) calls of a bench
workload with device-resident inputs and nothing else, so that ncu launch
lists map 1:1 onto the hot path.  python profiles/run_path.py C2 grid 2"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "ska-sdp-func_b200"))

import torch  # noqa: E402

import bench  # noqa: E402
from ska_sdp_func.grid_data import GridderUvwEsFft  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "C2"
    direction = sys.argv[2] if len(sys.argv) > 2 else "grid"
    calls = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    wl = bench.WORKLOADS[name]
    case = wl["make"](0)
    dev = torch.device("cuda", 0)
    n = case["image_size"]
    t = {k: torch.from_numpy(case[k]).to(dev) for k in ("uvw", "freq_hz", "vis", "weight")}
    img = torch.zeros((n, n), dtype=t["uvw"].dtype, device=dev)
    if direction == "degrid":
        img.uniform_(-0.5, 0.5)
    g = GridderUvwEsFft(t["uvw"], t["freq_hz"], t["vis"], t["weight"], img,
                        case["pixel_size"], case["pixel_size"], wl["eps"], wl["wstack"])
    g.enable_timing(True)
    for _ in range(calls):
        if direction == "grid":
            img.zero_()
            g.grid_uvw_es_fft(t["uvw"], t["freq_hz"], t["vis"], t["weight"], img)
        else:
            t["vis"].zero_()
            g.ifft_grid_uvw_es(t["uvw"], t["freq_hz"], t["vis"], t["weight"], img)
        torch.cuda.synchronize()
        print({k: (round(v, 3) if isinstance(v, float) else v)
               for k, v in g.stats().items()})
    print(g.info())


if __name__ == "__main__":
    main()
