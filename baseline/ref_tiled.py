#!/usr/bin/env python
"""Drives the UNMODIFIED reference (baseline/_ref, built for sm_100 by
baseline/build_ref_gpu.sh) through its own Python API of the tiled functions,
``ska_sdp_func.visibility.tiled_functions``, on device arrays -- the
reference's CUDA kernels on this GPU. Harness only; runs in its OWN process
(the reference library and the product library share file name and symbols).

  python baseline/ref_tiled.py --case '{"seed": 11}' --out x.npz [--steps 3]
stores num / offsets / counts / skipped / sort_* / index_* and prints one JSON
line {"available": true, "ms_count": .., "ms_sort": .., "ms_index": ..}.
"""
import argparse
import ctypes
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
    from ska_sdp_func.visibility.tiled_functions import (
        bucket_sort, count_and_prefix_sum, tiled_indexing)
    assert os.path.realpath(ska_sdp_func.__file__).startswith(os.path.realpath(REF))
    sys.path.insert(0, ROOT)
    from tests.synth_tiled import geometry_args, num_tiles, tiled_case

    c = tiled_case(**json.loads(args.case))
    dev = torch.device("cuda", 0)
    nt = num_tiles(c)
    t = {k: torch.from_numpy(c[k]).to(dev) for k in ("uvw", "freqs", "vis", "weights")}
    d = {k: cupy.asarray(v) for k, v in t.items()}
    real = t["uvw"].dtype

    def timed(fn):
        best = None
        for _ in range(args.steps):
            torch.cuda.synchronize()
            e0, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
            e0.record()
            out = fn()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
            best = ms if best is None else min(best, ms)
        return out, best

    def count():
        offsets = torch.zeros(nt + 1, dtype=torch.int32, device=dev)
        counts = torch.zeros(nt, dtype=torch.int32, device=dev)
        skipped = torch.zeros(1, dtype=torch.int32, device=dev)
        n = ctypes.c_int(0)
        count_and_prefix_sum(d["uvw"], d["freqs"], d["vis"], *geometry_args(c),
                             n, cupy.asarray(offsets), cupy.asarray(counts),
                             cupy.asarray(skipped))
        return n.value, offsets, counts, skipped

    (m, offsets, counts, skipped), ms_count = timed(count)

    def sort():
        s = {k: torch.zeros(m, dtype=real, device=dev)
             for k in ("uu", "vv", "vis", "weight")}
        s["tile"] = torch.zeros(m, dtype=torch.int32, device=dev)
        s["offsets"] = offsets.clone()
        bucket_sort(d["uvw"], d["freqs"], d["vis"], d["weights"],
                    *geometry_args(c), cupy.asarray(s["uu"]), cupy.asarray(s["vv"]),
                    cupy.asarray(s["weight"]), cupy.asarray(s["tile"]),
                    cupy.asarray(s["vis"]), cupy.asarray(s["offsets"]))
        return s

    def index():
        s = {k: torch.zeros(m, dtype=real, device=dev) for k in ("uu", "vv")}
        s["tile"] = torch.zeros(m, dtype=torch.int32, device=dev)
        s["index"] = torch.zeros(m, dtype=torch.int32, device=dev)
        s["offsets"] = offsets.clone()
        tiled_indexing(d["uvw"], d["freqs"], *geometry_args(c),
                       c["freqs"].shape[0], c["uvw"].shape[1], c["uvw"].shape[0],
                       cupy.asarray(s["tile"]), cupy.asarray(s["uu"]),
                       cupy.asarray(s["vv"]), cupy.asarray(s["index"]),
                       cupy.asarray(s["offsets"]))
        return s

    s, ms_sort = timed(sort)
    x, ms_index = timed(index)
    if args.out:
        out = dict(num=np.int64(m), offsets=offsets.cpu().numpy(),
                   counts=counts.cpu().numpy(), skipped=skipped.cpu().numpy())
        out.update({f"sort_{k}": v.cpu().numpy() for k, v in s.items()})
        out.update({f"index_{k}": v.cpu().numpy() for k, v in x.items()})
        np.savez(args.out, **out)
    print(json.dumps({"available": True, "num_visibilities": m,
                      "ms_count": ms_count, "ms_sort": ms_sort,
                      "ms_index": ms_index}))


if __name__ == "__main__":
    main()
