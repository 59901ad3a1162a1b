#!/usr/bin/env python
"""Generates tests/golden/golden_tiled.npz: outputs of the UNMODIFIED reference
library's CPU path of sdp_count_and_prefix_sum / sdp_bucket_sort /
sdp_tiled_indexing (baseline/_ref, built by baseline/build_ref_gpu.sh from
/root/reference; its host loops need no GPU) on seeded inputs. Run in the
authoring container:
    python tests/golden/make_golden_tiled.py
The reference's host branch counts in fp64 only (sdp_tiled_functions.cpp:367)
and passes tile_v_min where tile_u_min belongs (:84-89), so these vectors pin
oracle/tiled_oracle.py in its `host_branch_slip=True` mode. The fp32 bucket
sort / indexing case has no reference count to start from (fp32 positions can
round to another cell than fp64 ones): its input offsets come from the oracle
and are stored with the outputs, which then show whether they were right.
"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = os.path.join(ROOT, "baseline", "_ref")
sys.path.insert(0, REF)
os.environ["SKA_SDP_FUNC_LIB_DIR"] = REF

from ska_sdp_func.visibility.tiled_functions import (  # noqa: E402
    bucket_sort, count_and_prefix_sum, tiled_indexing)

sys.path.insert(0, ROOT)
from tests.synth_tiled import geometry_args, num_tiles, tiled_case  # noqa: E402
from oracle import tiled_oracle  # noqa: E402

CASES = {
    "small_f64": dict(seed=1, positive=True),
    "centre_f64": dict(seed=2, positive=True, core=0.6, num_chan=4),
    "tile10_f64": dict(seed=3, positive=True, tile_size_u=10, tile_size_v=10,
                       grid_size=200, support=3, spread=0.4),
    "small_f32": dict(seed=4, positive=True, single=True),
}


def main():
    out = {}
    for name, kw in CASES.items():
        c = tiled_case(**kw)
        nt = num_tiles(c)
        offsets = np.zeros(nt + 1, dtype=np.int32)
        if not kw.get("single"):
            counts = np.zeros(nt, dtype=np.int32)
            skipped = np.zeros(1, dtype=np.int32)
            n = ctypes.c_int(0)
            count_and_prefix_sum(c["uvw"], c["freqs"], c["vis"],
                                 *geometry_args(c), n, offsets, counts, skipped)
            out[f"{name}_num"] = np.int64(n.value)
            out[f"{name}_offsets"] = offsets.copy()
            out[f"{name}_counts"] = counts.copy()
            out[f"{name}_skipped"] = skipped.copy()
            m = n.value
        else:
            # the host count exists in fp64 only (.cpp:367)
            m, offsets, _, _ = tiled_oracle.count_and_prefix_sum(
                c["uvw"], c["freqs"], *geometry_args(c), host_branch_slip=True)
        real = c["uvw"].dtype
        s = {k: np.zeros(m, dtype=real) for k in ("uu", "vv", "vis", "weight")}
        s["tile"] = np.zeros(m, dtype=np.int32)
        offs = offsets.copy()
        if kw.get("single"):
            out[f"{name}_offsets_in"] = offsets.copy()
        bucket_sort(c["uvw"], c["freqs"], c["vis"], c["weights"],
                    *geometry_args(c), s["uu"], s["vv"], s["weight"], s["tile"],
                    s["vis"], offs)
        for k, v in s.items():
            out[f"{name}_sort_{k}"] = v
        out[f"{name}_sort_offsets"] = offs
        t = {k: np.zeros(m, dtype=real) for k in ("uu", "vv")}
        t["tile"] = np.zeros(m, dtype=np.int32)
        t["index"] = np.zeros(m, dtype=np.int32)
        offs = offsets.copy()
        tiled_indexing(c["uvw"], c["freqs"], *geometry_args(c),
                       c["freqs"].shape[0],
                       c["uvw"].shape[1], c["uvw"].shape[0],
                       t["tile"], t["uu"], t["vv"], t["index"], offs)
        for k, v in t.items():
            out[f"{name}_index_{k}"] = v
        out[f"{name}_index_offsets"] = offs
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "golden_tiled.npz"),
                        **out)
    print("wrote", len(out), "arrays;",
          {k: int(v) for k, v in out.items() if k.endswith("_num")})


if __name__ == "__main__":
    main()
