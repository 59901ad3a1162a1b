#!/usr/bin/env python
"""Generates tests/golden/golden_weighting.npz: outputs of the UNMODIFIED
reference library's CPU path of sdp_weighting_uniform / sdp_weighting_briggs
(baseline/_ref, built by baseline/build_ref_gpu.sh from /root/reference; its
host loops need no GPU) on seeded inputs. Run in the authoring container:
    python tests/golden/make_golden_weighting.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = os.path.join(ROOT, "baseline", "_ref")
sys.path.insert(0, REF)
os.environ["SKA_SDP_FUNC_LIB_DIR"] = REF

from ska_sdp_func.visibility.weighting import (  # noqa: E402
    briggs_weights, uniform_weights)

sys.path.insert(0, ROOT)
from tests.synth_weighting import weighting_case  # noqa: E402


def main():
    out = {}
    for name, kw in (("small_f64", dict(seed=1, single=False)),
                     ("small_f32", dict(seed=2, single=True)),
                     ("pols4_f64", dict(seed=3, single=False, num_pols=4,
                                        grid_size=64)),
                     ("offgrid_f64", dict(seed=4, single=False, shrink=0.6,
                                          positive=True))):
        c = weighting_case(**kw)
        for kind in ("uniform", "briggs"):
            grid = np.zeros_like(c["grid"])
            outw = c["output0"].copy()
            if kind == "uniform":
                uniform_weights(c["uvw"], c["freq_hz"], c["max_abs_uv"], grid,
                                c["input"], outw)
            else:
                briggs_weights(c["uvw"], c["freq_hz"], c["max_abs_uv"],
                               c["robust"], grid, c["input"], outw)
            out[f"{name}_{kind}_grid"] = grid
            out[f"{name}_{kind}_out"] = outw
    np.savez_compressed(os.path.join(ROOT, "tests", "golden",
                                     "golden_weighting.npz"), **out)
    print("wrote", len(out), "arrays")


if __name__ == "__main__":
    main()
