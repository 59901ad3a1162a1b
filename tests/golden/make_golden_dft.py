#!/usr/bin/env python
"""Generates tests/golden/golden_dft.npz: outputs of the UNMODIFIED reference
library's CPU path of sdp_dft_point_v00 / sdp_dft_point_v01 (baseline/_ref,
built by baseline/build_ref_gpu.sh from /root/reference; its OpenMP loops need
no GPU) on seeded inputs. Run in the authoring container:
    python tests/golden/make_golden_dft.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = os.path.join(ROOT, "baseline", "_ref")
sys.path.insert(0, REF)
os.environ["SKA_SDP_FUNC_LIB_DIR"] = REF

from ska_sdp_func.visibility import dft_point_v00, dft_point_v01  # noqa: E402

sys.path.insert(0, ROOT)
from tests.synth_dft import dft_case  # noqa: E402

CASES = {
    "pols4": dict(seed=1, num_baselines=30, num_times=2),
    "pols1": dict(seed=2, num_pols=1, num_channels=13, num_components=33,
                  num_baselines=40, num_times=2),
}


def main():
    out = {}
    for name, kw in CASES.items():
        c = dft_case(**kw)
        for dtype, tag in ((np.complex128, "c128"), (np.complex64, "c64")):
            vis = np.zeros(c["vis_shape"], dtype=dtype)
            dft_point_v00(c["directions"], c["fluxes"], c["uvw_lambda"], vis)
            out[f"{name}_v00_{tag}"] = vis
            vis = np.zeros(c["vis_shape"], dtype=dtype)
            dft_point_v01(c["directions"], c["fluxes"], c["uvw"],
                          c["channel_start_hz"], c["channel_step_hz"], vis)
            out[f"{name}_v01_{tag}"] = vis
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "golden_dft.npz"),
                        **out)
    print("wrote", len(out), "arrays",
          {k: float(np.abs(v).max()) for k, v in out.items()})


if __name__ == "__main__":
    main()
