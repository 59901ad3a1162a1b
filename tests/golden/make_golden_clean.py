#!/usr/bin/env python
"""Generates tests/golden/golden_clean.npz: outputs of the UNMODIFIED reference
library's CPU path of sdp_hogbom_clean (baseline/_ref, built by
baseline/build_ref_gpu.sh from /root/reference; its host loop needs no GPU)
on seeded inputs. Run in the authoring container:
    python tests/golden/make_golden_clean.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = os.path.join(ROOT, "baseline", "_ref")
sys.path.insert(0, REF)
os.environ["SKA_SDP_FUNC_LIB_DIR"] = REF

from ska_sdp_func.clean import hogbom_clean  # noqa: E402

sys.path.insert(0, ROOT)
from tests.synth_clean import clean_case  # noqa: E402

CASES = {
    "f64": dict(seed=1, cycles=300),
    "f32": dict(seed=2, single=True, cycles=300),
    "f64_limit": dict(seed=3, n=48, beam=(1.5, 1.5, 0.0, 10.0), cycles=25,
                      threshold=-1.0),
}


def main():
    out = {}
    for name, kw in CASES.items():
        kw = dict(kw)
        cycles = kw.pop("cycles")
        threshold = kw.pop("threshold", None)
        c = clean_case(**kw)
        if threshold is not None:
            c["threshold"] = threshold
        model, residual, sky = (np.zeros_like(c["dirty"]) for _ in range(3))
        hogbom_clean(c["dirty"], c["psf"], c["cbeam_details"], c["loop_gain"],
                     c["threshold"], cycles, model, residual, sky)
        out[f"{name}_model"] = model
        out[f"{name}_residual"] = residual
        out[f"{name}_skymodel"] = sky
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "golden_clean.npz"),
                        **out)
    print("wrote", len(out), "arrays",
          {k: int(np.count_nonzero(v)) for k, v in out.items() if "model" in k})


if __name__ == "__main__":
    main()
