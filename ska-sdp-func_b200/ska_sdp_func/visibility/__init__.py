"""Visibility-domain functions (same module layout as the reference's
``ska_sdp_func.visibility``): uv-grid weighting."""
from .weighting import (briggs_weights, briggs_weights_no_alloc, get_uv_range,
                        uniform_weights)

__all__ = ["briggs_weights", "briggs_weights_no_alloc", "get_uv_range",
           "uniform_weights"]
