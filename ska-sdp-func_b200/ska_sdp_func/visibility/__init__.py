"""Visibility-domain functions (same module layout as the reference's
``ska_sdp_func.visibility``): uv-grid weighting, the tiled presort, the
point-source DFT predict."""
from .dft import dft_point_v00, dft_point_v01
from .tiled_functions import bucket_sort, count_and_prefix_sum, tiled_indexing
from .weighting import (briggs_weights, briggs_weights_no_alloc, get_uv_range,
                        uniform_weights)

__all__ = ["briggs_weights", "briggs_weights_no_alloc", "bucket_sort",
           "count_and_prefix_sum", "dft_point_v00", "dft_point_v01",
           "get_uv_range", "tiled_indexing", "uniform_weights"]
