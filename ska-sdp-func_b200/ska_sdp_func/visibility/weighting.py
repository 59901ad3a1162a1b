"""Uniform / Briggs visibility weighting on the uv grid.

Same functions, argument order and meaning as the reference's
``src/ska_sdp_func/visibility/weighting.py:44-176`` (``get_uv_range``,
``briggs_weights``, ``uniform_weights``), bound to ``sdp_weighting_briggs`` /
``sdp_weighting_uniform`` of this library. Arrays may be numpy arrays (staged
through the GPU inside the call) or any ``__cuda_array_interface__`` object
(processed in place); all arrays of one call must live in the same place.
"""

import ctypes

import numpy

from ..utility import Lib, Mem

Lib.wrap_func(
    "sdp_weighting_briggs",
    restype=None,
    argtypes=[
        Mem.handle_type(),
        Mem.handle_type(),
        ctypes.c_double,
        ctypes.c_double,
        Mem.handle_type(),
        Mem.handle_type(),
        Mem.handle_type(),
    ],
    check_errcode=True,
)

Lib.wrap_func(
    "sdp_weighting_uniform",
    restype=None,
    argtypes=[
        Mem.handle_type(),
        Mem.handle_type(),
        ctypes.c_double,
        Mem.handle_type(),
        Mem.handle_type(),
        Mem.handle_type(),
    ],
    check_errcode=True,
)

Lib.wrap_func(
    "sdp_weighting_briggs_no_alloc",
    restype=None,
    argtypes=[
        Mem.handle_type(),
        Mem.handle_type(),
        ctypes.c_double,
        ctypes.c_double,
        ctypes.c_int64,
        Mem.handle_type(),
        Mem.handle_type(),
    ],
    check_errcode=True,
)


def get_uv_range(uvw, freq_hz):
    """Largest |u| in wavelengths at the last channel (the reference looks at
    the u column only, weighting.py:59-60).

    :param uvw: [num_times, num_baselines, 3], metres.
    :param freq_hz: [num_channels], Hz, ascending.
    """
    max_abs_uv = numpy.amax(numpy.abs(uvw[:, :, 0:1]))
    max_abs_uv *= freq_hz[-1] / 299792458.0
    return max_abs_uv


def briggs_weights(uvw, freq_hz, max_abs_uv, robust_param, grid_uv,
                   input_weights, output_weights):
    """Briggs (robust) weights.

    ``grid_uv`` [grid_size, grid_size, num_pols] must be zero on entry and
    holds the summed input weights per uv cell afterwards;
    ``output_weights = input_weights / (1 + f2 * grid_uv[cell])`` with
    ``f2 = (5 / 10**robust_param)**2 / (sum grid[cell]**2 / sum grid[cell])``.

    :param uvw: [num_times, num_baselines, 3], metres, float64.
    :param freq_hz: [num_channels], Hz, float64.
    :param max_abs_uv: uv extent of the grid in wavelengths.
    :param robust_param: -2 (close to uniform) ... 2 (close to natural).
    :param grid_uv: [grid_size, grid_size, num_pols], float64 or float32.
    :param input_weights: [num_times, num_baselines, num_channels, num_pols].
    :param output_weights: same shape and type as ``input_weights``.
    """
    Lib.sdp_weighting_briggs(
        Mem(uvw),
        Mem(freq_hz),
        max_abs_uv,
        robust_param,
        Mem(grid_uv),
        Mem(input_weights),
        Mem(output_weights),
    )


def uniform_weights(uvw, freq_hz, max_abs_uv, grid_uv, input_weights,
                    output_weights):
    """Uniform weights: ``output_weights = 1 / grid_uv[cell]`` where
    ``grid_uv`` (zero on entry) receives the summed input weights per uv cell.
    Arguments as for :func:`briggs_weights`."""
    Lib.sdp_weighting_uniform(
        Mem(uvw),
        Mem(freq_hz),
        max_abs_uv,
        Mem(grid_uv),
        Mem(input_weights),
        Mem(output_weights),
    )


def briggs_weights_no_alloc(uvw, freq_hz, max_abs_uv, robust_param, grid_size,
                            input_weights, output_weights):
    """:func:`briggs_weights` without any device allocation (no counterpart in
    the reference): the uv weight grid is kept in a static device array of the
    library instead of ``grid_uv``. Device arrays only;
    ``grid_size**2 * num_pols`` must not exceed 1,048,576; calls that overlap
    on the GPU race on the shared array."""
    Lib.sdp_weighting_briggs_no_alloc(
        Mem(uvw),
        Mem(freq_hz),
        max_abs_uv,
        robust_param,
        grid_size,
        Mem(input_weights),
        Mem(output_weights),
    )
