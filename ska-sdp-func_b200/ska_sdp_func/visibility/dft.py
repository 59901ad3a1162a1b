"""Point-source DFT predict.

Same functions, argument order and meaning as the reference's
``src/ska_sdp_func/visibility/dft.py:41-146`` (``dft_point_v00``,
``dft_point_v01``), bound to ``sdp_dft_point_v00`` / ``sdp_dft_point_v01`` of
this library. Arrays may be numpy arrays (staged through the GPU inside the
call) or any ``__cuda_array_interface__`` object (processed in place); all
arrays of one call must live in the same place.
"""

import ctypes

from ..utility import Lib, Mem

Lib.wrap_func(
    "sdp_dft_point_v00",
    restype=None,
    argtypes=[Mem.handle_type()] * 4,
    check_errcode=True,
)

Lib.wrap_func(
    "sdp_dft_point_v01",
    restype=None,
    argtypes=[Mem.handle_type()] * 3
    + [ctypes.c_double, ctypes.c_double, Mem.handle_type()],
    check_errcode=True,
)


def dft_point_v00(source_directions, source_fluxes, uvw_lambda, vis):
    """Visibilities of point sources by direct summation (deprecated layout:
    coordinates in wavelengths per channel).

    :param source_directions: direction cosines ``[num_components, 3]``, fp64
    :param source_fluxes: complex fluxes
        ``[num_components, num_channels, num_pols]``, complex fp64
    :param uvw_lambda: coordinates in wavelengths
        ``[num_times, num_baselines, num_channels, 3]``, fp64
    :param vis: output ``[num_times, num_baselines, num_channels, num_pols]``,
        complex fp64 or fp32 (overwritten)
    """
    Lib.sdp_dft_point_v00(
        Mem(source_directions), Mem(source_fluxes), Mem(uvw_lambda), Mem(vis)
    )


def dft_point_v01(
    source_directions,
    source_fluxes,
    uvw,
    channel_start_hz,
    channel_step_hz,
    vis,
):
    """Visibilities of point sources by direct summation.

    :param source_directions: direction cosines ``[num_components, 3]``, fp64
    :param source_fluxes: complex fluxes
        ``[num_components, num_channels, num_pols]``, complex fp64
    :param uvw: coordinates in metres ``[num_times, num_baselines, 3]``, fp64
    :param channel_start_hz: frequency of the first channel, Hz
    :param channel_step_hz: frequency increment between channels, Hz
    :param vis: output ``[num_times, num_baselines, num_channels, num_pols]``,
        complex fp64 or fp32 (overwritten)
    """
    Lib.sdp_dft_point_v01(
        Mem(source_directions),
        Mem(source_fluxes),
        Mem(uvw),
        channel_start_hz,
        channel_step_hz,
        Mem(vis),
    )
