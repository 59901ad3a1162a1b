"""Tile count + prefix sum, bucket sort and tiled indexing of visibilities.

Same functions, argument order and meaning as the reference's
``src/ska_sdp_func/visibility/tiled_functions.py:80-269``
(``count_and_prefix_sum``, ``bucket_sort``, ``tiled_indexing``), bound to
``sdp_count_and_prefix_sum`` / ``sdp_bucket_sort`` / ``sdp_tiled_indexing`` of
this library. Arrays may be numpy arrays (staged through the GPU inside the
call) or any ``__cuda_array_interface__`` object (processed in place); all
arrays of one call must live in the same place. Inside a tile the sorted
entries are in (time, baseline, channel) order on every run.
"""

import ctypes

from ..utility import Lib, Mem

Lib.wrap_func(
    "sdp_count_and_prefix_sum",
    restype=None,
    argtypes=[
        Mem.handle_type(),
        Mem.handle_type(),
        Mem.handle_type(),
        ctypes.c_int,
        ctypes.c_int64,
        ctypes.c_int64,
        ctypes.c_double,
        ctypes.c_int64,
        ctypes.POINTER(ctypes.c_int),
        Mem.handle_type(),
        Mem.handle_type(),
        Mem.handle_type(),
    ],
    check_errcode=True,
)

Lib.wrap_func(
    "sdp_bucket_sort",
    restype=None,
    argtypes=[Mem.handle_type()] * 4
    + [ctypes.c_int, ctypes.c_int64, ctypes.c_int64, ctypes.c_double,
       ctypes.c_int64]
    + [Mem.handle_type()] * 6,
    check_errcode=True,
)

Lib.wrap_func(
    "sdp_tiled_indexing",
    restype=None,
    argtypes=[Mem.handle_type()] * 2
    + [ctypes.c_int, ctypes.c_int64, ctypes.c_int64, ctypes.c_double]
    + [ctypes.c_int64] * 4
    + [Mem.handle_type()] * 5,
    check_errcode=True,
)


def count_and_prefix_sum(
    uvw,
    freqs,
    vis,
    grid_size,
    tile_size_u,
    tile_size_v,
    cell_size_rad,
    support,
    num_visibilities,
    tile_offsets,
    num_points_in_tiles,
    num_skipped,
):
    """Counts the (visibility, tile) pairs per uv tile and prefix-sums them.

    :param uvw: baseline coordinates in metres, ``[num_times, num_baselines, 3]``
    :param freqs: channel frequencies in Hz, ``[num_channels]``
    :param vis: complex visibilities,
        ``[num_times, num_baselines, num_channels, num_pols]`` (shape only)
    :param grid_size: cells per side of the (square) grid
    :param tile_size_u: tile size along u, in cells
    :param tile_size_v: tile size along v, in cells
    :param cell_size_rad: cell size in radians
    :param support: half-width of a visibility's footprint, in cells
    :param num_visibilities: ``ctypes.c_int`` that receives the number of
        pairs (the length the sorted arrays need)
    :param tile_offsets: int32 ``[num_tiles + 1]``, receives the prefix sum
    :param num_points_in_tiles: int32 ``[num_tiles]``, receives the counts
    :param num_skipped: int32 ``[1]``, receives the number of visibilities
        whose footprint leaves the grid
    """
    Lib.sdp_count_and_prefix_sum(
        Mem(uvw),
        Mem(freqs),
        Mem(vis),
        grid_size,
        tile_size_u,
        tile_size_v,
        cell_size_rad,
        support,
        ctypes.byref(num_visibilities),
        Mem(tile_offsets),
        Mem(num_points_in_tiles),
        Mem(num_skipped),
    )


def bucket_sort(
    uvw,
    freqs,
    vis,
    weights,
    grid_size,
    tile_size_u,
    tile_size_v,
    cell_size_rad,
    support,
    sorted_uu,
    sorted_vv,
    sorted_weight,
    sorted_tile,
    sorted_vis,
    tile_offsets,
):
    """Bucket-sorts positions, visibilities and weights by uv tile.

    Inputs and geometry as for :func:`count_and_prefix_sum`; ``weights`` has
    the shape of ``vis``. ``sorted_uu`` / ``sorted_vv`` (cells from the grid
    centre), ``sorted_weight``, ``sorted_vis`` (real: the reference's view of
    ``vis`` as reals; complex: the visibility itself) and ``sorted_tile``
    (``pv * 32768 + pu``) have ``num_visibilities`` elements.
    ``tile_offsets`` comes from :func:`count_and_prefix_sum` and is advanced
    by each tile's entries, as in the reference.
    """
    Lib.sdp_bucket_sort(
        Mem(uvw),
        Mem(freqs),
        Mem(vis),
        Mem(weights),
        grid_size,
        tile_size_u,
        tile_size_v,
        cell_size_rad,
        support,
        Mem(sorted_uu),
        Mem(sorted_vv),
        Mem(sorted_weight),
        Mem(sorted_tile),
        Mem(sorted_vis),
        Mem(tile_offsets),
    )


def tiled_indexing(
    uvw,
    freqs,
    grid_size,
    tile_size_u,
    tile_size_v,
    cell_size_rad,
    support,
    num_channels,
    num_baselines,
    num_times,
    sorted_tile,
    sorted_uu,
    sorted_vv,
    sorted_vis_index,
    tile_offsets,
):
    """Bucket-sorts visibility INDICES (flat over time, baseline, channel) by
    uv tile; other arguments as for :func:`bucket_sort`."""
    Lib.sdp_tiled_indexing(
        Mem(uvw),
        Mem(freqs),
        grid_size,
        tile_size_u,
        tile_size_v,
        cell_size_rad,
        support,
        num_channels,
        num_baselines,
        num_times,
        Mem(sorted_tile),
        Mem(sorted_uu),
        Mem(sorted_vv),
        Mem(sorted_vis_index),
        Mem(tile_offsets),
    )
