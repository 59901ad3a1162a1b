"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's tiled
visibility functions (tile count + prefix sum, bucket sort, tiled indexing),
the parity oracle of sdp_count_and_prefix_sum / sdp_bucket_sort /
sdp_tiled_indexing. Never imported by the product.

Follows /root/reference/src/ska-sdp-func/visibility/sdp_tiled_functions.cpp
and .cu operation by operation:
  tile geometry (centre tile, top-left offset)      .cpp:333-342  -> geometry
  grid position and on-grid test of a visibility    .cu:96-104 (= .cpp:64-77) -> positions
  tile range of a visibility (float arithmetic)     .cu:106-114 (= .cpp:11-19) -> tile_ranges
  count per tile + skipped count                    .cu:116-124   -> count_and_prefix_sum
  serial prefix sum                                 .cpp:106-122  -> count_and_prefix_sum
  bucket sort / indexing into the caller's offsets  .cu:191-207, :271-286 -> bucket_sort, tiled_indexing
Entry order inside a tile: the reference's GPU kernels take their slot with an
atomicAdd, so their order inside a tile is unspecified; this oracle (and the
product) emit in the order of the reference's HOST loops (.cpp:154-215: time,
baseline, channel; per visibility pv outer, pu inner).

`host_branch_slip=True` restates the reference's HOST branch literally: its
TILE_RANGES call passes tile_v_min where tile_u_min belongs (.cpp:84-89,
:189-194, :276-281), so tile_u_min stays 0 there. That mode exists only so
that this oracle can be PINNED on outputs of the unmodified reference
library's CPU path (tests/golden/golden_tiled.npz, made by
tests/golden/make_golden_tiled.py against baseline/_ref); the product follows
the GPU kernels (slip off), which is what the reference's own tests run.
"""
import numpy as np

C_0 = 299792458.0  # sdp_tiled_functions.cpp:21


def geometry(grid_size, tile_size_u, tile_size_v):
    """.cpp:333-342: tiles are laid out so that the grid centre is the middle
    of a tile; float32 reciprocals of the tile sizes."""
    centre = grid_size // 2
    ctile_u = centre // tile_size_u
    ctile_v = centre // tile_size_v
    return dict(
        centre=centre,
        inv_u=np.float32(1.0 / tile_size_u), inv_v=np.float32(1.0 / tile_size_v),
        num_tiles_u=(grid_size + tile_size_u - 1) // tile_size_u,
        num_tiles_v=(grid_size + tile_size_v - 1) // tile_size_v,
        top_left_u=centre - ctile_u * tile_size_u - tile_size_u // 2,
        top_left_v=centre - ctile_v * tile_size_v - tile_size_v // 2)


def _round_half_away(x):
    """C round(): halfway cases away from zero (np.rint rounds to even)."""
    return np.sign(x) * np.floor(np.abs(x) + x.dtype.type(0.5))


def positions(uvw, freqs, grid_size, cell_size_rad, support):
    """pos_u, pos_v in the precision of uvw (.cu:96-98: freq / c in double
    rounded to that type, u * that in that type, times the double grid scale,
    rounded to that type), grid_u / grid_v (:100-101) and the on-grid test
    (:103-104). Shapes [T, B, C]."""
    real = uvw.dtype.type
    grid_scale = np.float64(grid_size * cell_size_rad)
    inv_wavelength = (freqs.astype(np.float64) / C_0).astype(real)
    pos = []
    for k in (0, 1):
        p = uvw[:, :, k, None] * inv_wavelength[None, None, :]      # in `real`
        pos.append((p.astype(np.float64) * grid_scale).astype(real))
    centre = grid_size // 2
    grid_u = _round_half_away(pos[0]).astype(np.int64) + centre
    grid_v = _round_half_away(pos[1]).astype(np.int64) + centre
    inside = ((grid_u + support < grid_size) & (grid_u - support >= 0) &
              (grid_v + support < grid_size) & (grid_v - support >= 0))
    return pos[0], pos[1], grid_u, grid_v, inside


def tile_ranges(grid_u, grid_v, geo, support, host_branch_slip=False):
    """[u_min, u_max) x [v_min, v_max) in tiles, float32 products with the
    reciprocal tile size, floor / ceil (.cu:107-114)."""
    rel_u = (grid_u - geo["top_left_u"]).astype(np.int32).astype(np.int64)
    rel_v = (grid_v - geo["top_left_v"]).astype(np.int32).astype(np.int64)
    u1 = (rel_u - support).astype(np.float32) * geo["inv_u"]
    u2 = (rel_u + support + 1).astype(np.float32) * geo["inv_u"]
    v1 = (rel_v - support).astype(np.float32) * geo["inv_v"]
    v2 = (rel_v + support + 1).astype(np.float32) * geo["inv_v"]
    u_min = np.floor(u1).astype(np.int64)
    u_max = np.ceil(u2).astype(np.int64)
    v_min = np.floor(v1).astype(np.int64)
    v_max = np.ceil(v2).astype(np.int64)
    if host_branch_slip:
        u_min = np.zeros_like(u_min)
    return u_min, u_max, v_min, v_max


def entries(uvw, freqs, grid_size, tile_size_u, tile_size_v, cell_size_rad,
            support, host_branch_slip=False):
    """Every (visibility, tile) pair in the host loops' order. Returns
    (i_vis, tile index pu + pv * num_tiles_u, pu, pv, pos_u, pos_v) per entry,
    the number of skipped visibilities and the geometry. A tile column / row
    that does not exist (pu = num_tiles_u: the last half tile of a grid whose
    size is a multiple of the tile size) keeps the linear index the reference
    forms for it, i.e. it aliases onto a neighbouring tile, and its literal
    (pu, pv); only pairs whose linear index leaves [0, num_tiles) -- an
    out-of-bounds access in the reference -- are dropped."""
    geo = geometry(grid_size, tile_size_u, tile_size_v)
    pos_u, pos_v, grid_u, grid_v, inside = positions(
        uvw, freqs, grid_size, cell_size_rad, support)
    u_min, u_max, v_min, v_max = tile_ranges(grid_u, grid_v, geo, support,
                                             host_branch_slip)
    flat = np.flatnonzero(inside.ravel())           # i_vis = (t * B + b) * C + c
    u0, u1 = u_min.ravel()[flat], u_max.ravel()[flat]
    v0, v1 = v_min.ravel()[flat], v_max.ravel()[flat]
    nu = np.maximum(u1 - u0, 0)
    nv = np.maximum(v1 - v0, 0)
    n = nu * nv
    owner = np.repeat(np.arange(flat.size), n)
    first = np.cumsum(n) - n
    k = np.arange(owner.size) - first[owner]        # pv outer, pu inner
    pv = v0[owner] + k // np.maximum(nu[owner], 1)
    pu = u0[owner] + k % np.maximum(nu[owner], 1)
    tile = pu + pv * geo["num_tiles_u"]
    ok = (tile >= 0) & (tile < geo["num_tiles_u"] * geo["num_tiles_v"])
    i_vis = flat[owner][ok]
    return (i_vis, tile[ok], pu[ok], pv[ok], pos_u.ravel()[i_vis],
            pos_v.ravel()[i_vis], int(inside.size - flat.size), geo)


def count_and_prefix_sum(uvw, freqs, grid_size, tile_size_u, tile_size_v,
                         cell_size_rad, support, host_branch_slip=False):
    """-> (num_visibilities, tile_offsets[num_tiles + 1], num_points_in_tiles
    [num_tiles], num_skipped), int32 as in the reference."""
    _, tile, _, _, _, _, skipped, geo = entries(
        uvw, freqs, grid_size, tile_size_u, tile_size_v, cell_size_rad,
        support, host_branch_slip)
    num_tiles = geo["num_tiles_u"] * geo["num_tiles_v"]
    counts = np.bincount(tile, minlength=num_tiles).astype(np.int32)
    offsets = np.zeros(num_tiles + 1, dtype=np.int32)
    offsets[1:] = np.cumsum(counts)
    return int(offsets[-1]), offsets, counts, skipped


def _slots(tile, tile_offsets):
    """Destination of every entry: the caller's offset of its tile plus its
    rank inside the tile in emission order (.cpp:199-200); also the offsets
    as the reference leaves them (each advanced by its tile's entries)."""
    order = np.argsort(tile, kind="stable")
    sorted_tile = tile[order]
    start = np.searchsorted(sorted_tile, sorted_tile, side="left")
    rank = np.empty(tile.size, dtype=np.int64)
    rank[order] = np.arange(tile.size) - start
    dest = tile_offsets[tile].astype(np.int64) + rank
    after = tile_offsets.copy()
    np.add.at(after, tile, 1)
    return dest, after


def bucket_sort(uvw, freqs, vis, weights, grid_size, tile_size_u, tile_size_v,
                cell_size_rad, support, tile_offsets, num_out,
                host_branch_slip=False, complex_vis=False):
    """-> dict(sorted_uu, sorted_vv, sorted_vis, sorted_weight, sorted_tile,
    tile_offsets). `vis` and `weights` are read at flat element i_vis of the
    array taken as reals of uvw's precision, as the reference's kernels do
    (VIS_TYPE = double / float, .cu:203-204); complex_vis=True copies the
    complex element instead (the product's extension for complex outputs)."""
    real = uvw.dtype.type
    i_vis, tile, pu, pv, pos_u, pos_v, _, _ = entries(
        uvw, freqs, grid_size, tile_size_u, tile_size_v, cell_size_rad,
        support, host_branch_slip)
    dest, after = _slots(tile, tile_offsets)
    keep = (dest >= 0) & (dest < num_out)
    dest = dest[keep]
    out = dict(sorted_uu=np.zeros(num_out, dtype=real),
               sorted_vv=np.zeros(num_out, dtype=real),
               sorted_weight=np.zeros(num_out, dtype=real),
               sorted_tile=np.zeros(num_out, dtype=np.int32),
               tile_offsets=after)
    out["sorted_uu"][dest] = pos_u[keep]
    out["sorted_vv"][dest] = pos_v[keep]
    out["sorted_tile"][dest] = (pv * 32768 + pu)[keep]
    if complex_vis:
        pols = vis.shape[-1]
        out["sorted_vis"] = np.zeros(num_out, dtype=vis.dtype)
        out["sorted_vis"][dest] = vis.reshape(-1)[i_vis[keep] * pols]
    else:
        out["sorted_vis"] = np.zeros(num_out, dtype=real)
        out["sorted_vis"][dest] = vis.reshape(-1).view(real)[i_vis[keep]]
    out["sorted_weight"][dest] = weights.reshape(-1).view(real)[i_vis[keep]]
    return out


def tiled_indexing(uvw, freqs, grid_size, tile_size_u, tile_size_v,
                   cell_size_rad, support, tile_offsets, num_out,
                   host_branch_slip=False):
    """-> dict(sorted_tile, sorted_uu, sorted_vv, sorted_vis_index,
    tile_offsets). The reference's HOST branch leaves sorted_uu / sorted_vv
    zero (.cpp:282-294); its kernels fill them (.cu:281-282), as here unless
    host_branch_slip is set."""
    real = uvw.dtype.type
    i_vis, tile, pu, pv, pos_u, pos_v, _, _ = entries(
        uvw, freqs, grid_size, tile_size_u, tile_size_v, cell_size_rad,
        support, host_branch_slip)
    dest, after = _slots(tile, tile_offsets)
    keep = (dest >= 0) & (dest < num_out)
    dest = dest[keep]
    out = dict(sorted_uu=np.zeros(num_out, dtype=real),
               sorted_vv=np.zeros(num_out, dtype=real),
               sorted_tile=np.zeros(num_out, dtype=np.int32),
               sorted_vis_index=np.zeros(num_out, dtype=np.int32),
               tile_offsets=after)
    if not host_branch_slip:
        out["sorted_uu"][dest] = pos_u[keep]
        out["sorted_vv"][dest] = pos_v[keep]
    out["sorted_tile"][dest] = (pv * 32768 + pu)[keep]
    out["sorted_vis_index"][dest] = i_vis[keep]
    return out
