"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's uv-grid
visibility weighting, used as the parity oracle of sdp_weighting_uniform /
sdp_weighting_briggs. Never imported by the product.

Follows the reference's host loops operation by operation
(/root/reference/src/ska-sdp-func/visibility/sdp_weighting.cpp):
  weights_grid_write        :20-75    -> grid_write
  sum_weights_calc          :78-133   -> sums
  robustness_calc           :136-144  -> robustness
  uniform_weights_grid_read :150-211  -> uniform
  briggs_weights_grid_read  :214-279  -> briggs
numpy-vectorised: the cell arithmetic is the reference's fp64 expression
(freq / c, u * that, / max_abs_uv * (G/2), floor, + G/2); the grid is
accumulated with np.add.at, which adds in flattened (time, baseline, channel,
pol) order in the grid's own precision -- the order and precision of the
reference's loops. Pinned by tests/golden/golden_weighting.npz, outputs of the
unmodified reference library's CPU path (tests/golden/make_golden_weighting.py).
"""
import numpy as np

C_0 = 299792458.0  # sdp_weighting.cpp:10


def cells(uvw, freq_hz, max_abs_uv, grid_size):
    """(idx_u, idx_v, inside) per (time, baseline, channel), :44-53. `inside`
    also excludes negative indices (out of bounds in the reference)."""
    half = grid_size // 2
    inv_wavelength = freq_hz.astype(np.float64) / C_0
    grid_u = uvw[:, :, 0, None].astype(np.float64) * inv_wavelength[None, None, :]
    grid_v = uvw[:, :, 1, None].astype(np.float64) * inv_wavelength[None, None, :]
    idx_u = (np.floor(grid_u / max_abs_uv * half) + half).astype(np.int64)
    idx_v = (np.floor(grid_v / max_abs_uv * half) + half).astype(np.int64)
    inside = (idx_u < grid_size) & (idx_v < grid_size) & (idx_u >= 0) & (idx_v >= 0)
    return idx_u, idx_v, inside


def grid_write(uvw, freq_hz, max_abs_uv, grid, input_weights):
    """grid[idx_u, idx_v, pol] += input weight, in visit order (:20-75)."""
    idx_u, idx_v, inside = cells(uvw, freq_hz, max_abs_uv, grid.shape[0])
    t, b, c = np.nonzero(inside)
    for p in range(grid.shape[2]):
        np.add.at(grid[:, :, p], (idx_u[t, b, c], idx_v[t, b, c]),
                  input_weights[t, b, c, p])
    return idx_u, idx_v, inside


def uniform(uvw, freq_hz, max_abs_uv, grid, input_weights, output_weights):
    """sdp_weighting_uniform on host arrays; grid and output are updated in
    place (visibilities outside the grid keep their output value)."""
    idx_u, idx_v, inside = grid_write(uvw, freq_hz, max_abs_uv, grid, input_weights)
    t, b, c = np.nonzero(inside)
    one = grid.dtype.type(1.0)
    output_weights[t, b, c, :] = one / grid[idx_u[t, b, c], idx_v[t, b, c], :]


def briggs(uvw, freq_hz, max_abs_uv, robust_param, grid, input_weights,
           output_weights):
    """sdp_weighting_briggs on host arrays."""
    idx_u, idx_v, inside = grid_write(uvw, freq_hz, max_abs_uv, grid, input_weights)
    t, b, c = np.nonzero(inside)
    g = grid[idx_u[t, b, c], idx_v[t, b, c], :].astype(np.float64)
    # :78-133: running fp64 sums in visit order (np.cumsum keeps the order)
    flat = g.reshape(-1)
    sumweight = float(np.cumsum(flat)[-1]) if flat.size else 0.0
    sumweight2 = float(np.cumsum(flat * flat)[-1]) if flat.size else 0.0
    numerator = pow(5.0 * 1 / (pow(10.0, robust_param)), 2.0)
    robustness = numerator / (sumweight2 / sumweight)
    w = input_weights[t, b, c, :].astype(np.float64) / (1.0 + robustness * g)
    output_weights[t, b, c, :] = w.astype(grid.dtype)
    return robustness
