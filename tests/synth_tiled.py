"""Seeded inputs of the tiled-function tests (the reference's data model:
uvw [T, B, 3] metres, freqs [C] Hz, vis / weights [T, B, C, 1])."""
import numpy as np

C_0 = 299792458.0


def tiled_case(seed=1, single=False, num_times=3, num_baselines=50,
               num_chan=6, grid_size=400, tile_size_u=32, tile_size_v=16,
               support=4, spread=0.55, positive=False, core=0.0,
               spread_v=None):
    """Visibilities spread over `spread` x the half grid in u and v (> 0.5
    leaves some off the grid: they are counted as skipped); `core` is the
    fraction drawn from a 20x narrower distribution (dense centre tiles);
    positive: u, v >= 0 only; spread_v: another spread for v (the reference
    writes out of bounds for visibilities in the last half tile ROW of a grid
    whose size is a multiple of the tile size: cases meant for it keep v away
    from there)."""
    rng = np.random.default_rng(seed)
    real = np.float32 if single else np.float64
    freqs = np.linspace(1.0e9, 1.5e9, num_chan)
    cell_size_rad = 4.06e-5
    # pos = u * f / c * grid_size * cell_size_rad (cells from the centre)
    metres = spread * grid_size / (freqs[-1] / C_0 * grid_size * cell_size_rad)
    uvw = rng.uniform(0.0 if positive else -metres, metres,
                      (num_times, num_baselines, 3))
    if spread_v is not None:
        uvw[:, :, 1] *= spread_v / spread
    if core > 0.0:
        narrow = rng.uniform(size=(num_times, num_baselines)) < core
        uvw[narrow] *= 0.05
    shape = (num_times, num_baselines, num_chan, 1)
    vis = (rng.uniform(-1, 1, shape) + 1j * rng.uniform(-1, 1, shape))
    weights = rng.uniform(0.5, 2.0, shape)
    return dict(uvw=uvw.astype(real), freqs=freqs.astype(real),
                vis=vis.astype(np.complex64 if single else np.complex128),
                weights=weights.astype(real), grid_size=grid_size,
                tile_size_u=tile_size_u, tile_size_v=tile_size_v,
                cell_size_rad=cell_size_rad, support=support)


def num_tiles(c):
    return (((c["grid_size"] + c["tile_size_u"] - 1) // c["tile_size_u"]) *
            ((c["grid_size"] + c["tile_size_v"] - 1) // c["tile_size_v"]))


def geometry_args(c):
    return (c["grid_size"], c["tile_size_u"], c["tile_size_v"],
            c["cell_size_rad"], c["support"])
