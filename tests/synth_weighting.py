"""Seeded inputs of the uv-grid weighting tests (shapes of the reference's
data model: uvw [T, B, 3], weights [T, B, C, P], grid [G, G, P])."""
import numpy as np

C_0 = 299792458.0


def weighting_case(seed=1, single=False, num_times=3, num_baselines=40,
                   num_chan=5, num_pols=1, grid_size=32, shrink=1.0,
                   positive=False):
    """shrink < 1: the grid covers only that fraction of the uv extent, so
    some visibilities fall outside it (their output weight is untouched).
    positive: u, v >= 0 only -- the reference checks the upper bound of a
    cell index only (sdp_weighting.cpp:53), so a case meant for it must not
    leave the grid on the negative side."""
    rng = np.random.default_rng(seed)
    uvw = rng.uniform(0.0 if positive else -900.0, 900.0,
                      (num_times, num_baselines, 3))
    freq = np.linspace(1.0e8, 1.3e8, num_chan)
    wtype = np.float32 if single else np.float64
    inp = rng.uniform(0.5, 2.0, (num_times, num_baselines, num_chan, num_pols))
    max_abs_uv = float(np.amax(np.abs(uvw[:, :, 0:2])) * freq[-1] / C_0) * 1.0001
    return dict(uvw=uvw, freq_hz=freq, input=inp.astype(wtype),
                grid=np.zeros((grid_size, grid_size, num_pols), dtype=wtype),
                output0=np.full(inp.shape, 7.0, dtype=wtype),
                max_abs_uv=max_abs_uv * shrink, robust=-0.5 + 0.5 * seed % 2)
