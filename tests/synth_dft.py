"""Seeded inputs of the DFT tests (the reference's data model)."""
import numpy as np


def dft_case(seed=1, num_components=20, num_pols=4, num_channels=10,
             num_baselines=351, num_times=3, max_uvw_m=3000.0, fov=0.05,
             channel_start_hz=100e6, channel_step_hz=100e3):
    """Sources within `fov` radians of the phase centre, n = sqrt(1 - l^2 -
    m^2) - 1; baselines up to max_uvw_m metres (|phase| up to ~1e4 rad at the
    defaults, where the rounding of a phase in radians shows)."""
    rng = np.random.default_rng(seed)
    lm = rng.uniform(-fov, fov, (num_components, 2))
    n = np.sqrt(1.0 - lm[:, 0] ** 2 - lm[:, 1] ** 2) - 1.0
    directions = np.column_stack([lm, n])
    shape = (num_components, num_channels, num_pols)
    fluxes = rng.uniform(0.1, 1.0, shape) + 1j * rng.uniform(-0.3, 0.3, shape)
    uvw = rng.uniform(-max_uvw_m, max_uvw_m, (num_times, num_baselines, 3))
    uvw[..., 2] *= 0.1
    freqs = channel_start_hz + np.arange(num_channels) * channel_step_hz
    uvw_lambda = uvw[:, :, None, :] * (freqs / 299792458.0)[None, None, :, None]
    return dict(directions=directions, fluxes=fluxes, uvw=uvw,
                uvw_lambda=np.ascontiguousarray(uvw_lambda),
                channel_start_hz=channel_start_hz,
                channel_step_hz=channel_step_hz,
                vis_shape=(num_times, num_baselines, num_channels, num_pols))
