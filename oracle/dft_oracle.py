"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's point-source
DFT predict, the parity oracle of sdp_dft_point_v00 / sdp_dft_point_v01. Never
imported by the product.

Follows the reference's host loops (/root/reference/src/ska-sdp-func/
visibility/sdp_dft.cpp):
  dft_point_v00  :21-99    phase = -2 pi (l u + m v + n w), uvw in wavelengths
                           per channel; vis = sum over components of
                           flux * (cos phase + i sin phase), in component order
  dft_point_v01  :212-286  the same with uvw in metres and
                           inv_wavelength = (start + c step) / c0 per channel,
                           phase = -2 pi inv_wavelength (l u + m v + n w)
numpy-vectorised over the visibilities, sequential over the components (the
order of the reference's sums); fp64 throughout, the sum cast to the
visibility type at the end. `exact=True` keeps the phase in turns and reduces
it before the factor 2 pi (t - rint(t) is exact), which is what the product
does: the two differ by the reference's own rounding of the phase in radians
(~1e-16 x |phase|). Pinned by tests/golden/golden_dft.npz,
outputs of the unmodified reference library's CPU path
(tests/golden/make_golden_dft.py).
"""
import numpy as np

C_0 = 299792458.0  # sdp_dft.cpp:11


def _sum(dot_turns, fluxes, exact):
    """dot_turns[s][T, B, C] -> vis[T, B, C, P]."""
    num_components, _, num_pols = fluxes.shape
    vis = np.zeros(dot_turns[0].shape + (num_pols,), dtype=np.complex128)
    for s in range(num_components):
        t = dot_turns[s]
        if exact:
            phase = -2.0 * np.pi * (t - np.rint(t))
        else:
            phase = t  # already radians
        phasor = np.cos(phase) + 1j * np.sin(phase)
        vis += phasor[..., None] * fluxes[s][None, None, :, :]
    return vis


def dft_point_v00(directions, fluxes, uvw_lambda, vis_dtype=np.complex128,
                  exact=False):
    """:21-99 (phase in radians, in the reference's operation order)."""
    dots = []
    for s in range(directions.shape[0]):
        l, m, n = directions[s]
        d = l * uvw_lambda[..., 0] + m * uvw_lambda[..., 1] + n * uvw_lambda[..., 2]
        dots.append(d if exact else -2.0 * np.pi * d)
    return _sum(dots, fluxes, exact).astype(vis_dtype)


def dft_point_v01(directions, fluxes, uvw, channel_start_hz, channel_step_hz,
                  vis_dtype=np.complex128, exact=False):
    """:212-286."""
    num_channels = fluxes.shape[1]
    inv_wavelength = (channel_start_hz +
                      np.arange(num_channels) * channel_step_hz) / C_0
    dots = []
    for s in range(directions.shape[0]):
        l, m, n = directions[s]
        d = (l * uvw[..., 0] + m * uvw[..., 1] + n * uvw[..., 2])[..., None]
        if exact:
            dots.append(inv_wavelength[None, None, :] * d)
        else:
            dots.append((-2.0 * np.pi * inv_wavelength)[None, None, :] * d)
    return _sum(dots, fluxes, exact).astype(vis_dtype)
