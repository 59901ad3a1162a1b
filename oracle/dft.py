"""TEST INFRASTRUCTURE ONLY (oracle).

Explicit O(N_vis * N_pix) evaluation of the measurement equation the
reference's w-stacking ES gridder approximates (SURVEY.md section 0.6,
verified there against the reference's kernels):

  dirty[ix, iy] = sum_k wgt_k * Re( vis_k * exp(+2 pi i (f/c) (x u + y v - w (n - 1))) ) / n
  vis_k        += sum_pix img[ix, iy] / n * exp(-2 pi i (f/c) (x u + y v - w (n - 1)))   (weights ignored)

with x = (ix - N/2) pix <-> row <-> u, y = (iy - N/2) pix <-> column <-> v,
n = sqrt(1 - x^2 - y^2); do_w_stacking=False drops the w term and the 1/n.
Small sizes only. Used to pin signs, axis order and normalisation.
"""

import numpy as np

C_0 = 299792458.0


def _geometry(image_size, pixel_size):
    idx = (np.arange(image_size) - image_size // 2) * pixel_size
    x, y = np.meshgrid(idx, idx, indexing="ij")
    n = np.sqrt(1.0 - x * x - y * y)
    return x, y, n


def dft_grid(uvw, freq_hz, vis, weight, image_size, pixel_size,
             do_w_stacking=True):
    x, y, n = _geometry(image_size, pixel_size)
    dirty = np.zeros((image_size, image_size))
    for r in range(uvw.shape[0]):
        for c in range(freq_hz.shape[0]):
            k = freq_hz[c] / C_0
            u, v, w = uvw[r] * k
            phase = x * u + y * v
            if do_w_stacking:
                phase = phase - w * (n - 1.0)
            term = (vis[r, c] * weight[r, c]) * np.exp(2j * np.pi * phase)
            dirty += term.real / n if do_w_stacking else term.real
    return dirty


def dft_degrid(uvw, freq_hz, image, pixel_size, do_w_stacking=True):
    image_size = image.shape[0]
    x, y, n = _geometry(image_size, pixel_size)
    vis = np.zeros((uvw.shape[0], freq_hz.shape[0]), dtype=np.complex128)
    src = image / n if do_w_stacking else image
    for r in range(uvw.shape[0]):
        for c in range(freq_hz.shape[0]):
            k = freq_hz[c] / C_0
            u, v, w = uvw[r] * k
            phase = x * u + y * v
            if do_w_stacking:
                phase = phase - w * (n - 1.0)
            vis[r, c] = np.sum(src * np.exp(-2j * np.pi * phase))
    return vis
