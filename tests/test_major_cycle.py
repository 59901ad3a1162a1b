"""One major cycle on the device, end to end through the public Python API, in
the shape of the reference's own integration test
(tests/clean/test_hogbom_clean.py:25-170: DFT predict -> ES gridder (dirty
image and PSF) -> Hogbom CLEAN), closed with the degridder: point sources ->
dft_point_v01 -> GridderUvwEsFft.grid_uvw_es_fft -> hogbom_clean ->
GridderUvwEsFft.ifft_grid_uvw_es -> visibilities again. Every array between the
steps is a device array; nothing returns to the host until the checks.

Differences from the reference's test, on purpose: the PSF is gridded with the
dirty image's pixel size on twice the extent (the reference grids it on the
same field with half the pixel size, which is not the PSF Hogbom's subtraction
needs), and the sources sit on pixel centres so that the recovered model can
be compared with the truth."""
import numpy as np
import pytest

from oracle import clean_oracle as co
from ska_sdp_func.clean import hogbom_clean
from ska_sdp_func.grid_data import GridderUvwEsFft
from ska_sdp_func.visibility import dft_point_v01

C_0 = 299792458.0


@pytest.mark.gpu
def test_predict_grid_clean_degrid_on_the_device():
    import torch
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(12)
    n, num_baselines, num_sources = 256, 2000, 10
    freq0, dfreq = 100e6, 100e3
    pix = 2.0 * np.pi / 180 / n                       # 2 degree field
    # baselines in a disc of 3 km, w = 0 (as the reference's test)
    radius = 3000.0 * rng.uniform(0, 1, num_baselines)
    angle = 2 * np.pi * rng.uniform(0, 1, num_baselines)
    uvw = np.zeros((1, num_baselines, 3))
    uvw[0, :, 0] = radius * np.cos(angle)
    uvw[0, :, 1] = radius * np.sin(angle)
    # sources on pixel centres, at least 12 pixels apart, away from the edge
    cells = []
    while len(cells) < num_sources:
        c = rng.integers(n // 8, n - n // 8, 2)
        if all(np.abs(c - o).max() >= 12 for o in cells):
            cells.append(c)
    cells = np.array(cells)
    flux = rng.uniform(1.0, 10.0, num_sources)
    lm = (cells - n // 2) * pix
    directions = np.column_stack([lm, np.sqrt(1 - (lm ** 2).sum(axis=1)) - 1.0])
    fluxes = flux[:, None, None] + 0j
    one = np.array([[0.0, 0.0, 0.0]])

    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    d_uvw3, d_dir, d_flux = t(uvw), t(directions), t(fluxes)
    vis = torch.zeros((1, num_baselines, 1, 1), dtype=torch.complex128, device=dev)
    vis_psf = torch.zeros_like(vis)
    dft_point_v01(d_dir, d_flux, d_uvw3, freq0, dfreq, vis)
    dft_point_v01(t(one), t(np.ones((1, 1, 1)) + 0j), d_uvw3, freq0, dfreq, vis_psf)

    d_uvw = d_uvw3[0]
    d_freq = t(np.array([freq0]))
    d_weight = torch.ones((num_baselines, 1), dtype=torch.float64, device=dev)
    dirty = torch.zeros((n, n), dtype=torch.float64, device=dev)
    psf = torch.zeros((2 * n, 2 * n), dtype=torch.float64, device=dev)
    g = GridderUvwEsFft(d_uvw, d_freq, vis[0, :, :, 0], d_weight, dirty, pix,
                        pix, 1e-12, False)
    g.grid_uvw_es_fft(d_uvw, d_freq, vis[0, :, :, 0].contiguous(), d_weight, dirty)
    g_psf = GridderUvwEsFft(d_uvw, d_freq, vis_psf[0, :, :, 0], d_weight, psf,
                            pix, pix, 1e-12, False)
    g_psf.grid_uvw_es_fft(d_uvw, d_freq, vis_psf[0, :, :, 0].contiguous(),
                          d_weight, psf)
    dirty /= num_baselines
    psf /= num_baselines

    model, residual, sky = (torch.zeros_like(dirty) for _ in range(3))
    beam = np.array([2.0, 2.0, 0.0, 32.0])
    cycles = 3000
    hogbom_clean(dirty, psf, beam, 0.1, 1e-3, cycles, model, residual, sky)

    predicted = torch.zeros((num_baselines, 1), dtype=torch.complex128, device=dev)
    image = model.clone()        # the degridder corrects its image in place
    g.ifft_grid_uvw_es(d_uvw, d_freq, predicted, d_weight, image)
    torch.cuda.synchronize()

    # ---- checks (host from here on)
    h_psf, h_dirty, h_model = psf.cpu().numpy(), dirty.cpu().numpy(), model.cpu().numpy()
    assert abs(h_psf[n, n] - 1.0) < 1e-9 and np.argmax(h_psf) == n * 2 * n + n
    # the CLEAN of this library on the device = the oracle on the same images
    want = co.hogbom_clean(h_dirty, h_psf, beam, 0.1, 1e-3, cycles)
    assert np.array_equal(h_model, want["clean_model"])
    assert np.array_equal(residual.cpu().numpy(), want["residual"])
    assert 100 < want["cycles"] < cycles          # stopped by the threshold
    # every source is recovered at its pixel (+-1) with its flux
    for (x, y), f in zip(cells, flux):
        got = h_model[x - 1:x + 2, y - 1:y + 2].sum()
        assert abs(got - f) / f < 0.02, (x, y, f, got)
    assert abs(h_model.sum() - flux.sum()) / flux.sum() < 0.01
    assert np.abs(residual.cpu().numpy()).max() < 2e-3
    # and the model predicts the visibilities it was made from
    true = vis[0, :, 0, 0].cpu().numpy()
    got = predicted[:, 0].cpu().numpy()
    err = np.sqrt(np.mean(np.abs(got - true) ** 2) / np.mean(np.abs(true) ** 2))
    print("major cycle: cycles", want["cycles"], "relative rms of the "
          "predicted visibilities", err)
    assert err < 0.02
