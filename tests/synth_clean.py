"""Seeded inputs of the CLEAN tests: a PSF with sidelobes (peak 1 at [n, n])
and a dirty image made of shifted copies of it plus noise."""
import numpy as np


def clean_case(seed=1, n=64, num_sources=6, single=False, noise=1e-3,
               beam=(2.0, 1.5, 30.0, 16.0)):
    rng = np.random.default_rng(seed)
    real = np.float32 if single else np.float64
    d = np.arange(2 * n, dtype=np.float64) - n
    r2 = d[:, None] ** 2 + d[None, :] ** 2
    r = np.sqrt(r2)
    # main lobe + decaying rings + a few spokes: a synthesised-beam look-alike
    psf = (np.exp(-r2 / (2 * 2.2 ** 2))
           + 0.08 * np.cos(0.9 * r) * np.exp(-r / (0.35 * n))
           + 0.02 * np.cos(0.31 * d[:, None] + 0.17 * d[None, :])
           * np.exp(-r / (0.5 * n)))
    psf /= psf[n, n]
    dirty = rng.normal(0.0, noise, (n, n))
    for _ in range(num_sources):
        x, y = rng.integers(n // 8, n - n // 8, 2)
        dirty += rng.uniform(0.5, 3.0) * psf[n - x:2 * n - x, n - y:2 * n - y]
    return dict(dirty=dirty.astype(real), psf=psf.astype(real),
                cbeam_details=np.array(beam, dtype=real), loop_gain=0.1,
                threshold=0.01, n=n)
