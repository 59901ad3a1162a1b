"""Seeded synthetic inputs for the ES gridder path.

Shared by the parity tests and by bench.py. Everything is generated on the
host with numpy; ``scale`` arguments shrink a configuration for tests while
keeping its geometry (same fractional uv coverage, same field of view).
"""

import numpy as np

C_0 = 299792458.0


def reference_test_case(nxydirty=256, num_vis=1000, num_chan=10, seed=40,
                        single=False):
    """The generator of the reference's own adjointness test
    (tests/grid_data/test_gridder_uvw_es_fft.py:535-561), BASELINE config C1
    at nxydirty=256. About half of the upper-channel visibilities fall off
    the grid by design (exercises the clipping rules)."""
    np.random.seed(seed)
    pixel_size_rad = 2 * np.pi / 180 / nxydirty
    f_0 = 1e9
    freqs = f_0 + np.arange(num_chan) * (f_0 / num_chan)
    uvw = (np.random.rand(num_vis, 3) - 0.5) / (pixel_size_rad * f_0 / C_0)
    vis = (np.random.rand(num_vis, num_chan) - 0.5
           + 1j * (np.random.rand(num_vis, num_chan) - 0.5))
    image = np.random.rand(nxydirty, nxydirty) - 0.5
    weight = np.ones([num_vis, num_chan])
    if single:
        vis = vis.astype(np.complex64)
        image = image.astype(np.float32)
        freqs = freqs.astype(np.float32)
        uvw = uvw.astype(np.float32)
        weight = weight.astype(np.float32)
    return dict(uvw=uvw, freq_hz=freqs, vis=vis, weight=weight, image=image,
                pixel_size=pixel_size_rad, image_size=nxydirty,
                name=f"C1-ref-test-{nxydirty}")


def uniform_box(num_rows=1000, num_chan=1024, image_size=4096, fov_deg=2.0,
                seed=41, single=True, w_factor=1.0):
    """BASELINE config C2 ("uniform box"): u, v, w ~ U(-0.45, 0.45) / (pix f_max / c)
    so that every footprint is on the grid; f = 1 GHz (1 + 0.25 c / C)."""
    rng = np.random.default_rng(seed)
    pix = fov_deg * np.pi / 180 / image_size
    freqs = 1e9 * (1.0 + 0.25 * np.arange(num_chan) / num_chan)
    scale = 1.0 / (pix * freqs[-1] / C_0)
    uvw = rng.uniform(-0.45, 0.45, (num_rows, 3)) * scale
    uvw[:, 2] *= w_factor
    vis = (rng.uniform(-0.5, 0.5, (num_rows, num_chan))
           + 1j * rng.uniform(-0.5, 0.5, (num_rows, num_chan)))
    weight = rng.uniform(0.5, 1.5, (num_rows, num_chan))
    image = rng.uniform(-0.5, 0.5, (image_size, image_size))
    out = dict(uvw=uvw, freq_hz=freqs, vis=vis, weight=weight, image=image,
               pixel_size=pix, name=f"C2-uniform-box-{num_rows}x{num_chan}-{image_size}")
    return _cast(out, single)


def xyz_to_uvw(xyz, ha, dec):
    """Station/baseline ENU-like xyz -> uvw for hour angle and declination
    (the standard rotation, as in the reference's test fixture
    tests/grid_data/test_gridder_wtower_uvw.py:326-340)."""
    x, y, z = xyz[:, 0], xyz[:, 1], xyz[:, 2]
    u = np.sin(ha) * x + np.cos(ha) * y
    v0 = -np.cos(ha) * x + np.sin(ha) * y
    w = np.cos(dec) * v0 + np.sin(dec) * z
    v = np.sin(dec) * v0 - np.cos(dec) * z  # sign irrelevant for synthetic data
    return np.stack([u, -v, w], axis=1)


def _array_baselines(rng, num_stations, r_core, core_frac, r_max, z_sigma,
                     ha, dec):
    n_core = int(num_stations * core_frac)
    r = np.concatenate([
        r_core * np.sqrt(rng.uniform(0, 1, n_core)),
        np.exp(rng.uniform(np.log(r_core), np.log(r_max),
                           num_stations - n_core)),
    ])
    az = rng.uniform(0, 2 * np.pi, num_stations)
    xyz = np.stack([r * np.cos(az), r * np.sin(az),
                    rng.normal(0, z_sigma, num_stations)], axis=1)
    i, j = np.triu_indices(num_stations, 1)
    return xyz_to_uvw(xyz[j] - xyz[i], ha, dec)


def ska_low_snapshot(num_stations=512, num_chan=1024, image_size=16384,
                     seed=43, single=True, r_max=35e3):
    """BASELINE config C4 ("SKA-Low AA*-like snapshot"): core-heavy station
    layout (half within 500 m, rest log-uniform to r_max), one time step,
    f = 130-165 MHz, pixel size 0.45 / max(|u|, |v|) in wavelengths at f_max."""
    rng = np.random.default_rng(seed)
    uvw = _array_baselines(rng, num_stations, 500.0, 0.5, r_max, 5.0,
                           ha=0.3, dec=-0.5)
    freqs = np.linspace(130e6, 165e6, num_chan)
    max_uv = np.max(np.abs(uvw[:, :2])) * freqs[-1] / C_0
    pix = 0.45 / max_uv
    num_rows = uvw.shape[0]
    vis = (rng.standard_normal((num_rows, num_chan))
           + 1j * rng.standard_normal((num_rows, num_chan)))
    weight = rng.uniform(0.5, 1.5, (num_rows, num_chan))
    image = rng.uniform(-0.5, 0.5, (image_size, image_size)) if image_size <= 4096 \
        else None
    out = dict(uvw=uvw, freq_hz=freqs, vis=vis, weight=weight, image=image,
               pixel_size=pix, image_size=image_size,
               name=f"C4-ska-low-{num_stations}st-{num_chan}ch-{image_size}")
    return _cast(out, single)


def mid_like_predict(num_stations=125, num_times=2, num_chan=1024,
                     image_size=8192, seed=42, single=False, r_max=80e3):
    """BASELINE config C3 ("Mid-like array", degridding): log-uniform station
    radii 30 m - r_max, f = 0.95-1.05 GHz, model image = 200 point sources."""
    rng = np.random.default_rng(seed)
    parts = []
    n_core = 0
    r = np.exp(rng.uniform(np.log(30.0), np.log(r_max), num_stations))
    az = rng.uniform(0, 2 * np.pi, num_stations)
    xyz = np.stack([r * np.cos(az), r * np.sin(az),
                    rng.normal(0, 20.0, num_stations)], axis=1)
    i, j = np.triu_indices(num_stations, 1)
    for t in range(num_times):
        parts.append(xyz_to_uvw(xyz[j] - xyz[i], 0.3 + 0.05 * t, -0.5))
    del n_core
    uvw = np.concatenate(parts, axis=0)
    freqs = np.linspace(0.95e9, 1.05e9, num_chan)
    max_uv = np.max(np.abs(uvw[:, :2])) * freqs[-1] / C_0
    pix = 0.45 / max_uv
    num_rows = uvw.shape[0]
    image = rng.normal(0, 1e-3, (image_size, image_size))
    sx = rng.integers(image_size // 8, 7 * image_size // 8, 200)
    sy = rng.integers(image_size // 8, 7 * image_size // 8, 200)
    image[sx, sy] += rng.uniform(0.1, 1.0, 200)
    vis = (rng.standard_normal((num_rows, num_chan))
           + 1j * rng.standard_normal((num_rows, num_chan)))
    weight = np.ones((num_rows, num_chan))
    out = dict(uvw=uvw, freq_hz=freqs, vis=vis, weight=weight, image=image,
               pixel_size=pix,
               name=f"C3-mid-like-{num_rows}x{num_chan}-{image_size}")
    return _cast(out, single)


def wide_field(num_rows=31250, num_chan=1024, image_size=8192, fov_deg=20.0,
               max_w_lambda=2e4, seed=44, single=True):
    """BASELINE config C5 ("wide-field large-|w|"): u, v uniform in a disc of
    radius 0.45 / pix wavelengths, w ~ U(-1, 1) max_w_lambda, f = 1.0-1.1 GHz."""
    rng = np.random.default_rng(seed)
    pix = fov_deg * np.pi / 180 / image_size
    freqs = np.linspace(1.0e9, 1.1e9, num_chan)
    lam_min = C_0 / freqs[-1]
    rad = 0.45 / pix * np.sqrt(rng.uniform(0, 1, num_rows)) * lam_min
    ang = rng.uniform(0, 2 * np.pi, num_rows)
    w = rng.uniform(-1, 1, num_rows) * max_w_lambda * lam_min
    uvw = np.stack([rad * np.cos(ang), rad * np.sin(ang), w], axis=1)
    vis = (rng.uniform(-0.5, 0.5, (num_rows, num_chan))
           + 1j * rng.uniform(-0.5, 0.5, (num_rows, num_chan)))
    weight = rng.uniform(0.5, 1.5, (num_rows, num_chan))
    image = rng.uniform(-0.5, 0.5, (image_size, image_size))
    out = dict(uvw=uvw, freq_hz=freqs, vis=vis, weight=weight, image=image,
               pixel_size=pix,
               name=f"C5-wide-field-{num_rows}x{num_chan}-{image_size}")
    return _cast(out, single)


def _cast(case, single):
    if "image_size" not in case and case.get("image") is not None:
        case["image_size"] = case["image"].shape[0]
    real = np.float32 if single else np.float64
    cplx = np.complex64 if single else np.complex128
    for key in ("uvw", "freq_hz", "weight", "image"):
        if case[key] is not None:
            case[key] = np.ascontiguousarray(case[key], dtype=real)
    case["vis"] = np.ascontiguousarray(case["vis"], dtype=cplx)
    return case
