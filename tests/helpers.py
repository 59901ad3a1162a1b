"""Shared helpers of the GPU parity tests: run the product through its public
Python class (which goes through the C-ABI) and the oracle on the same inputs."""
import numpy as np

from oracle import es_oracle as eo
from ska_sdp_func.grid_data import GridderUvwEsFft

TOL = {np.dtype(np.float32): 1e-5, np.dtype(np.float64): 1e-10}


def rel_max(a, b):
    """max|a - b| / max|b| -- the parity metric of BASELINE.json."""
    return float(np.max(np.abs(np.asarray(a, dtype=np.complex128)
                               - np.asarray(b, dtype=np.complex128)))
                 / np.max(np.abs(b)))


def product_grid(case, eps, wstack, gridder=None):
    """Dirty image from numpy arrays through the C-ABI (host staging path)."""
    dirty = np.zeros((case["image"].shape[0],) * 2, dtype=case["uvw"].dtype)
    if gridder is None:
        gridder = GridderUvwEsFft(case["uvw"], case["freq_hz"], case["vis"],
                                  case["weight"], dirty, case["pixel_size"],
                                  case["pixel_size"], eps, wstack)
    gridder.grid_uvw_es_fft(case["uvw"], case["freq_hz"], case["vis"],
                            case["weight"], dirty)
    return dirty, gridder


def product_degrid(case, eps, wstack, gridder=None):
    """(predicted vis, corrected image) from numpy arrays through the C-ABI."""
    image = case["image"].copy()
    vis = np.zeros_like(case["vis"])
    if gridder is None:
        gridder = GridderUvwEsFft(case["uvw"], case["freq_hz"], vis,
                                  case["weight"], image, case["pixel_size"],
                                  case["pixel_size"], eps, wstack)
    gridder.ifft_grid_uvw_es(case["uvw"], case["freq_hz"], vis,
                             case["weight"], image)
    return vis, image, gridder


def oracle_for(case, eps, wstack):
    return eo.OracleGridder(case["uvw"], case["freq_hz"], case["vis"],
                            case["weight"], case["image"], case["pixel_size"],
                            case["pixel_size"], eps, wstack)


def oracle_grid(case, eps, wstack):
    o = oracle_for(case, eps, wstack)
    return o.grid_uvw_es_fft(case["uvw"], case["freq_hz"], case["vis"],
                             case["weight"])[0]


def oracle_degrid(case, eps, wstack):
    o = oracle_for(case, eps, wstack)
    vis, img, _ = o.ifft_grid_uvw_es(case["uvw"], case["freq_hz"], case["image"])
    return vis, img


def check_parity(case, eps, wstack, do_grid=True, do_degrid=True):
    """Returns dict of parity errors; asserts against the stated tolerance:
    1e-5 (fp32) / 1e-10 (fp64) of max|oracle| (BASELINE.json north_star)."""
    tol = TOL[np.dtype(case["uvw"].dtype)]
    out = {}
    gridder = None
    if do_grid:
        dirty, gridder = product_grid(case, eps, wstack)
        out["grid"] = rel_max(dirty, oracle_grid(case, eps, wstack))
        assert out["grid"] <= tol, out
    if do_degrid:
        vis, img, gridder = product_degrid(case, eps, wstack, gridder)
        o_vis, o_img = oracle_degrid(case, eps, wstack)
        out["degrid"] = rel_max(vis, o_vis)
        out["image"] = rel_max(img, o_img)
        assert out["degrid"] <= tol and out["image"] <= tol, out
    return out


def check_parity_planes(case, eps, count, do_grid=True, do_degrid=True,
                        image_size=None, wstack=True):
    """Parity on a block of `count` w-planes from the middle of the lattice
    (all planes if count is None), all visibilities of the case, at the
    case's real grid / image size: product restricted with set_plane_range
    against the oracle restricted to the same planes. The oracle's cost is
    one G x G fp64 FFT per plane, which is what bounds `count` at the grid
    sizes bench.py times (G = 6160 ... 24640). Returns (errors, plan info)."""
    n = image_size or case["image"].shape[0]
    real = case["uvw"].dtype
    tol = TOL[np.dtype(real)]
    dirty = np.zeros((n, n), dtype=real)
    gridder = GridderUvwEsFft(case["uvw"], case["freq_hz"], case["vis"],
                              case["weight"], dirty, case["pixel_size"],
                              case["pixel_size"], eps, wstack)
    info = gridder.info()
    planes = info["num_total_w_grids"]
    o = eo.OracleGridder(case["uvw"], case["freq_hz"], case["vis"],
                         case["weight"], dirty, case["pixel_size"],
                         case["pixel_size"], eps, wstack)
    assert o.plan["grid_size"] == info["grid_size"]
    assert o.plan["num_total_w_grids"] == planes
    # one lattice on both sides: an ulp of difference in max |w| (fp32) moves
    # plane p by 6e-8 p planes, which alone exceeds the tolerance at p ~ 700
    assert o.plan["w_scale"] == info["w_scale"] and \
        o.plan["min_plane_w"] == info["min_plane_w"], (o.plan, info)
    if count is None or count >= planes:
        lo, hi = 0, planes
    else:
        lo = planes // 2 - count // 2
        hi = lo + count
    gridder.set_plane_range(lo, hi, apply_correction=True)
    out = {}
    if do_grid:
        gridder.grid_uvw_es_fft(case["uvw"], case["freq_hz"], case["vis"],
                                case["weight"], dirty)
        want, _ = o.grid_uvw_es_fft(case["uvw"], case["freq_hz"], case["vis"],
                                    case["weight"], planes=(lo, hi))
        out["grid"] = rel_max(dirty, want)
        assert np.max(np.abs(want)) > 0
        assert out["grid"] <= tol, (out, info)
    if do_degrid:
        image = case.get("image")
        if image is None or image.shape[0] != n:
            image = np.random.default_rng(3).uniform(-0.5, 0.5, (n, n)).astype(real)
        img = image.copy()
        vis = np.zeros_like(case["vis"])
        gridder.ifft_grid_uvw_es(case["uvw"], case["freq_hz"], vis,
                                 case["weight"], img)
        want, want_img, _ = o.ifft_grid_uvw_es(case["uvw"], case["freq_hz"],
                                               image, planes=(lo, hi))
        out["degrid"] = rel_max(vis, want)
        out["image"] = rel_max(img, want_img)
        assert np.max(np.abs(want)) > 0
        assert out["degrid"] <= tol and out["image"] <= tol, (out, info)
    return out, info
