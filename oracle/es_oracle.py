"""TEST INFRASTRUCTURE ONLY (oracle) -- never imported by the product package.

CPU restatement of the reference's sdp_GridderUvwEsFft path:

* plan maths in numpy/Python, restating
  ``sdp_gridder_uvw_es_fft.cpp:301-424`` (plan scalars, w-plane lattice) and
  ``sdp_gridder_uvw_es_fft_utils.cpp:122-218,225-537`` (KernelDB selection,
  good FFT sizes, Gauss-Legendre correction tables incl. the reference's
  Newton iteration and its stopping rule);
* the per-visibility / per-pixel loops in plain C (``es_oracle.c``), fp64
  throughout with fp64 coordinates ("oracle64" for fp32 configurations:
  ``plan_prec=4`` selects the fp32 plan, the arithmetic stays fp64).

Pinned (tests/test_oracle.py) against oracle/_ref (the reference's own kernel
source compiled for the host), against the explicit DFT and against the golden
vectors in tests/golden/ that were generated from oracle/_ref.
"""

import ctypes
import math
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle_es.so")
C_0 = 299792458.0
QUADRATURE_SUPPORT_BOUND = 32  # es_fft_utils.h:13

_KERNEL_DB = None


def kernel_db():
    """ES kernel parameter table (see oracle/tools/extract_kernel_db.py)."""
    global _KERNEL_DB
    if _KERNEL_DB is None:
        _KERNEL_DB = np.loadtxt(os.path.join(_HERE, "kernel_db.csv"),
                                delimiter=",", comments="#")
    return _KERNEL_DB


def good_size_complex(n):
    """Smallest 11-smooth integer >= n (es_fft_utils.cpp:181-218)."""
    if n <= 12:
        return n
    best = 2 * n
    f11 = 1
    while f11 < best:
        f117 = f11
        while f117 < best:
            f1175 = f117
            while f1175 < best:
                x = f1175
                while x < n:
                    x *= 2
                while True:
                    if x < n:
                        x *= 3
                    elif x > n:
                        if x < best:
                            best = x
                        if x & 1:
                            break
                        x >>= 1
                    else:
                        return n
                f1175 *= 5
            f117 *= 7
        f11 *= 11
    return best


def params_from_epsilon(epsilon, image_size, is_double):
    """(grid_size, support, beta_over_support) -- es_fft_utils.cpp:488-536."""
    db = kernel_db()
    num_opts = 20
    idx = [len(db) - 1] * num_opts
    os_factors = [2.6] * num_opts
    max_support = 16 if is_double else 8
    for i, (w, sigma, eps_row, _beta, corr_range) in enumerate(db):
        this_support = int(math.floor(w))
        if (w <= max_support and eps_row <= epsilon and corr_range <= 10.0
                and sigma <= os_factors[this_support]):
            idx[this_support] = i
            os_factors[this_support] = sigma
    min_nu, min_idx = 80000, len(db)
    for i in range(num_opts):
        sigma = db[idx[i]][1]
        nu = 2 * good_size_complex(int(image_size * sigma * 0.5) + 1)
        if nu <= min_nu:
            min_nu, min_idx = nu, idx[i]
    w, sigma, _eps, beta, _corr = db[min_idx]
    nu = 2 * good_size_complex(int(image_size * sigma * 0.5) + 1)
    return nu, int(math.floor(w)), float(beta)


def _legendre(x, n):
    """(P_n(x), P_n'(x)) by the three-term recurrence, es_fft_utils.cpp:13-57."""
    if n == 0:
        return 1.0, 0.0
    if n == 1:
        return x, 1.0
    p_im2, p_im1, p_i = 1.0, x, 0.0
    for i in range(2, n + 1):
        p_i = x * p_im1 + ((float(i) - 1.0) / float(i)) * (x * p_im1 - p_im2)
        p_im2, p_im1 = p_im1, p_i
    # after the loop p_im2 holds P_{n-1}
    deriv = float(n) / (1.0 - x * x) * (p_im2 - x * p_i)
    return p_i, deriv


def _legendre_root(i, n, accuracy=1e-16, max_iter=100):
    """i-th largest root of P_n and its Gauss-Legendre weight,
    es_fft_utils.cpp:60-110. Newton from Tricomi's estimate; the stopping
    test is fdim(next, current) > accuracy, i.e. it also stops as soon as an
    update moves DOWN -- reproduced because it decides the last bits."""
    improvement = 1.0 - (1.0 - 1.0 / float(n)) / (8.0 * n * n)
    nxt = improvement * math.cos(math.pi * (4.0 * float(i) - 1.0)
                                 / (4.0 * n + 2.0))
    iterations = 0
    while True:
        est = nxt
        p_n, deriv = _legendre(est, n)
        nxt = est - p_n / deriv
        iterations += 1
        if not (max(nxt - est, 0.0) > accuracy and iterations < max_iter):
            break
    _p, deriv = _legendre(nxt, n)
    weight = 2.0 / ((1.0 - nxt * nxt) * deriv * deriv)
    return nxt, weight


def gauss_legendre_tables(image_size, grid_size, support, beta):
    """es_fft_utils.cpp:122-175: (kernel, nodes, weights)[32], corr[N/2+1], C(0)."""
    p = int(1.5 * support + 2.0)
    qk = np.zeros(QUADRATURE_SUPPORT_BOUND)
    qn = np.zeros(QUADRATURE_SUPPORT_BOUND)
    qw = np.zeros(QUADRATURE_SUPPORT_BOUND)
    for i in range(1, p + 1):
        x_i, w_i = _legendre_root(i, 2 * p)
        qn[i - 1] = x_i
        qw[i - 1] = w_i
        qk[i - 1] = math.exp(beta * (math.sqrt(1.0 - x_i * x_i) - 1.0))
    norm = 0.0
    for i in range(p):
        norm += qk[i] * qw[i]
    norm *= float(support)
    corr = np.zeros(image_size // 2 + 1)
    l_m = np.arange(image_size // 2 + 1) * (1.0 / grid_size)
    for i in range(p):  # same summation order as the reference
        corr += qk[i] * qw[i] * np.cos(np.pi * l_m * support * qn[i])
    corr = corr * support / norm
    return qk, qn, qw, corr, norm


def get_w_range(uvw, freq_hz):
    """gridder_uvw_es_fft.py:90-106, in the reference's own operation order
    and precision (|w| extremes times (frequency / c), all in the arrays'
    dtype): a different order moves max_abs_w by an ulp in fp32, which shifts
    every w-plane of the lattice by ~6e-8 of its index -- 4e-5 planes at plane
    700, enough to break 1e-5 parity between two plans that should be one."""
    min_abs_w = np.amin(np.abs(uvw[:, 2]))
    max_abs_w = np.amax(np.abs(uvw[:, 2]))
    min_abs_w = min_abs_w * (freq_hz[0] / C_0)
    max_abs_w = max_abs_w * (freq_hz[-1] / C_0)
    return float(min_abs_w), float(max_abs_w)


def make_plan(image_size, pixel_size, epsilon, do_w_stacking, min_abs_w,
              max_abs_w, is_double):
    """Plan scalars, es_fft.cpp:305-424, as a dict."""
    grid_size, support, beta_row = params_from_epsilon(
        epsilon, image_size, is_double)
    beta = beta_row * support
    plan = dict(image_size=image_size, grid_size=grid_size, support=support,
                beta=beta, pixel_size=pixel_size,
                uv_scale=grid_size * pixel_size,
                do_wstacking=int(bool(do_w_stacking)))
    if do_w_stacking:
        x0 = -0.5 * image_size * pixel_size
        s = x0 * x0 + x0 * x0
        nmin = math.sqrt(max(1.0 - s, 0.0)) - 1.0
        if s > 1.0:
            nmin = -math.sqrt(abs(1.0 - s)) - 1.0
        w_scale = 0.25 / abs(nmin)
        nw = int((max_abs_w - min_abs_w) / w_scale + 2)  # C truncation
        if max_abs_w > min_abs_w:
            w_scale = 1.0 / ((1.0 + 1e-13) * (max_abs_w - min_abs_w) / (nw - 1))
        else:
            # A single |w| value divides by zero in the reference
            # (es_fft.cpp:358, undefined there). The drop-in keeps the
            # unstretched plane spacing; the oracle defines the same.
            nw = 2
            w_scale = 1.0 / w_scale
        plan.update(
            min_abs_w=min_abs_w, max_abs_w=max_abs_w, w_scale=w_scale,
            min_plane_w=min_abs_w - (0.5 * support - 1.0) / w_scale,
            max_plane_w=max_abs_w + (0.5 * support - 1.0) / w_scale,
            num_total_w_grids=nw + support - 2)
    else:
        plan.update(min_abs_w=0.0, max_abs_w=0.0, w_scale=1.0,
                    min_plane_w=0.0, max_plane_w=0.0, num_total_w_grids=1)
    plan["inv_w_scale"] = 1.0 / plan["w_scale"]
    qk, qn, qw, corr, norm = gauss_legendre_tables(
        image_size, grid_size, support, beta)
    plan.update(quad_kernel=qk, quad_nodes=qn, quad_weights=qw,
                conv_corr=corr, conv_corr_norm_factor=norm,
                num_quad=int(1.5 * support + 2.0))
    return plan


class _Params(ctypes.Structure):
    _fields_ = [
        ("num_rows", ctypes.c_int32), ("num_chan", ctypes.c_int32),
        ("image_size", ctypes.c_int32), ("grid_size", ctypes.c_int32),
        ("support", ctypes.c_int32), ("num_total_w_grids", ctypes.c_int32),
        ("do_wstacking", ctypes.c_int32), ("num_quad", ctypes.c_int32),
        ("beta", ctypes.c_double), ("pixel_size", ctypes.c_double),
        ("uv_scale", ctypes.c_double), ("w_scale", ctypes.c_double),
        ("inv_w_scale", ctypes.c_double), ("min_plane_w", ctypes.c_double),
        ("conv_corr_norm_factor", ctypes.c_double),
        ("quad_kernel", ctypes.c_void_p), ("quad_nodes", ctypes.c_void_p),
        ("quad_weights", ctypes.c_void_p), ("conv_corr", ctypes.c_void_p),
    ]


_lib = None


def build():
    """Compile es_oracle.c (building the checker is not using it)."""
    subprocess.run(["make", "-C", _HERE, "liboracle_es.so"], check=True,
                   stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.eso_grid.restype = ctypes.c_int
        _lib.eso_grid.argtypes = [ctypes.c_void_p] * 6 + [ctypes.c_int] * 3 + [
            ctypes.c_void_p]
        _lib.eso_degrid.restype = ctypes.c_int
        _lib.eso_degrid.argtypes = [ctypes.c_void_p] * 5 + [ctypes.c_int] * 3 + [
            ctypes.c_void_p]
        _lib.eso_fft2.restype = ctypes.c_int
        _lib.eso_fft2.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                  ctypes.c_int]
        _lib.eso_max_threads.restype = ctypes.c_int
    return _lib


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def fft2(a, sign, nthreads=0):
    """Unnormalised 2-D FFT with the oracle's own C FFT (for its unit test)."""
    a = np.ascontiguousarray(a, dtype=np.complex128).copy()
    assert a.ndim == 2 and a.shape[0] == a.shape[1]
    assert lib().eso_fft2(_ptr(a), a.shape[0], sign, nthreads) == 0
    return a


class OracleGridder:
    """Drop-in shaped like the reference's Python class, on numpy, in fp64.

    ``is_double`` chooses the plan branch the way the visibility precision
    does in the reference (False: supports up to 8); inputs of either precision
    are promoted to fp64 and all arithmetic is fp64.
    """

    def __init__(self, uvw, freq_hz, vis, weight, dirty_image,
                 pixel_size_x_rad, pixel_size_y_rad, epsilon, do_w_stacking,
                 is_double=None):
        assert pixel_size_x_rad == pixel_size_y_rad
        if is_double is None:
            is_double = np.asarray(vis).dtype == np.complex128
        if do_w_stacking:
            # same arithmetic as the Python wrapper: in the arrays' own dtype
            min_abs_w, max_abs_w = get_w_range(np.asarray(uvw),
                                               np.asarray(freq_hz))
        else:
            min_abs_w = max_abs_w = 0.0
        self.num_rows, self.num_chan = np.asarray(vis).shape
        self.plan = make_plan(np.asarray(dirty_image).shape[0],
                              pixel_size_x_rad, epsilon, do_w_stacking,
                              min_abs_w, max_abs_w, is_double)

    def _params(self):
        pl = self.plan
        keep = [np.ascontiguousarray(pl[k], dtype=np.float64) for k in
                ("quad_kernel", "quad_nodes", "quad_weights", "conv_corr")]
        prm = _Params(
            self.num_rows, self.num_chan, pl["image_size"], pl["grid_size"],
            pl["support"], pl["num_total_w_grids"], pl["do_wstacking"],
            pl["num_quad"], pl["beta"], pl["pixel_size"], pl["uv_scale"],
            pl["w_scale"], pl["inv_w_scale"], pl["min_plane_w"],
            pl["conv_corr_norm_factor"], *[a.ctypes.data for a in keep])
        return prm, keep

    def grid_uvw_es_fft(self, uvw, freq_hz, vis, weight, dirty_image=None,
                        planes=(0, -1), nthreads=0):
        """Returns (dirty fp64 [N,N], timings[4]); dirty_image (if given) is the
        initial content that is accumulated into, as in the reference."""
        n = self.plan["image_size"]
        out = (np.zeros((n, n)) if dirty_image is None else
               np.array(dirty_image, dtype=np.float64))
        uvw = np.ascontiguousarray(uvw, dtype=np.float64)
        freq_hz = np.ascontiguousarray(freq_hz, dtype=np.float64)
        vis = np.ascontiguousarray(vis, dtype=np.complex128)
        weight = np.ascontiguousarray(weight, dtype=np.float64)
        prm, keep = self._params()
        t = np.zeros(4)
        rc = lib().eso_grid(ctypes.addressof(prm), _ptr(uvw), _ptr(freq_hz),
                            _ptr(vis), _ptr(weight), _ptr(out), planes[0],
                            planes[1], nthreads, _ptr(t))
        assert rc == 0
        del keep
        return out, t

    def ifft_grid_uvw_es(self, uvw, freq_hz, dirty_image, vis=None,
                         planes=(0, -1), nthreads=0):
        """Returns (vis complex128 [R,C], corrected image, timings[4])."""
        out = (np.zeros((self.num_rows, self.num_chan), dtype=np.complex128)
               if vis is None else np.array(vis, dtype=np.complex128))
        img = np.array(dirty_image, dtype=np.float64)
        uvw = np.ascontiguousarray(uvw, dtype=np.float64)
        freq_hz = np.ascontiguousarray(freq_hz, dtype=np.float64)
        prm, keep = self._params()
        t = np.zeros(4)
        rc = lib().eso_degrid(ctypes.addressof(prm), _ptr(uvw), _ptr(freq_hz),
                              _ptr(out), _ptr(img), planes[0], planes[1],
                              nthreads, _ptr(t))
        assert rc == 0
        del keep
        return out, img, t
