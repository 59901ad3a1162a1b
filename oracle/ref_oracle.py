"""TEST INFRASTRUCTURE ONLY (oracle) -- never imported by the product package.

ctypes front-end to ``oracle/_ref/libes_ref_host.so``: the reference's own
kernel source (``sdp_gridder_uvw_es_fft_kernels.cu``), planning maths
(``..._es_fft_utils.cpp``) and CPU FFT backend (vendored PocketFFT) compiled
for the host, driven by ``oracle/ref_host/ref_driver.cpp`` (a restatement of
``sdp_gridder_uvw_es_fft.cpp:301-424,578-741,789-952``).

The library is built in the authoring container by ``make -C oracle ref``
(needs /root/reference) and travels to the GPU box as a prebuilt ``.so``;
nothing here reads /root/reference at run time.
"""

import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libes_ref_host.so")

C_0 = 299792458.0


class PlanInfo(ctypes.Structure):
    """Mirror of oref_PlanInfo (ref_driver.cpp)."""

    _fields_ = [
        ("grid_size", ctypes.c_int),
        ("support", ctypes.c_int),
        ("num_total_w_grids", ctypes.c_int),
        ("image_size", ctypes.c_int),
        ("beta", ctypes.c_double),
        ("pixel_size", ctypes.c_double),
        ("uv_scale", ctypes.c_double),
        ("w_scale", ctypes.c_double),
        ("inv_w_scale", ctypes.c_double),
        ("min_plane_w", ctypes.c_double),
        ("max_plane_w", ctypes.c_double),
        ("conv_corr_norm_factor", ctypes.c_double),
        ("min_abs_w", ctypes.c_double),
        ("max_abs_w", ctypes.c_double),
        ("inv_w_range", ctypes.c_double),
    ]

    def as_dict(self):
        return {name: getattr(self, name) for name, _ in self._fields_}


_lib = None


def available():
    """True if the prebuilt reference-host oracle is present."""
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.oref_create.restype = ctypes.c_void_p
        _lib.oref_create.argtypes = [ctypes.c_int] * 5 + [ctypes.c_double] * 4 + [
            ctypes.c_int
        ]
        _lib.oref_info.argtypes = [ctypes.c_void_p, ctypes.POINTER(PlanInfo)]
        _lib.oref_tables.argtypes = [ctypes.c_void_p] * 5
        _lib.oref_free.argtypes = [ctypes.c_void_p]
        for name in ("oref_grid", "oref_degrid"):
            fn = getattr(_lib, name)
            fn.restype = ctypes.c_int
            fn.argtypes = [ctypes.c_void_p] * 6 + [ctypes.c_int] * 4 + [
                ctypes.c_void_p
            ]
        _lib.oref_max_threads.restype = ctypes.c_int
    return _lib


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


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


class RefHostGridder:
    """The reference path on the CPU.

    ``plan_prec`` (4|8) picks the KernelDB branch the way the visibility
    precision does in the reference; ``compute_prec`` (4|8) picks the kernel
    instantiations. (4, 4) and (8, 8) are the two literal reference paths;
    (4, 8) is "oracle64": the fp32-selected plan evaluated in fp64.
    """

    def __init__(self, num_rows, num_chan, image_size, pixel_size, epsilon,
                 do_w_stacking, min_abs_w=0.0, max_abs_w=0.0, plan_prec=8,
                 compute_prec=8):
        if not do_w_stacking:
            min_abs_w = max_abs_w = 0.0
        self.compute_prec = compute_prec
        self.shape = (num_rows, num_chan, image_size)
        self._h = lib().oref_create(
            num_rows, num_chan, image_size, plan_prec, compute_prec,
            pixel_size, epsilon, min_abs_w, max_abs_w, int(bool(do_w_stacking)))
        info = PlanInfo()
        lib().oref_info(self._h, ctypes.byref(info))
        self.info = info.as_dict()

    def tables(self):
        n = self.shape[2] // 2 + 1
        qk, qn, qw = (np.zeros(32) for _ in range(3))
        cc = np.zeros(n)
        lib().oref_tables(self._h, _ptr(qk), _ptr(qn), _ptr(qw), _ptr(cc))
        return qk, qn, qw, cc

    def _cast(self, uvw, freq_hz, vis, weight, image):
        r = np.float64 if self.compute_prec == 8 else np.float32
        c = np.complex128 if self.compute_prec == 8 else np.complex64
        return (np.ascontiguousarray(uvw, dtype=r),
                np.ascontiguousarray(freq_hz, dtype=r),
                np.ascontiguousarray(vis, dtype=c),
                np.ascontiguousarray(weight, dtype=r),
                np.ascontiguousarray(image, dtype=r))

    def grid(self, uvw, freq_hz, vis, weight, dirty=None, planes=(0, -1),
             fast=True, nthreads=0):
        """Returns (dirty_image, timings[4])."""
        n = self.shape[2]
        if dirty is None:
            dirty = np.zeros((n, n))
        uvw, freq_hz, vis, weight, dirty = self._cast(
            uvw, freq_hz, vis, weight, dirty)
        dirty = dirty.copy()
        t = np.zeros(4)
        lib().oref_grid(self._h, _ptr(uvw), _ptr(freq_hz), _ptr(vis),
                        _ptr(weight), _ptr(dirty), planes[0], planes[1],
                        int(fast), nthreads, _ptr(t))
        return dirty, t

    def degrid(self, uvw, freq_hz, image, vis=None, weight=None,
               planes=(0, -1), fast=True, nthreads=0):
        """Returns (vis, corrected_image, timings[4]); image is not modified."""
        nr, nc, _ = self.shape
        if vis is None:
            vis = np.zeros((nr, nc), dtype=np.complex128)
        if weight is None:
            weight = np.ones((nr, nc))
        uvw, freq_hz, vis, weight, image = self._cast(
            uvw, freq_hz, vis, weight, image)
        vis = vis.copy()
        image = image.copy()
        t = np.zeros(4)
        lib().oref_degrid(self._h, _ptr(uvw), _ptr(freq_hz), _ptr(vis),
                          _ptr(weight), _ptr(image), planes[0], planes[1],
                          int(fast), nthreads, _ptr(t))
        return vis, image, t

    def __del__(self):
        try:
            lib().oref_free(self._h)
        except Exception:  # pylint: disable=broad-except
            pass
