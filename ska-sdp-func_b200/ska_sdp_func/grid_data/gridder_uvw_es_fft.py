"""Python front-end of the w-stacking ES gridder / degridder.

Mirrors the reference's ``src/ska_sdp_func/grid_data/gridder_uvw_es_fft.py``
(class name, constructor arguments, ``get_w_range``, ``grid_uvw_es_fft`` and
``ifft_grid_uvw_es``) on top of the B200-native ``libska_sdp_func.so``.
Arrays may be numpy arrays (staged through the device inside the call -- an
extension, the reference only accepts device arrays) or device arrays exposing
``__cuda_array_interface__`` (cupy, torch, numba ...; processed in place on the
legacy default stream without synchronising, like the reference).
"""

import ctypes

import numpy as np

from ..utility import Lib, Mem, StructWrapper

SPEED_OF_LIGHT = 299792458.0


class EsPlanInfo(ctypes.Structure):
    """sdp_EsPlanInfo (include/.../sdp_gridder_uvw_es_fft_b200.h)."""

    _fields_ = [
        ("num_rows", ctypes.c_int32),
        ("num_chan", ctypes.c_int32),
        ("image_size", ctypes.c_int32),
        ("grid_size", ctypes.c_int32),
        ("support", ctypes.c_int32),
        ("num_total_w_grids", ctypes.c_int32),
        ("do_wstacking", ctypes.c_int32),
        ("is_double", ctypes.c_int32),
        ("epsilon", ctypes.c_double),
        ("pixel_size", ctypes.c_double),
        ("beta", ctypes.c_double),
        ("uv_scale", ctypes.c_double),
        ("min_abs_w", ctypes.c_double),
        ("max_abs_w", ctypes.c_double),
        ("min_plane_w", ctypes.c_double),
        ("max_plane_w", ctypes.c_double),
        ("w_scale", ctypes.c_double),
        ("inv_w_scale", ctypes.c_double),
        ("inv_w_range", ctypes.c_double),
        ("conv_corr_norm_factor", ctypes.c_double),
    ]

    def as_dict(self):
        """Plain dict of the fields."""
        return {name: getattr(self, name) for name, _ in self._fields_}


class EsCallStats(ctypes.Structure):
    """sdp_EsCallStats (include/.../sdp_gridder_uvw_es_fft_b200.h)."""

    _fields_ = [
        ("kernel_launches", ctypes.c_int64),
        ("fft_execs", ctypes.c_int64),
        ("planes_processed", ctypes.c_int64),
        ("num_entries", ctypes.c_int64),
        ("h2d_bytes", ctypes.c_int64),
        ("d2h_bytes", ctypes.c_int64),
        ("ms_total", ctypes.c_double),
        ("ms_presort", ctypes.c_double),
        ("ms_scatter_gather", ctypes.c_double),
        ("ms_fft", ctypes.c_double),
        ("ms_screen", ctypes.c_double),
        ("plane_batch", ctypes.c_int32),
        ("timing_enabled", ctypes.c_int32),
        ("pruned_fft", ctypes.c_int32),
        ("reserved_", ctypes.c_int32),
        ("occupied_tiles", ctypes.c_int64),
        ("occupied_tile_rows", ctypes.c_int64),
    ]

    def as_dict(self):
        """Plain dict of the fields."""
        return {name: getattr(self, name) for name, _ in self._fields_}


class GridderUvwEsFft(StructWrapper):
    """Processing function GridderUvwEsFft."""

    def __init__(
        self,
        uvw,
        freq_hz,
        vis,
        weight,
        dirty_image,
        pixel_size_x_rad,
        pixel_size_y_rad,
        epsilon: float,
        do_w_stacking: bool,
        w_range=None,
    ):
        """Create a plan for (de)gridding with the given buffers' shapes.

        uvw [R, 3] metres, freq_hz [C] Hz ascending, vis [R, C] complex (its
        precision selects fp32 or fp64 processing), weight [R, C],
        dirty_image [N, N] square; one precision and one location for all.
        epsilon is the requested accuracy; with do_w_stacking the full
        improved w-stacking algorithm is run, otherwise w is taken as zero.
        w_range (extension, default None = derive from uvw like the reference):
        (min |w|, max |w|) in wavelengths, for callers that need one global
        w-plane lattice across several plans (multi-GPU).
        """
        if do_w_stacking and w_range is not None:
            min_abs_w, max_abs_w = w_range
        elif do_w_stacking:
            min_abs_w, max_abs_w = GridderUvwEsFft.get_w_range(uvw, freq_hz)
        else:
            min_abs_w = 0
            max_abs_w = 0
        create_args = (
            Mem(uvw),
            Mem(freq_hz),
            Mem(vis),
            Mem(weight),
            Mem(dirty_image),
            pixel_size_x_rad,
            pixel_size_y_rad,
            epsilon,
            min_abs_w,
            max_abs_w,
            do_w_stacking,
        )
        super().__init__(
            Lib.sdp_gridder_uvw_es_fft_create_plan,
            create_args,
            Lib.sdp_gridder_uvw_es_fft_free_plan,
        )

    @staticmethod
    def get_w_range(uvw, freq_hz):
        """(min |w|, max |w|) in wavelengths: min at the first channel, max at
        the last one (freq_hz ascending). Arithmetic is done in the arrays'
        own precision, as the reference does (py:90-106)."""
        if isinstance(uvw, np.ndarray):
            min_abs_w = np.amin(np.abs(uvw[:, 2]))
            max_abs_w = np.amax(np.abs(uvw[:, 2]))
            first, last = freq_hz[0], freq_hz[-1]
        elif hasattr(uvw, "__cuda_array_interface__"):
            # cupy / torch / ...: reduce on the device, finish on the host in
            # the same precision
            dtype = np.dtype(uvw.__cuda_array_interface__["typestr"]).type
            abs_w = abs(uvw[:, 2])
            min_abs_w = dtype(float(abs_w.min()))
            max_abs_w = dtype(float(abs_w.max()))
            first = dtype(float(freq_hz[0]))
            last = dtype(float(freq_hz[-1]))
        else:
            print(f"Unsupported uvw type of {type(uvw)}.")
            return -1, -1
        min_abs_w = min_abs_w * (first / SPEED_OF_LIGHT)
        max_abs_w = max_abs_w * (last / SPEED_OF_LIGHT)
        return min_abs_w, max_abs_w

    def grid_uvw_es_fft(self, uvw, freq_hz, vis, weight, dirty_image):
        """Generate a dirty image from visibility data (accumulated into
        dirty_image, which is then grid-corrected: pass zeros)."""
        Lib.sdp_grid_uvw_es_fft(
            self,
            Mem(uvw),
            Mem(freq_hz),
            Mem(vis),
            Mem(weight),
            Mem(dirty_image),
        )

    def ifft_grid_uvw_es(self, uvw, freq_hz, vis, weight, dirty_image):
        """Predict visibilities from an image: vis += prediction; the grid
        correction is applied to dirty_image in place; weight is unused."""
        Lib.sdp_ifft_degrid_uvw_es(
            self,
            Mem(uvw),
            Mem(freq_hz),
            Mem(vis),
            Mem(weight),
            Mem(dirty_image),
        )

    # ---- B200 extensions (not in the reference) ----

    def info(self):
        """Planning scalars of this plan as a dict."""
        out = EsPlanInfo()
        Lib.sdp_gridder_uvw_es_fft_get_info(self, ctypes.byref(out))
        return out.as_dict()

    @staticmethod
    def plan_info(num_rows, num_chan, image_size, is_double, pixel_size_rad,
                  epsilon, min_abs_w, max_abs_w, do_w_stacking):
        """Host-only planning: (info dict, quad_kernel, quad_nodes,
        quad_weights, conv_corr). Touches no device."""
        out = EsPlanInfo()
        qk, qn, qw = (np.zeros(32) for _ in range(3))
        cc = np.zeros(image_size // 2 + 1)
        as_p = lambda a: a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))
        Lib.sdp_gridder_uvw_es_fft_plan_info(
            num_rows, num_chan, image_size, int(bool(is_double)),
            pixel_size_rad, epsilon, min_abs_w, max_abs_w,
            int(bool(do_w_stacking)), ctypes.byref(out), as_p(qk), as_p(qn),
            as_p(qw), as_p(cc))
        return out.as_dict(), qk, qn, qw, cc

    def set_plane_range(self, plane_begin, plane_end, apply_correction=True):
        """Own only w-planes [plane_begin, plane_end) of the global lattice."""
        Lib.sdp_gridder_uvw_es_fft_set_plane_range(
            self, plane_begin, plane_end, int(bool(apply_correction)))

    def correct_image(self, dirty_image):
        """Apply the grid correction to an image (after a multi-GPU reduce)."""
        Lib.sdp_gridder_uvw_es_fft_correct_image(self, Mem(dirty_image))

    def plane_histogram(self, uvw, freq_hz):
        """Visibilities per first w-plane, int64 [num_total_w_grids]."""
        counts = np.zeros(self.info()["num_total_w_grids"], dtype=np.int64)
        Lib.sdp_gridder_uvw_es_fft_plane_histogram(
            self, Mem(uvw), Mem(freq_hz),
            counts.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)))
        return counts

    def enable_timing(self, enable=True):
        """Record per-call device timings (adds a sync per call)."""
        Lib.sdp_gridder_uvw_es_fft_enable_timing(self, int(bool(enable)))

    def batch_times(self, capacity=65536):
        """(first_plane, num_planes, ms) arrays: device time of every w-plane
        batch of the last call (timing must be enabled)."""
        first = np.zeros(capacity, dtype=np.int32)
        count = np.zeros(capacity, dtype=np.int32)
        ms = np.zeros(capacity, dtype=np.float64)
        n = Lib.sdp_gridder_uvw_es_fft_get_batch_times(
            self, capacity,
            first.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)),
            count.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)),
            ms.ctypes.data_as(ctypes.POINTER(ctypes.c_double)))
        return first[:n].copy(), count[:n].copy(), ms[:n].copy()

    def stats(self):
        """Statistics of the last grid / degrid call as a dict."""
        out = EsCallStats()
        Lib.sdp_gridder_uvw_es_fft_get_stats(self, ctypes.byref(out))
        return out.as_dict()


_PLAN = GridderUvwEsFft.handle_type()
_MEM = Mem.handle_type()

Lib.wrap_func(
    "sdp_gridder_uvw_es_fft_create_plan",
    restype=_PLAN,
    argtypes=[_MEM] * 5 + [ctypes.c_double] * 5 + [ctypes.c_int],
    check_errcode=True,
)
Lib.wrap_func(
    "sdp_gridder_uvw_es_fft_free_plan", restype=None, argtypes=[_PLAN]
)
Lib.wrap_func(
    "sdp_grid_uvw_es_fft",
    restype=None,
    argtypes=[_PLAN] + [_MEM] * 5,
    check_errcode=True,
)
Lib.wrap_func(
    "sdp_ifft_degrid_uvw_es",
    restype=None,
    argtypes=[_PLAN] + [_MEM] * 5,
    check_errcode=True,
)
Lib.wrap_func(
    "sdp_gridder_uvw_es_fft_get_info",
    restype=None,
    argtypes=[_PLAN, ctypes.POINTER(EsPlanInfo)],
)
Lib.wrap_func(
    "sdp_gridder_uvw_es_fft_plan_info",
    restype=None,
    argtypes=[ctypes.c_int32] * 4 + [ctypes.c_double] * 4 + [ctypes.c_int32]
    + [ctypes.POINTER(EsPlanInfo)] + [ctypes.POINTER(ctypes.c_double)] * 4,
    check_errcode=True,
)
Lib.wrap_func(
    "sdp_gridder_uvw_es_fft_set_plane_range",
    restype=None,
    argtypes=[_PLAN, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32],
    check_errcode=True,
)
Lib.wrap_func(
    "sdp_gridder_uvw_es_fft_correct_image",
    restype=None,
    argtypes=[_PLAN, _MEM],
    check_errcode=True,
)
Lib.wrap_func(
    "sdp_gridder_uvw_es_fft_plane_histogram",
    restype=None,
    argtypes=[_PLAN, _MEM, _MEM, ctypes.POINTER(ctypes.c_int64)],
    check_errcode=True,
)
Lib.wrap_func(
    "sdp_gridder_uvw_es_fft_enable_timing",
    restype=None,
    argtypes=[_PLAN, ctypes.c_int32],
)
Lib.wrap_func(
    "sdp_gridder_uvw_es_fft_get_batch_times",
    restype=ctypes.c_int32,
    argtypes=[_PLAN, ctypes.c_int32, ctypes.POINTER(ctypes.c_int32),
              ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_double)],
)
Lib.wrap_func(
    "sdp_gridder_uvw_es_fft_get_stats",
    restype=None,
    argtypes=[_PLAN, ctypes.POINTER(EsCallStats)],
)
