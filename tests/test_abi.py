"""The drop-in boundary without a GPU: the shared library loads, exports every
symbol include/*.h declares, and honours the reference's argument contract
(same status codes in the same order as es_fft.cpp:72-262; the reference pins
these with 30 constructor-misuse cases in
tests/grid_data/test_gridder_uvw_es_fft.py:71-431 -- there on cupy arrays,
here on numpy arrays: validation happens before anything touches a device).
"""
import ctypes
import glob
import os
import re

import numpy as np
import pytest

from ska_sdp_func.grid_data import GridderUvwEsFft
from ska_sdp_func.utility import CError, Lib, Mem
from tests.conftest import HAVE_GPU

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = set()
    for header in glob.glob(os.path.join(ROOT, "include", "**", "*.h"),
                            recursive=True):
        text = open(header).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names.update(re.findall(r"\b(sdp_[a-z0-9_]+)\s*\(", text))
    names.discard("sdp_log_message")  # variadic, checked separately
    return sorted(names)


def test_library_loads_and_exports_every_declared_symbol():
    lib = Lib.handle()
    assert os.path.basename(Lib.lib_path).startswith("libska_sdp_func")
    missing = [n for n in declared_symbols() + ["sdp_log_message"]
               if not hasattr(lib, n)]
    assert not missing, missing
    assert len(declared_symbols()) >= 40


def test_library_has_sm100a_code_and_no_fallback_symbols():
    """The product path is CUDA only: the .so carries sm_100a SASS."""
    import subprocess
    out = subprocess.run(["cuobjdump", "--list-elf", Lib.lib_path],
                         capture_output=True, text=True, check=False).stdout
    if out:  # cuobjdump present
        assert "sm_100a" in out


def make_arrays(real=np.float64, rows=50, chans=4, n=64):
    cplx = np.complex128 if real == np.float64 else np.complex64
    return dict(
        uvw=np.zeros((rows, 3), dtype=real),
        freq=np.linspace(1e9, 1.1e9, chans).astype(real),
        vis=np.zeros((rows, chans), dtype=cplx),
        weight=np.ones((rows, chans), dtype=real),
        image=np.zeros((n, n), dtype=real))


def create(a, pix_x=1e-4, pix_y=1e-4, eps=1e-5, wstack=False, **override):
    b = dict(a)
    b.update(override)
    return GridderUvwEsFft(b["uvw"], b["freq"], b["vis"], b["weight"],
                           b["image"], pix_x, pix_y, eps, wstack)


TYPE_ERR = "Unsupported data type"
ARG_ERR = "Invalid function argument"


@pytest.mark.parametrize("override,message", [
    (dict(uvw=np.zeros((50, 3), dtype=np.complex128)), TYPE_ERR),
    (dict(freq=np.zeros(4, dtype=np.complex128)), TYPE_ERR),
    (dict(vis=np.zeros((50, 4), dtype=np.float64)), TYPE_ERR),
    (dict(weight=np.zeros((50, 4), dtype=np.complex128)), TYPE_ERR),
    (dict(image=np.zeros((64, 64), dtype=np.complex128)), TYPE_ERR),
    (dict(uvw=np.zeros((51, 3))), ARG_ERR),
    (dict(uvw=np.zeros((50, 4))), ARG_ERR),
    (dict(freq=np.zeros(5)), ARG_ERR),
    (dict(weight=np.zeros((50, 5))), ARG_ERR),
    (dict(weight=np.zeros((49, 4))), ARG_ERR),
    (dict(image=np.zeros((64, 65))), ARG_ERR),
    (dict(uvw=np.zeros((50, 3), dtype=np.float32)), TYPE_ERR),
    (dict(freq=np.zeros(4, dtype=np.float32)), TYPE_ERR),
    (dict(vis=np.zeros((50, 4), dtype=np.complex64)), TYPE_ERR),
    (dict(weight=np.zeros((50, 4), dtype=np.float32)), TYPE_ERR),
    (dict(image=np.zeros((64, 64), dtype=np.float32)), TYPE_ERR),
    (dict(uvw=np.zeros((3, 50)).T), ARG_ERR),               # not C contiguous
    (dict(vis=np.zeros((4, 50), dtype=complex).T), ARG_ERR),
    (dict(image=np.zeros((64, 128))[:, ::2]), ARG_ERR),
])
def test_create_plan_rejects_misuse_with_reference_codes(override, message):
    with pytest.raises(CError, match=message):
        create(make_arrays(), **override)


def test_create_plan_single_precision_family_checks():
    a = make_arrays(np.float32)
    with pytest.raises(CError, match=TYPE_ERR):
        create(a, vis=np.zeros((50, 4), dtype=np.complex128))
    with pytest.raises(CError, match=TYPE_ERR):
        create(a, image=np.zeros((64, 64)))


def test_pixel_sizes_must_match():
    with pytest.raises(CError, match=ARG_ERR):
        create(make_arrays(), pix_x=1e-4, pix_y=2e-4)


def test_read_only_image_rejected():
    a = make_arrays()
    a["image"].setflags(write=False)
    with pytest.raises(CError, match=ARG_ERR):
        create(a)


def test_check_order_location_then_type_then_shape():
    """First failing check wins: kinds (3) are tested before shapes (2),
    precision (3) after shapes."""
    a = make_arrays()
    with pytest.raises(CError, match=TYPE_ERR):   # complex uvw AND bad shape
        create(a, uvw=np.zeros((7, 2), dtype=complex))
    with pytest.raises(CError, match=ARG_ERR):    # bad shape AND mixed precision
        create(a, freq=np.zeros(9, dtype=np.float32))


@pytest.mark.skipif(HAVE_GPU, reason="only meaningful without a device")
def test_no_device_means_allocation_failure_not_a_cpu_fallback():
    """Without a GPU a valid request fails like the reference does
    (GPU allocation failure, code 4) -- there is no CPU path."""
    with pytest.raises(CError, match="Memory allocation failure"):
        create(make_arrays())


def test_get_w_range_numpy_and_bad_input():
    """tests/grid_data/test_gridder_uvw_es_fft.py:434-497."""
    np.random.seed(40)
    pix = 2 * np.pi / 180 / 1024
    freq = 1e9 + np.arange(10) * 1e8
    uvw = (np.random.rand(1000, 3) - 0.5) / (pix * 1e9 / 299792458.0)
    lo = np.amin(np.abs(uvw[:, 2])) * freq[0] / 299792458.0
    hi = np.amax(np.abs(uvw[:, 2])) * freq[-1] / 299792458.0
    got_lo, got_hi = GridderUvwEsFft.get_w_range(uvw, freq)
    assert abs(got_lo - lo) / lo < 1e-15 and abs(got_hi - hi) / hi < 1e-15
    assert GridderUvwEsFft.get_w_range(None, None) == (-1, -1)
    lo32, _ = GridderUvwEsFft.get_w_range(uvw.astype(np.float32),
                                          freq.astype(np.float32))
    assert isinstance(lo32, np.float32)


def test_mem_wrapper_semantics():
    lib = Lib.handle()
    lib.sdp_mem_num_elements.restype = ctypes.c_int64
    lib.sdp_mem_shape_dim.restype = ctypes.c_int64
    lib.sdp_mem_stride_bytes_dim.restype = ctypes.c_int64
    for fn in ("sdp_mem_num_elements", "sdp_mem_is_c_contiguous",
               "sdp_mem_is_read_only", "sdp_mem_is_complex", "sdp_mem_type",
               "sdp_mem_location", "sdp_mem_is_double"):
        getattr(lib, fn).argtypes = [Mem.handle_type()]
    lib.sdp_mem_shape_dim.argtypes = [Mem.handle_type(), ctypes.c_int32]
    lib.sdp_mem_stride_bytes_dim.argtypes = [Mem.handle_type(), ctypes.c_int32]
    a = np.zeros((6, 5), dtype=np.complex64)
    m = Mem(a)
    assert lib.sdp_mem_num_elements(m) == 30
    assert lib.sdp_mem_type(m) == 36 and lib.sdp_mem_location(m) == 0
    assert lib.sdp_mem_is_c_contiguous(m) == 1 and lib.sdp_mem_is_complex(m) == 1
    assert lib.sdp_mem_shape_dim(m, 1) == 5 and lib.sdp_mem_shape_dim(m, 2) == 0
    assert lib.sdp_mem_stride_bytes_dim(m, 0) == 40
    t = Mem(a.T)
    assert lib.sdp_mem_is_c_contiguous(t) == 0
    a.setflags(write=False)
    assert lib.sdp_mem_is_read_only(Mem(a)) == 1
    empty = Mem()
    assert lib.sdp_mem_num_elements(empty) == 0
    assert lib.sdp_mem_is_read_only(empty) == 1   # no data => read-only
    with pytest.raises(TypeError):
        Mem(np.zeros(3, dtype=np.int64))
    with pytest.raises(TypeError):
        Mem([1, 2, 3])


def test_mem_create_copy_clear_on_host():
    lib = Lib.handle()
    lib.sdp_mem_create.restype = ctypes.c_void_p
    lib.sdp_mem_create.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int32,
                                   ctypes.POINTER(ctypes.c_int64),
                                   ctypes.POINTER(ctypes.c_int)]
    lib.sdp_mem_data.restype = ctypes.c_void_p
    lib.sdp_mem_data.argtypes = [ctypes.c_void_p]
    lib.sdp_mem_free.argtypes = [ctypes.c_void_p]
    lib.sdp_mem_random_fill.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int)]
    lib.sdp_mem_clear_contents.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int)]
    shape = (ctypes.c_int64 * 2)(4, 3)
    status = ctypes.c_int(0)
    m = lib.sdp_mem_create(8, 0, 2, shape, ctypes.byref(status))
    assert status.value == 0 and m
    buf = (ctypes.c_double * 12).from_address(lib.sdp_mem_data(m))
    assert all(v == 0.0 for v in buf)        # host memory is zero-filled
    lib.sdp_mem_random_fill(m, ctypes.byref(status))
    assert all(0.0 <= v <= 1.0 for v in buf) and any(v > 0 for v in buf)
    lib.sdp_mem_clear_contents(m, ctypes.byref(status))
    assert all(v == 0.0 for v in buf)
    lib.sdp_mem_free(m)
    bad = ctypes.c_int(0)
    m = lib.sdp_mem_create(3, 0, 2, shape, ctypes.byref(bad))  # not a type
    assert bad.value == 3
    lib.sdp_mem_free(m)


def test_status_in_out_convention():
    """A non-zero status on entry makes the call a no-op (es_fft.cpp:291)."""
    lib = Lib.handle()
    fn = lib.sdp_gridder_uvw_es_fft_create_plan
    a = make_arrays()
    mems = [Mem(a[k]) for k in ("uvw", "freq", "vis", "weight", "image")]
    status = ctypes.c_int(5)
    raw = ctypes.CDLL(Lib.lib_path).sdp_gridder_uvw_es_fft_create_plan
    raw.restype = ctypes.c_void_p
    raw.argtypes = [Mem.handle_type()] * 5 + [ctypes.c_double] * 5 + [
        ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
    plan = raw(*[m._as_parameter_ for m in mems], 1e-4, 1e-4, 1e-5, 0.0, 0.0, 0,
               ctypes.byref(status))
    assert plan is None and status.value == 5
    del fn
    # exec with a NULL plan is a silent no-op (es_fft.cpp:543,756)
    grid = ctypes.CDLL(Lib.lib_path).sdp_grid_uvw_es_fft
    grid.argtypes = [ctypes.c_void_p] + [Mem.handle_type()] * 5 + [
        ctypes.POINTER(ctypes.c_int)]
    ok = ctypes.c_int(0)
    grid(None, *[m._as_parameter_ for m in mems], ctypes.byref(ok))
    assert ok.value == 0
    ctypes.CDLL(Lib.lib_path).sdp_gridder_uvw_es_fft_free_plan(None)
