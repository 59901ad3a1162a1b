"""uv-grid visibility weighting (sdp_weighting_uniform / sdp_weighting_briggs,
SURVEY 8(f) #1): the oracle against outputs of the unmodified reference
library's CPU path (golden), the argument contract (error codes and check
order of sdp_weighting.cpp:282-336, no GPU needed), and -- on the B200 -- the
CUDA kernels against the oracle through the C-ABI, with host and
device-resident arrays, in the shapes of the reference's own tests
(tests/visibility/test_weighting.py:18-31,352-434)."""
import os

import numpy as np
import pytest

from oracle import weighting_oracle as wo
from tests.synth_weighting import weighting_case
from ska_sdp_func.utility import Lib
from ska_sdp_func.utility.error_checking import CError
from ska_sdp_func.visibility import briggs_weights, get_uv_range, uniform_weights

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = {
    "small_f64": dict(seed=1, single=False),
    "small_f32": dict(seed=2, single=True),
    "pols4_f64": dict(seed=3, single=False, num_pols=4, grid_size=64),
    "offgrid_f64": dict(seed=4, single=False, shrink=0.6, positive=True),
}
# fp64: the kernels' sums run in another order than the host loops (1e-16);
# fp32: one ulp where fused multiply-adds round the Briggs denominator once
RTOL = {np.dtype(np.float64): 1e-13, np.dtype(np.float32): 2.5e-7}


def run_oracle(c, kind):
    grid = np.zeros_like(c["grid"])
    out = c["output0"].copy()
    if kind == "uniform":
        wo.uniform(c["uvw"], c["freq_hz"], c["max_abs_uv"], grid, c["input"], out)
    else:
        wo.briggs(c["uvw"], c["freq_hz"], c["max_abs_uv"], c["robust"], grid,
                  c["input"], out)
    return grid, out


@pytest.mark.parametrize("name", sorted(CASES))
@pytest.mark.parametrize("kind", ["uniform", "briggs"])
def test_oracle_equals_the_reference_cpu_path(name, kind):
    z = np.load(os.path.join(ROOT, "tests", "golden", "golden_weighting.npz"))
    c = weighting_case(**CASES[name])
    grid, out = run_oracle(c, kind)
    g, o = z[f"{name}_{kind}_grid"], z[f"{name}_{kind}_out"]
    assert np.array_equal(grid, g)  # same additions in the same order
    tol = 0.0 if out.dtype == np.float64 else 2.5e-7
    assert np.max(np.abs(out - o) / np.abs(o)) <= tol
    if name == "offgrid_f64":
        # visibilities outside the grid keep the caller's output value
        assert np.sum(o == 7.0) > 0 and np.array_equal(o == 7.0, out == 7.0)


def test_library_exports_the_weighting_entry_points():
    lib = Lib.handle()
    for name in ("sdp_weighting_uniform", "sdp_weighting_briggs"):
        assert hasattr(lib, name)


def test_get_uv_range_is_the_reference_formula():
    c = weighting_case(seed=5)
    want = np.amax(np.abs(c["uvw"][:, :, 0:1])) * (c["freq_hz"][-1] / 299792458.0)
    assert get_uv_range(c["uvw"], c["freq_hz"]) == want


def _call(c, **over):
    a = dict(uvw=c["uvw"], freq_hz=c["freq_hz"], grid=np.zeros_like(c["grid"]),
             input=c["input"], output=c["output0"].copy())
    a.update(over)
    uniform_weights(a["uvw"], a["freq_hz"], c["max_abs_uv"], a["grid"],
                    a["input"], a["output"])


@pytest.mark.parametrize("over, code", [
    (dict(input=np.zeros((3, 40, 5))), 1),                     # weights not 4-D: RUNTIME
    (dict(output=np.zeros((3, 40, 5, 1), dtype=np.complex128)), 3),  # complex weights
    (dict(uvw=np.zeros((3, 40, 3), dtype=np.complex128)), 3),  # complex uvw
    (dict(uvw=np.zeros((3, 41, 3))), 2),                       # uvw shape
    (dict(uvw=np.zeros((3, 40, 6))[:, :, ::2]), 2),            # uvw not contiguous
    (dict(grid=np.zeros((32, 32))), 2),                        # grid not 3-D
    (dict(grid=np.zeros((32, 30, 1))), 2),                     # grid not square
    (dict(uvw=np.zeros((3, 40, 3), dtype=np.float32)), 3),     # fp32 coordinates
    (dict(grid=np.zeros((32, 32, 1), dtype=np.float32)), 3),   # grid type != weights type
    (dict(freq_hz=np.zeros(4)), 2),                            # channels mismatch
])
def test_argument_contract(over, code):
    """Checks and codes of sdp_weighting.cpp:297-336,342-399 (they run before
    any CUDA call, so this needs no GPU)."""
    c = weighting_case(seed=1)
    with pytest.raises(CError, match=f"Error {code}"):
        _call(c, **over)


def test_read_only_output_is_rejected():
    c = weighting_case(seed=1)
    out = c["output0"].copy()
    out.flags.writeable = False
    with pytest.raises(CError, match="Error 2"):
        _call(c, output=out)


# ------------------------------------------------------------------ GPU

@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
@pytest.mark.parametrize("kind", ["uniform", "briggs"])
def test_kernels_match_the_oracle_host_arrays(name, kind):
    c = weighting_case(**CASES[name])
    want_grid, want_out = run_oracle(c, kind)
    grid = np.zeros_like(c["grid"])
    out = c["output0"].copy()
    if kind == "uniform":
        uniform_weights(c["uvw"], c["freq_hz"], c["max_abs_uv"], grid, c["input"], out)
    else:
        briggs_weights(c["uvw"], c["freq_hz"], c["max_abs_uv"], c["robust"], grid,
                       c["input"], out)
    rtol = RTOL[out.dtype]
    assert np.allclose(grid, want_grid, rtol=rtol, atol=0)
    assert np.allclose(out, want_out, rtol=rtol, atol=0)
    assert np.array_equal(out == 7.0, want_out == 7.0)


@pytest.mark.gpu
@pytest.mark.parametrize("single", [False, True])
def test_device_resident_arrays_larger_case(single):
    """Arrays owned by torch on the device (any __cuda_array_interface__
    object is accepted), 2e5 visibilities x 2 polarisations, many hits per
    cell; nothing is copied."""
    import torch
    c = weighting_case(seed=9, single=single, num_times=4, num_baselines=500,
                       num_chan=100, num_pols=2, grid_size=128)
    for kind in ("uniform", "briggs"):
        want_grid, want_out = run_oracle(c, kind)
        dev = torch.device("cuda", 0)
        t = {k: torch.from_numpy(c[k]).to(dev) for k in ("uvw", "freq_hz", "input")}
        grid = torch.zeros(c["grid"].shape, dtype=t["input"].dtype, device=dev)
        out = torch.from_numpy(c["output0"]).to(dev)
        if kind == "uniform":
            uniform_weights(t["uvw"], t["freq_hz"], c["max_abs_uv"], grid,
                            t["input"], out)
        else:
            briggs_weights(t["uvw"], t["freq_hz"], c["max_abs_uv"], c["robust"],
                           grid, t["input"], out)
        torch.cuda.synchronize()
        # fp32 grid cells are sums of ~12 weights added in any order
        rtol = 1e-12 if not single else 2e-6
        assert np.allclose(grid.cpu().numpy(), want_grid, rtol=rtol, atol=0)
        assert np.allclose(out.cpu().numpy(), want_out, rtol=rtol, atol=0)


@pytest.mark.gpu
def test_reference_test_case_prime_3x3():
    """The reference's own test input (tests/visibility/test_weighting.py:18-31)
    against the loops its test restates (= the oracle)."""
    freqs = np.array([1e9, 1.1e9, 1.2e9])
    uvw = np.asarray([[[2, 3, 5], [7, 11, 13], [17, 19, 23]]], dtype=np.float64)
    inp = np.asarray([[[[10], [31], [21]]] * 3], dtype=np.float64)
    max_abs_uv = 16011.076569511299
    for kind, robust in (("uniform", None), ("briggs", -2.0)):
        want_grid = np.zeros((3, 3, 1))
        want = np.zeros_like(inp)
        grid = np.zeros((3, 3, 1))
        got = np.zeros_like(inp)
        if kind == "uniform":
            wo.uniform(uvw, freqs, max_abs_uv, want_grid, inp, want)
            uniform_weights(uvw, freqs, max_abs_uv, grid, inp, got)
        else:
            wo.briggs(uvw, freqs, max_abs_uv, robust, want_grid, inp, want)
            briggs_weights(uvw, freqs, max_abs_uv, robust, grid, inp, got)
        assert np.allclose(got, want) and np.allclose(grid, want_grid)


@pytest.mark.gpu
def test_location_mismatch_is_reported():
    import torch
    c = weighting_case(seed=1)
    d_uvw = torch.from_numpy(c["uvw"]).to("cuda:0")
    with pytest.raises(CError, match="Error 6"):
        _call(c, uvw=d_uvw)
