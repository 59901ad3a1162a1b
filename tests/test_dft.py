"""Point-source DFT predict (sdp_dft_point_v00 / sdp_dft_point_v01, SURVEY
8(f) #4): the oracle against outputs of the unmodified reference library's
CPU path (golden), the argument contract (error codes and check order of
sdp_dft.cpp:102-147,289-344, no GPU needed), and -- on the B200 -- the CUDA
kernels against the oracle through the C-ABI, with host and device-resident
arrays, in the shapes of the reference's own tests
(tests/visibility/test_dft.py:38-160), and against the reference's CUDA build.

Tolerances (max|out - oracle| / max|oracle|): 1e-10 for complex fp64
visibilities, 1e-5 for complex fp32 ones -- the north_star's tolerances for the
degridder, whose exact counterpart this is."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle import dft_oracle as do
from tests.synth_dft import dft_case
from ska_sdp_func.utility import Lib
from ska_sdp_func.utility.error_checking import CError
from ska_sdp_func.visibility import dft_point_v00, dft_point_v01

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN_CASES = {
    "pols4": dict(seed=1, num_baselines=30, num_times=2),
    "pols1": dict(seed=2, num_pols=1, num_channels=13, num_components=33,
                  num_baselines=40, num_times=2),
}
TOL = {np.dtype(np.complex128): 1e-10, np.dtype(np.complex64): 1e-5}


def rel_max(a, b):
    return float(np.max(np.abs(a.astype(np.complex128) - b)) / np.max(np.abs(b)))


@pytest.mark.parametrize("name", sorted(GOLDEN_CASES))
def test_oracle_equals_the_reference_cpu_path(name):
    z = np.load(os.path.join(ROOT, "tests", "golden", "golden_dft.npz"))
    c = dft_case(**GOLDEN_CASES[name])
    for exact in (False, True):
        v00 = do.dft_point_v00(c["directions"], c["fluxes"], c["uvw_lambda"],
                               exact=exact)
        v01 = do.dft_point_v01(c["directions"], c["fluxes"], c["uvw"],
                               c["channel_start_hz"], c["channel_step_hz"],
                               exact=exact)
        # literal order: libm vs numpy sin / cos, a few ulp; phase in turns:
        # the reference's rounding of a ~1e4 rad phase, ~1e-12 per term
        tol = 1e-14 if not exact else 2e-12
        assert rel_max(v00, z[f"{name}_v00_c128"]) <= tol
        assert rel_max(v01, z[f"{name}_v01_c128"]) <= tol
        assert rel_max(z[f"{name}_v01_c64"], v01) <= 2e-7


def test_library_exports_the_dft_entry_points():
    lib = Lib.handle()
    for name in ("sdp_dft_point_v00", "sdp_dft_point_v01"):
        assert hasattr(lib, name)


def _v01(c, **over):
    a = dict(directions=c["directions"], fluxes=c["fluxes"], uvw=c["uvw"],
             vis=np.zeros(c["vis_shape"], dtype=np.complex128))
    a.update(over)
    dft_point_v01(a["directions"], a["fluxes"], a["uvw"], c["channel_start_hz"],
                  c["channel_step_hz"], a["vis"])
    return a["vis"]


def _v00(c, **over):
    a = dict(directions=c["directions"], fluxes=c["fluxes"],
             uvw_lambda=c["uvw_lambda"],
             vis=np.zeros(c["vis_shape"], dtype=np.complex128))
    a.update(over)
    dft_point_v00(a["directions"], a["fluxes"], a["uvw_lambda"], a["vis"])
    return a["vis"]


@pytest.mark.parametrize("over, code", [
    (dict(vis=np.zeros((3, 351, 10, 4))), 3),                          # vis not complex
    (dict(vis=np.zeros((3, 351, 10), dtype=np.complex128)), 2),        # vis not 4-D
    (dict(vis=np.zeros((3, 351, 10, 2), dtype=np.complex128)), 2),     # 2 polarisations
    (dict(uvw=np.zeros((3, 351, 3), dtype=np.complex128)), 3),         # complex uvw
    (dict(uvw=np.zeros((3, 350, 3))), 2),                              # uvw shape
    (dict(uvw=np.zeros((3, 351, 6))[:, :, ::2]), 2),                   # uvw not contiguous
    (dict(directions=np.zeros((20, 2))), 2),                           # directions shape
    (dict(fluxes=np.zeros((20, 10, 4))), 3),                           # real fluxes
    (dict(fluxes=np.zeros((20, 9, 4), dtype=np.complex128)), 2),       # fluxes shape
    (dict(fluxes=np.zeros((20, 10, 4), dtype=np.complex64)), 3),       # fp32 fluxes
    (dict(uvw=np.zeros((3, 351, 3), dtype=np.float32)), 3),            # fp32 coordinates
])
def test_v01_argument_contract(over, code):
    c = dft_case(seed=1)
    with pytest.raises(CError, match=f"Error {code}"):
        _v01(c, **over)


@pytest.mark.parametrize("over, code", [
    (dict(vis=np.zeros((3, 351, 10, 4))), 3),
    (dict(uvw_lambda=np.zeros((3, 351, 10, 6))[..., ::2]), 2),         # not contiguous
    (dict(uvw_lambda=np.zeros((3, 351, 9, 3))), 2),                    # shape
    (dict(fluxes=np.zeros((20, 10, 4))), 3),
    (dict(directions=np.zeros((20, 3), dtype=np.float32)), 3),
])
def test_v00_argument_contract(over, code):
    c = dft_case(seed=1)
    with pytest.raises(CError, match=f"Error {code}"):
        _v00(c, **over)


def test_read_only_output_is_rejected():
    c = dft_case(seed=1)
    vis = np.zeros(c["vis_shape"], dtype=np.complex128)
    vis.flags.writeable = False
    with pytest.raises(CError, match="Error 2"):
        _v01(c, vis=vis)


# ------------------------------------------------------------------ GPU

CASES = {
    "reference_shape": dict(seed=3, num_times=10),     # test_dft.py:40-44
    "pols1_ragged": dict(seed=4, num_pols=1, num_channels=13, num_components=70,
                         num_baselines=130),
    "one_channel": dict(seed=5, num_channels=1, num_components=3),
    "long_baselines": dict(seed=6, num_pols=1, num_channels=20, max_uvw_m=6e4,
                           num_baselines=100),
}


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
@pytest.mark.parametrize("dtype", [np.complex128, np.complex64])
def test_kernels_match_the_oracle_host_arrays(name, dtype):
    c = dft_case(**CASES[name])
    want00 = do.dft_point_v00(c["directions"], c["fluxes"], c["uvw_lambda"],
                              exact=True)
    want01 = do.dft_point_v01(c["directions"], c["fluxes"], c["uvw"],
                              c["channel_start_hz"], c["channel_step_hz"],
                              exact=True)
    tol = TOL[np.dtype(dtype)]
    # the output is overwritten, not accumulated into
    got01 = _v01(c, vis=np.full(c["vis_shape"], 7 + 7j, dtype=dtype))
    got00 = _v00(c, vis=np.full(c["vis_shape"], 7 + 7j, dtype=dtype))
    errs = (rel_max(got01, want01), rel_max(got00, want00))
    print("dft", name, np.dtype(dtype).name, "v01 / v00 errors", errs)
    assert max(errs) <= tol
    # and against the literal (phase in radians) evaluation of the reference
    lit = do.dft_point_v01(c["directions"], c["fluxes"], c["uvw"],
                           c["channel_start_hz"], c["channel_step_hz"])
    assert rel_max(got01, lit) <= tol


@pytest.mark.gpu
def test_reference_test_inputs():
    """The reference's own tests (tests/visibility/test_dft.py:38-72,
    103-150): uniform random directions, fluxes and coordinates in [0, 1),
    against the formula its tests restate, to its 6 decimals and to 1e-10."""
    rng = np.random.default_rng(0)
    fluxes = rng.random((20, 10, 4)) + 0j
    directions = rng.random((20, 3))
    uvw_lambda = rng.random((10, 351, 10, 3))
    uvw = rng.random((10, 351, 3))
    vis = np.zeros((10, 351, 10, 4), dtype=np.complex128)
    dft_point_v00(directions, fluxes, uvw_lambda, vis)
    want = do.dft_point_v00(directions, fluxes, uvw_lambda)
    np.testing.assert_array_almost_equal(vis, want)
    assert rel_max(vis, want) <= 1e-10
    dft_point_v01(directions, fluxes, uvw, 100e6, 100e3, vis)
    want = do.dft_point_v01(directions, fluxes, uvw, 100e6, 100e3)
    np.testing.assert_array_almost_equal(vis, want)
    assert rel_max(vis, want) <= 1e-10


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.complex128, np.complex64])
def test_device_resident_arrays_larger_case(dtype):
    """Arrays owned by torch on the device; 1.3e5 visibilities x 300 sources;
    nothing is copied, the call does not synchronise."""
    import torch
    c = dft_case(seed=7, num_components=300, num_pols=4, num_channels=32,
                 num_baselines=1024, num_times=4)
    want = do.dft_point_v01(c["directions"], c["fluxes"], c["uvw"],
                            c["channel_start_hz"], c["channel_step_hz"],
                            exact=True)
    dev = torch.device("cuda", 0)
    t = {k: torch.from_numpy(c[k]).to(dev) for k in ("directions", "fluxes", "uvw")}
    vis = torch.full(c["vis_shape"], 7.0, device=dev,
                     dtype=torch.complex128 if dtype == np.complex128
                     else torch.complex64)
    dft_point_v01(t["directions"], t["fluxes"], t["uvw"], c["channel_start_hz"],
                  c["channel_step_hz"], vis)
    torch.cuda.synchronize()
    err = rel_max(vis.cpu().numpy(), want)
    print("dft device-resident", np.dtype(dtype).name, err)
    assert err <= TOL[np.dtype(dtype)]


@pytest.mark.gpu
def test_location_mismatch_is_reported():
    import torch
    c = dft_case(seed=1)
    with pytest.raises(CError, match="Error 6"):
        _v01(c, uvw=torch.from_numpy(c["uvw"]).to("cuda:0"))


@pytest.mark.gpu
def test_against_the_reference_cuda_build(tmp_path):
    """The unmodified reference library (baseline/_ref, sm_100 cubins) through
    its own Python functions on the same inputs, in its own process; prints
    both sides' device times."""
    ref = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.exists(os.path.join(ref, "libska_sdp_func.so")):
        pytest.skip("baseline/_ref not installed (baseline/build_ref_gpu.sh)")
    kw = dict(seed=8, num_components=200, num_pols=4, num_channels=16,
              num_baselines=500, num_times=2)
    path = str(tmp_path / "ref_dft.npz")
    r = subprocess.run(
        [sys.executable, os.path.join(ROOT, "baseline", "ref_dft.py"),
         "--case", json.dumps(kw), "--out", path, "--steps", "2"],
        capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    res = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert res["available"], res
    z = np.load(path)
    c = dft_case(**kw)
    for dtype, tag in ((np.complex128, "c128"), (np.complex64, "c64")):
        got01 = _v01(c, vis=np.zeros(c["vis_shape"], dtype=dtype))
        got00 = _v00(c, vis=np.zeros(c["vis_shape"], dtype=dtype))
        errs = (rel_max(got01, z[f"v01_{tag}"].astype(np.complex128)),
                rel_max(got00, z[f"v00_{tag}"].astype(np.complex128)))
        print("dft new-vs-refgpu", tag, errs)
        assert max(errs) <= TOL[np.dtype(dtype)]
    print("reference CUDA build, device ms:", res)
