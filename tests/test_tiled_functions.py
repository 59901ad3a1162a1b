"""Tiled visibility functions (sdp_count_and_prefix_sum / sdp_bucket_sort /
sdp_tiled_indexing): the oracle against outputs of the
unmodified reference library's CPU path (golden), the argument contract (no
GPU needed), and -- on the B200 -- the CUDA path against the oracle through the
C-ABI, bit for bit, with host and device-resident arrays, on the reference's
own test input (tests/visibility/test_tiled_functions.py:102-216) and against
the reference's CUDA build (baseline/_ref) tile by tile."""
import ctypes
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle import tiled_oracle as to
from tests.synth_tiled import geometry_args, num_tiles, tiled_case
from ska_sdp_func.utility import Lib
from ska_sdp_func.utility.error_checking import CError
from ska_sdp_func.visibility import (bucket_sort, count_and_prefix_sum,
                                     tiled_indexing)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# the cases of tests/golden/make_golden_tiled.py
GOLDEN_CASES = {
    "small_f64": dict(seed=1, positive=True),
    "centre_f64": dict(seed=2, positive=True, core=0.6, num_chan=4),
    "tile10_f64": dict(seed=3, positive=True, tile_size_u=10, tile_size_v=10,
                       grid_size=200, support=3, spread=0.4),
    "small_f32": dict(seed=4, positive=True, single=True),
}
# what the GPU tests run (both signs, off-grid visibilities, dense centre)
CASES = {
    "small_f64": dict(seed=11),
    "small_f32": dict(seed=12, single=True),
    "centre_f64": dict(seed=13, core=0.7, num_baselines=300, num_chan=16),
    "tile10_f32": dict(seed=14, single=True, tile_size_u=10, tile_size_v=10,
                       grid_size=200, support=3, spread=0.4),
    "wide_support_f64": dict(seed=15, support=20, grid_size=512, spread=0.45),
}


def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "golden_tiled.npz"))


@pytest.mark.parametrize("name", sorted(GOLDEN_CASES))
def test_oracle_equals_the_reference_cpu_path(name):
    """Bit for bit, in the mode that restates the reference's host branch."""
    z = golden()
    c = tiled_case(**GOLDEN_CASES[name])
    m, offsets, counts, skipped = to.count_and_prefix_sum(
        c["uvw"], c["freqs"], *geometry_args(c), host_branch_slip=True)
    if f"{name}_num" in z:
        assert m == int(z[f"{name}_num"])
        assert np.array_equal(offsets, z[f"{name}_offsets"])
        assert np.array_equal(counts, z[f"{name}_counts"])
        assert skipped == int(z[f"{name}_skipped"][0])
        assert skipped > 0 or GOLDEN_CASES[name].get("spread", 1) < 0.5
    else:
        assert np.array_equal(offsets, z[f"{name}_offsets_in"])
    got = to.bucket_sort(c["uvw"], c["freqs"], c["vis"], c["weights"],
                         *geometry_args(c), offsets, m, host_branch_slip=True)
    for k in ("uu", "vv", "vis", "weight", "tile"):
        assert np.array_equal(got[f"sorted_{k}"], z[f"{name}_sort_{k}"]), k
    assert np.array_equal(got["tile_offsets"], z[f"{name}_sort_offsets"])
    assert np.count_nonzero(got["sorted_tile"]) > 0.9 * m
    got = to.tiled_indexing(c["uvw"], c["freqs"], *geometry_args(c), offsets, m,
                            host_branch_slip=True)
    assert np.array_equal(got["sorted_tile"], z[f"{name}_index_tile"])
    assert np.array_equal(got["sorted_vis_index"], z[f"{name}_index_index"])
    assert np.array_equal(got["sorted_uu"], z[f"{name}_index_uu"])
    assert np.array_equal(got["tile_offsets"], z[f"{name}_index_offsets"])


def test_oracle_kernel_mode_differs_from_the_host_branch_only_in_u_min():
    """Without the slip every visibility lands in 1-4 tiles around its own
    position (what the reference's kernels and its tests compute)."""
    c = tiled_case(**GOLDEN_CASES["small_f64"])
    i_vis, tile, pu, pv, pos_u, _, skipped, geo = to.entries(
        c["uvw"], c["freqs"], *geometry_args(c))
    on_grid = c["uvw"].shape[0] * c["uvw"].shape[1] * c["freqs"].size - skipped
    assert on_grid <= tile.size <= 4 * on_grid
    grid_u = np.floor(np.abs(pos_u) + 0.5) * np.sign(pos_u) + geo["centre"]
    assert np.all(np.abs((pu + 0.5) * c["tile_size_u"] + geo["top_left_u"]
                         - grid_u) <= c["tile_size_u"] / 2 + c["support"] + 1)
    slip = to.entries(c["uvw"], c["freqs"], *geometry_args(c),
                      host_branch_slip=True)
    assert slip[1].size > tile.size and slip[6] == skipped


def test_library_exports_the_tiled_entry_points():
    lib = Lib.handle()
    for name in ("sdp_count_and_prefix_sum", "sdp_bucket_sort",
                 "sdp_tiled_indexing"):
        assert hasattr(lib, name)


def _count_args(c, **over):
    nt = num_tiles(c)
    a = dict(uvw=c["uvw"], freqs=c["freqs"], vis=c["vis"],
             offsets=np.zeros(nt + 1, dtype=np.int32),
             counts=np.zeros(nt, dtype=np.int32),
             skipped=np.zeros(1, dtype=np.int32))
    a.update(over)
    return a


def _count(c, **over):
    a = _count_args(c, **over)
    n = ctypes.c_int(0)
    count_and_prefix_sum(a["uvw"], a["freqs"], a["vis"], *geometry_args(c), n,
                         a["offsets"], a["counts"], a["skipped"])
    return n.value, a


@pytest.mark.parametrize("over, code", [
    (dict(vis=np.zeros((3, 50, 6, 1))), 3),                      # vis not complex
    (dict(vis=np.zeros((3, 50, 6), dtype=np.complex128)), 2),    # vis not 4-D
    (dict(vis=np.zeros((3, 50, 6, 2), dtype=np.complex128)), 2),  # 2 polarisations
    (dict(uvw=np.zeros((3, 50, 3), dtype=np.complex128)), 3),    # complex uvw
    (dict(uvw=np.zeros((3, 51, 3))), 2),                         # uvw shape
    (dict(uvw=np.zeros((3, 50, 6))[:, :, ::2]), 2),              # uvw not contiguous
    (dict(freqs=np.zeros(6, dtype=np.float32)), 3),              # uvw / freqs types differ
    (dict(freqs=np.zeros(5)), 2),                                # channels mismatch
    (dict(offsets=np.zeros(10, dtype=np.int32)), 2),             # offsets too short
    (dict(counts=np.zeros(10, dtype=np.int32)), 2),              # counts too short
    (dict(offsets=np.zeros(326, dtype=np.float32)), 3),          # offsets not int32
    (dict(skipped=np.zeros(0, dtype=np.int32)), 2),              # no room for the count
])
def test_count_argument_contract(over, code):
    """Checks of sdp_tiled_functions.cpp:345-363,414-428 (reference codes)
    plus the sizes the kernels index; they run before any CUDA call."""
    c = tiled_case(seed=1)
    with pytest.raises(CError, match=f"Error {code}"):
        _count(c, **over)


def _sort_arrays(c, m, complex_vis=False):
    real = c["uvw"].dtype
    s = dict(uu=np.zeros(m, dtype=real), vv=np.zeros(m, dtype=real),
             weight=np.zeros(m, dtype=real), tile=np.zeros(m, dtype=np.int32),
             svis=np.zeros(m, dtype=c["vis"].dtype if complex_vis else real))
    return s


def _sort(c, offsets, s, **over):
    a = dict(uvw=c["uvw"], freqs=c["freqs"], vis=c["vis"], weights=c["weights"])
    a.update(s)
    a.update(over)
    bucket_sort(a["uvw"], a["freqs"], a["vis"], a["weights"], *geometry_args(c),
                a["uu"], a["vv"], a["weight"], a["tile"], a["svis"], offsets)


@pytest.mark.parametrize("over, code", [
    (dict(weights=np.zeros((3, 50, 6, 1), dtype=np.complex128)), 3),  # complex weights
    (dict(weights=np.zeros((3, 50, 5, 1))), 2),                  # weights shape
    (dict(weights=np.zeros((3, 50, 6, 1), dtype=np.float32)), 3),  # weights type
    (dict(uu=np.zeros(100, dtype=np.float32)), 3),               # sorted_uu type
    (dict(tile=np.zeros(100, dtype=np.float32)), 3),             # sorted_tile type
    (dict(svis=np.zeros(100, dtype=np.complex64)), 3),           # sorted_vis precision
])
def test_sort_argument_contract(over, code):
    c = tiled_case(seed=1)
    offsets = np.zeros(num_tiles(c) + 1, dtype=np.int32)
    with pytest.raises(CError, match=f"Error {code}"):
        _sort(c, offsets, _sort_arrays(c, 100), **over)


def test_status_set_on_entry_returns_at_once():
    """.cpp:512,731 (and the count, which would otherwise clear its outputs)."""
    c = tiled_case(seed=1)
    lib = Lib.handle()
    status = ctypes.c_int(4)
    lib.sdp_tiled_indexing.restype = None
    lib.sdp_tiled_indexing(None, None, 400, ctypes.c_int64(32),
                           ctypes.c_int64(16), ctypes.c_double(1e-5),
                           ctypes.c_int64(4), ctypes.c_int64(1),
                           ctypes.c_int64(1), ctypes.c_int64(1), None, None,
                           None, None, None, ctypes.byref(status))
    assert status.value == 4 and c is not None


# ------------------------------------------------------------------ GPU

def _oracle_all(c, complex_vis=False):
    m, offsets, counts, skipped = to.count_and_prefix_sum(
        c["uvw"], c["freqs"], *geometry_args(c))
    sort = to.bucket_sort(c["uvw"], c["freqs"], c["vis"], c["weights"],
                          *geometry_args(c), offsets, m, complex_vis=complex_vis)
    index = to.tiled_indexing(c["uvw"], c["freqs"], *geometry_args(c), offsets, m)
    return m, offsets, counts, skipped, sort, index


def _index(c, offsets, m):
    real = c["uvw"].dtype
    t = dict(tile=np.zeros(m, dtype=np.int32), uu=np.zeros(m, dtype=real),
             vv=np.zeros(m, dtype=real), index=np.zeros(m, dtype=np.int32))
    tiled_indexing(c["uvw"], c["freqs"], *geometry_args(c), c["freqs"].shape[0],
                   c["uvw"].shape[1], c["uvw"].shape[0], t["tile"], t["uu"],
                   t["vv"], t["index"], offsets)
    return t


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_kernels_match_the_oracle_host_arrays(name):
    """Counts, offsets, skipped count, every sorted array and the advanced
    offsets, bit for bit (integer work; positions are the reference's
    expression in its precision)."""
    c = tiled_case(**CASES[name])
    m, offsets, counts, skipped, sort, index = _oracle_all(c)
    assert m > 0
    n, a = _count(c)
    assert n == m
    assert np.array_equal(a["offsets"], offsets)
    assert np.array_equal(a["counts"], counts)
    assert a["skipped"][0] == skipped
    s = _sort_arrays(c, m)
    offs = a["offsets"].copy()
    _sort(c, offs, s)
    for k, key in (("uu", "sorted_uu"), ("vv", "sorted_vv"), ("svis", "sorted_vis"),
                   ("weight", "sorted_weight"), ("tile", "sorted_tile")):
        assert np.array_equal(s[k], sort[key]), k
    assert np.array_equal(offs, sort["tile_offsets"])
    assert np.array_equal(offs[:-1], offsets[1:])  # each advanced to the next tile
    offs = a["offsets"].copy()
    t = _index(c, offs, m)
    assert np.array_equal(t["index"], index["sorted_vis_index"])
    assert np.array_equal(t["tile"], index["sorted_tile"])
    assert np.array_equal(t["uu"], index["sorted_uu"])
    assert np.array_equal(t["vv"], index["sorted_vv"])
    assert np.array_equal(offs, index["tile_offsets"])


@pytest.mark.gpu
def test_reference_test_input():
    """The reference's own test (tests/visibility/test_tiled_functions.py:
    102-216,219-298): 8 x 8 x 6 visibilities of one baseline vector, 400^2
    grid, 32 x 16 tiles, support 4, tile arrays sized for 10 x 10 tiles; its
    assertion (every visibility counted once) and the positions entry by
    entry."""
    freqs = np.array([1e9, 1.1e9, 1.2e9, 1.3e9, 1.4e9, 1.5e9])
    uvw = np.tile(np.array([10.0, 31.0, 21.0]), (8, 8, 1))
    vis = np.full((8, 8, 6, 1), 1j, dtype=complex)
    weights = np.ones(vis.shape)
    geo = (400, 32, 16, 4.06e-5, 4)
    nt10 = 40 * 40
    offsets = np.zeros(nt10 + 1, dtype=np.int32)
    counts = np.zeros(nt10, dtype=np.int32)
    skipped = np.zeros(1, dtype=np.int32)
    n = ctypes.c_int(0)
    count_and_prefix_sum(uvw, freqs, vis, *geo, n, offsets, counts, skipped)
    assert n.value == vis.size and skipped[0] == 0
    s = {k: np.zeros(n.value) for k in ("uu", "vv", "vis", "weight")}
    tile = np.zeros(n.value, dtype=np.int32)
    bucket_sort(uvw, freqs, vis, weights, *geo, s["uu"], s["vv"], s["weight"],
                tile, s["vis"], offsets)
    want_u = np.tile(10.0 * (freqs / to.C_0) * (400 * 4.06e-5), 64)
    want_v = np.tile(31.0 * (freqs / to.C_0) * (400 * 4.06e-5), 64)
    assert np.array_equal(s["uu"], want_u) and np.array_equal(s["vv"], want_v)
    assert np.all(s["weight"] == 1.0)
    # all in the centre tile: top-left (-8, 0), cell (201, 202 | 203) ->
    # tile (209 / 32, 202 / 16) = (6, 12)
    assert np.all(tile == 12 * 32768 + 6)


@pytest.mark.gpu
@pytest.mark.parametrize("single", [False, True])
def test_device_resident_arrays_larger_case(single):
    """Arrays owned by torch on the device, 6.1e5 visibilities with a dense
    core, complex sorted_vis (the extension); two calls in a row give the same
    bytes (the reference's order inside a tile changes from run to run)."""
    import torch
    c = tiled_case(seed=21, single=single, num_times=4, num_baselines=1200,
                   num_chan=128, grid_size=2048, core=0.5, spread=0.52)
    m, offsets, counts, skipped, sort, index = _oracle_all(c, complex_vis=True)
    dev = torch.device("cuda", 0)
    t = {k: torch.from_numpy(c[k]).to(dev)
         for k in ("uvw", "freqs", "vis", "weights")}
    nt = num_tiles(c)
    d_offsets = torch.full((nt + 1,), 7, dtype=torch.int32, device=dev)
    d_counts = torch.full((nt,), 7, dtype=torch.int32, device=dev)
    d_skipped = torch.full((1,), 7, dtype=torch.int32, device=dev)
    n = ctypes.c_int(0)
    count_and_prefix_sum(t["uvw"], t["freqs"], t["vis"], *geometry_args(c), n,
                         d_offsets, d_counts, d_skipped)
    assert n.value == m and skipped > 0
    assert np.array_equal(d_offsets.cpu().numpy(), offsets)
    assert np.array_equal(d_counts.cpu().numpy(), counts)
    assert int(d_skipped.cpu()[0]) == skipped
    real = t["uvw"].dtype
    runs = []
    for _ in range(2):
        s = dict(uu=torch.full((m,), 7, dtype=real, device=dev),
                 vv=torch.full((m,), 7, dtype=real, device=dev),
                 weight=torch.full((m,), 7, dtype=real, device=dev),
                 tile=torch.full((m,), 7, dtype=torch.int32, device=dev),
                 vis=torch.full((m,), 7, dtype=t["vis"].dtype, device=dev))
        offs = d_offsets.clone()
        bucket_sort(t["uvw"], t["freqs"], t["vis"], t["weights"],
                    *geometry_args(c), s["uu"], s["vv"], s["weight"], s["tile"],
                    s["vis"], offs)
        torch.cuda.synchronize()
        runs.append({k: v.cpu().numpy() for k, v in s.items()})
        assert np.array_equal(offs.cpu().numpy(), sort["tile_offsets"])
    for k, key in (("uu", "sorted_uu"), ("vv", "sorted_vv"), ("vis", "sorted_vis"),
                   ("weight", "sorted_weight"), ("tile", "sorted_tile")):
        assert np.array_equal(runs[0][k], sort[key]), k
        assert np.array_equal(runs[0][k], runs[1][k]), k
    x = dict(tile=torch.zeros(m, dtype=torch.int32, device=dev),
             uu=torch.zeros(m, dtype=real, device=dev),
             vv=torch.zeros(m, dtype=real, device=dev),
             index=torch.zeros(m, dtype=torch.int32, device=dev))
    offs = d_offsets.clone()
    tiled_indexing(t["uvw"], t["freqs"], *geometry_args(c), c["freqs"].shape[0],
                   c["uvw"].shape[1], c["uvw"].shape[0], x["tile"], x["uu"],
                   x["vv"], x["index"], offs)
    torch.cuda.synchronize()
    assert np.array_equal(x["index"].cpu().numpy(), index["sorted_vis_index"])
    assert np.array_equal(x["tile"].cpu().numpy(), index["sorted_tile"])
    assert np.array_equal(x["uu"].cpu().numpy(), index["sorted_uu"])


@pytest.mark.gpu
def test_short_sorted_arrays_and_location_mismatch():
    """Slots beyond the sorted arrays are skipped (the reference would write
    past their end); arrays in different places are reported."""
    import torch
    c = tiled_case(**CASES["small_f64"])
    m, offsets, _, _, sort, _ = _oracle_all(c)
    s = _sort_arrays(c, m // 2)
    _sort(c, offsets.copy(), s)
    assert np.array_equal(s["uu"], sort["sorted_uu"][:m // 2])
    assert np.array_equal(s["tile"], sort["sorted_tile"][:m // 2])
    with pytest.raises(CError, match="Error 6"):
        _count(c, uvw=torch.from_numpy(c["uvw"]).to("cuda:0"))


def _per_tile_rows(offsets, cols):
    """Rows (one per entry, the given columns) sorted inside every tile."""
    rows = np.stack([np.asarray(x, dtype=np.float64) for x in cols], axis=1)
    out = rows.copy()
    for t in range(offsets.size - 1):
        lo, hi = offsets[t], offsets[t + 1]
        if hi - lo > 1:
            seg = rows[lo:hi]
            out[lo:hi] = seg[np.lexsort(seg.T[::-1])]
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("single", [False, True])
def test_against_the_reference_cuda_build(single, tmp_path):
    """The unmodified reference library (baseline/_ref, sm_100 cubins) through
    its own Python functions on the same inputs, in its own process: counts,
    offsets, skipped count and advanced offsets equal; the sorted arrays equal
    tile by tile as multisets (the reference's order inside a tile is that of
    its atomics). Prints both sides' device times."""
    ref = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.exists(os.path.join(ref, "libska_sdp_func.so")):
        pytest.skip("baseline/_ref not installed (baseline/build_ref_gpu.sh)")
    # u reaches the last half tile column (pu = num_tiles_u, whose pairs the
    # reference files under the next row's first tile: reproduced); v stays
    # off the last half tile ROW, where the reference writes out of bounds
    kw = dict(seed=31, single=single, num_times=4, num_baselines=600,
              num_chan=64, grid_size=1024, core=0.5, spread_v=0.45)
    path = str(tmp_path / "ref_tiled.npz")
    r = subprocess.run(
        [sys.executable, os.path.join(ROOT, "baseline", "ref_tiled.py"),
         "--case", json.dumps(kw), "--out", path, "--steps", "3"],
        capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    res = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
    assert res["available"], res
    z = np.load(path)
    c = tiled_case(**kw)
    n, a = _count(c)
    assert n == int(z["num"]) and n > 0
    assert np.array_equal(a["offsets"], z["offsets"])
    assert np.array_equal(a["counts"], z["counts"])
    assert np.array_equal(a["skipped"], z["skipped"])
    assert np.any((to.entries(c["uvw"], c["freqs"], *geometry_args(c))[2]) == 32)
    s = _sort_arrays(c, n)
    offs = a["offsets"].copy()
    _sort(c, offs, s)
    assert np.array_equal(offs, z["sort_offsets"])
    ours = _per_tile_rows(a["offsets"], [s["tile"], s["uu"], s["vv"], s["svis"],
                                         s["weight"]])
    theirs = _per_tile_rows(a["offsets"], [z["sort_tile"], z["sort_uu"],
                                           z["sort_vv"], z["sort_vis"],
                                           z["sort_weight"]])
    assert np.array_equal(ours, theirs)
    offs = a["offsets"].copy()
    t = _index(c, offs, n)
    assert np.array_equal(offs, z["index_offsets"])
    ours = _per_tile_rows(a["offsets"], [t["tile"], t["index"], t["uu"], t["vv"]])
    theirs = _per_tile_rows(a["offsets"], [z["index_tile"], z["index_index"],
                                           z["index_uu"], z["index_vv"]])
    assert np.array_equal(ours, theirs)
    print("reference CUDA build, device ms (count, sort, index):",
          res["ms_count"], res["ms_sort"], res["ms_index"])
