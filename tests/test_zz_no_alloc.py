"""sdp_weighting_briggs_no_alloc: the Briggs weighting entry point that
allocates nothing (uv weight grid + sums in a file-scope device array of the
library; no counterpart in the reference).
Checked through the C-ABI against the existing entry point
sdp_weighting_briggs on the same device-resident inputs, on two calls in a
row (the second with another grid size and precision: the shared array keeps
the first call's contents and must be cleared), plus the stated limits.

Collected last on purpose (file name).
"""
import numpy as np
import pytest

from tests.synth_weighting import weighting_case
from ska_sdp_func.utility import Lib
from ska_sdp_func.utility.error_checking import CError
from ska_sdp_func.visibility import briggs_weights, briggs_weights_no_alloc

MAX_CELLS = 1024 * 1024  # ES_WT_STATIC_CELLS, include/.../sdp_weighting.h


def test_entry_point_is_exported():
    assert hasattr(Lib.handle(), "sdp_weighting_briggs_no_alloc")


def _device_case(torch, **kw):
    c = weighting_case(**kw)
    dev = torch.device("cuda", 0)
    t = {k: torch.from_numpy(c[k]).to(dev) for k in ("uvw", "freq_hz", "input")}
    return c, t, dev


def _both(torch, **kw):
    """(output of sdp_weighting_briggs, output of the no-alloc variant)."""
    c, t, dev = _device_case(torch, **kw)
    grid = torch.zeros(c["grid"].shape, dtype=t["input"].dtype, device=dev)
    want = torch.from_numpy(c["output0"]).to(dev)
    got = torch.from_numpy(c["output0"]).to(dev)
    briggs_weights(t["uvw"], t["freq_hz"], c["max_abs_uv"], c["robust"], grid,
                   t["input"], want)
    torch.cuda.synchronize()
    # every tensor exists before the call: nothing is allocated inside it
    briggs_weights_no_alloc(t["uvw"], t["freq_hz"], c["max_abs_uv"], c["robust"],
                            c["grid"].shape[0], t["input"], got)
    torch.cuda.synchronize()
    return want.cpu().numpy(), got.cpu().numpy()


@pytest.mark.gpu
def test_two_calls_in_a_row_match_the_existing_entry_point():
    import torch
    # first call: fp64, 128^2 x 2 grid, many hits per cell
    want, got = _both(torch, seed=11, single=False, num_times=4, num_baselines=300,
                      num_chan=64, num_pols=2, grid_size=128)
    # tolerances of tests/test_weighting.py (atomics add in any order)
    assert np.allclose(got, want, rtol=1e-12, atol=0)
    assert np.all(got != 7.0)  # everything inside the grid was written
    # second call: fp32, smaller grid over the same (now dirty) scratch, some
    # visibilities outside the grid keep the caller's value
    want, got = _both(torch, seed=12, single=True, num_times=3, num_baselines=200,
                      num_chan=32, num_pols=1, grid_size=64, shrink=0.7,
                      positive=True)
    assert np.allclose(got, want, rtol=2e-6, atol=0)
    assert np.array_equal(got == 7.0, want == 7.0) and np.any(got == 7.0)
    # third: the first case again gives the first result again
    want3, got3 = _both(torch, seed=11, single=False, num_times=4, num_baselines=300,
                        num_chan=64, num_pols=2, grid_size=128)
    assert np.allclose(got3, want3, rtol=1e-12, atol=0)


@pytest.mark.gpu
def test_limits_are_reported():
    import torch
    c, t, dev = _device_case(torch, seed=13, num_pols=4)
    out = torch.from_numpy(c["output0"]).to(dev)
    # 1024^2 x 4 pols > 1 Mi cells
    with pytest.raises(CError, match="Error 2"):
        briggs_weights_no_alloc(t["uvw"], t["freq_hz"], c["max_abs_uv"], 0.0, 1024,
                                t["input"], out)
    # the largest grid that fits with 4 polarisations works
    briggs_weights_no_alloc(t["uvw"], t["freq_hz"], c["max_abs_uv"], 0.0, 512,
                            t["input"], out)
    torch.cuda.synchronize()
    assert 512 * 512 * 4 == MAX_CELLS
    # host arrays would need staging, i.e. an allocation
    with pytest.raises(CError, match="Error 6"):
        briggs_weights_no_alloc(c["uvw"], c["freq_hz"], c["max_abs_uv"], 0.0, 32,
                                c["input"], c["output0"].copy())
