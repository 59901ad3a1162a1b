"""Parity at the grid sizes bench.py TIMES: the line-FFT
plan depends on G -- C2's G = 6160 runs radices 28.20.11 with 4-column /
2-row bundles, C4's G = 24640 four passes with one 197 KB line per CTA, C5's
G = 11520 drops the row bundle, C3's G = 14400 fp64 keeps its twiddles in
global memory -- so each of those kernel shapes is compared with the oracle
here, on the B200, through the C-ABI. The oracle costs one G x G fp64 FFT per
w-plane on the host, so the big grids are checked on a block of planes in the
middle of the lattice (all visibilities that touch them, both directions)."""
import numpy as np
import pytest

from tests import synth
from tests.helpers import check_parity_planes

pytestmark = pytest.mark.gpu


def test_c2_full_size_all_visibilities():
    """BASELINE configs[1] exactly as bench.py --workload C2 runs it: 1000 x
    1024 visibilities, 4096^2 image, G = 6160, all 72 planes (gridding);
    degridding on a block of 12 planes."""
    case = synth.uniform_box(1000, 1024, 4096, seed=41, single=True)
    out, info = check_parity_planes(case, 1e-5, None, do_degrid=False)
    assert info["grid_size"] == 6160 and info["support"] == 8
    out2, _ = check_parity_planes(case, 1e-5, 12, do_grid=False)
    print("C2 full size", out, out2)


def test_c4_grid_size_24640_single_line_mode():
    """BASELINE configs[3] shapes: 16384^2 image, G = 24640 (one line per CTA,
    four radix passes), SKA-Low generator with 512 stations x 2 channels at
    the band edges (2.6e5 visibilities, the full w range -> 1488 planes): two
    planes gridded, one degridded."""
    case = synth.ska_low_snapshot(512, 2, 16384, seed=43, single=True)
    out, info = check_parity_planes(case, 1e-5, 2, do_degrid=False,
                                    image_size=16384)
    assert info["grid_size"] == 24640 and info["image_size"] == 16384
    out2, _ = check_parity_planes(case, 1e-5, 1, do_grid=False, image_size=16384)
    print("C4 grid size", out, out2, info["num_total_w_grids"])


@pytest.mark.parametrize("generic", [False, True])
def test_c5_grid_size_11520_eps_1e4(generic, monkeypatch):
    """BASELINE configs[4] shapes: 8192^2, eps = 1e-4 (W = 7, G = 11520: one
    row per CTA in the row pass), wide field, 3 planes, both directions; with
    the kernel set built for this stage list (32, 24, 15) and with the generic
    set (radix dispatch at run time) that every other length uses."""
    if generic:
        monkeypatch.setenv("SKA_SDP_FUNC_ES_FFT_GENERIC", "1")
    case = synth.wide_field(2000, 32, 8192, seed=44, single=True)
    out, info = check_parity_planes(case, 1e-4, 3)
    assert info["grid_size"] == 11520 and info["support"] == 7
    print("C5 grid size", out, info["num_total_w_grids"])


def test_c3_grid_size_14400_fp64():
    """BASELINE configs[2] shapes: 8192^2, fp64, eps = 1e-10 (W = 12,
    G = 14400: 230 KB line, twiddles in global memory), Mid-like uvw, 2 planes,
    both directions."""
    case = synth.mid_like_predict(125, 1, 8, 8192, seed=42, single=False)
    out, info = check_parity_planes(case, 1e-10, 2)
    assert info["grid_size"] == 14400 and info["support"] == 12
    print("C3 grid size", out, info["num_total_w_grids"])
