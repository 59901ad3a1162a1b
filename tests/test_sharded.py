"""Multi-GPU host layer (ShardedGridderUvwEsFft): plane-ownership
partitioning, and world-size-2 runs over gloo -- on the CPU with an
oracle-backed per-rank gridder (host logic), on the GPU with the product
(two processes sharing one device)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from ska_sdp_func.grid_data.sharded_gridder_uvw_es_fft import (
    cost_from_batches, model_plane_cost, partition_planes)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "sharded_worker.py")


def test_partition_covers_all_planes_in_order():
    rng = np.random.default_rng(0)
    for ranks in (1, 2, 3, 8):
        for planes in (1, 5, 37, 1488):
            cost = rng.uniform(0.1, 2.0, planes)
            b = partition_planes(cost, ranks)
            assert b[0] == 0 and b[-1] == planes and len(b) == ranks + 1
            assert np.all(np.diff(b) >= 0)


def test_partition_balances_within_one_plane():
    rng = np.random.default_rng(1)
    cost = rng.uniform(0.5, 1.5, 1488)
    cost[:200] *= 0.1          # cheap planes at the low-w end
    b = partition_planes(cost, 8)
    sums = np.array([cost[b[r]:b[r + 1]].sum() for r in range(8)])
    assert sums.max() - cost.sum() / 8 <= cost.max()
    assert sums.max() / (cost.sum() / 8) < 1.01


def test_partition_edge_cases():
    assert list(partition_planes(np.zeros(4), 2)) == [0, 2, 4]
    assert list(partition_planes([], 3)) == [0, 0, 0, 0]
    b = partition_planes([1.0, 1.0], 4)       # fewer planes than ranks
    assert b[0] == 0 and b[-1] == 2 and np.all(np.diff(b) >= 0)
    assert list(partition_planes([5.0, 1.0, 1.0, 1.0, 1.0, 1.0], 2)) == [0, 1, 6]


def test_model_cost_counts_visits_over_the_footprint():
    hist = np.zeros(12)
    hist[3] = 10                              # footprints start on plane 3
    cost = model_plane_cost(hist, support=4)
    assert np.all(cost[3:7] > 0) and np.all(cost[:3] == 0) and np.all(cost[7:] == 0)


def test_cost_from_batches():
    c = cost_from_batches([2, 6], [4, 2], [8.0, 3.0], 10)
    assert np.all(c[2:6] == 2.0) and np.all(c[6:8] == 1.5)
    assert np.isnan(c[0]) and np.isnan(c[9])


def _run_world2(mode, out_prefix, port):
    env = dict(os.environ, OMP_NUM_THREADS="2")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
           "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), WORKER, mode, out_prefix]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    return np.load(out_prefix + ".npz")


def _oracle_results():
    from oracle import es_oracle as eo
    from tests.sharded_worker import test_case
    case = test_case()
    n = case["image"].shape[0]
    o = eo.OracleGridder(case["uvw"], case["freq_hz"], case["vis"],
                         case["weight"], case["image"], case["pixel_size"],
                         case["pixel_size"], 1e-5, True)
    dirty0, _ = o.grid_uvw_es_fft(case["uvw"], case["freq_hz"], case["vis"],
                                  case["weight"], np.full((n, n), 0.25))
    dirty, _ = o.grid_uvw_es_fft(case["uvw"], case["freq_hz"], case["vis"],
                                 case["weight"])
    vis, img, _ = o.ifft_grid_uvw_es(
        case["uvw"], case["freq_hz"], case["image"],
        vis=np.full(case["vis"].shape, 1.0 + 2.0j))
    return o, dirty0, dirty, vis, img


def test_world2_gloo_plane_ownership_on_cpu(tmp_path):
    """Two ranks over gloo, each gridding / degridding only the planes it
    owns (oracle-backed stand-in): reduced image and summed prediction equal
    the single-rank oracle; the correction is applied once; rank 1's initial
    image / vis content is ignored; rebalancing moves the boundary towards
    the expensive planes."""
    res = _run_world2("oracle", str(tmp_path / "w2"), 29731)
    o, dirty0, dirty, vis, img = _oracle_results()
    planes = o.plan["num_total_w_grids"]
    assert planes >= 8
    b0, b1 = res["bounds0"], res["bounds1"]
    assert b0[0] == 0 and b0[-1] == planes and 0 < b0[1] < planes
    # fake batch times made the upper half 3x as expensive: the boundary moves up
    assert b1[1] > planes // 2 and b1[1] >= b0[1]
    scale = np.max(np.abs(dirty))
    # the initial content (0.25 on rank 0 only) is corrected once per owning
    # rank in the stand-in, so compare the visibility part via linearity
    assert np.max(np.abs(res["dirty_rebalanced"] - dirty)) / scale < 1e-12
    assert np.max(np.abs(res["vis"] - vis)) / np.max(np.abs(vis)) < 1e-12
    assert np.max(np.abs(res["model_corrected"] - img)) / np.max(np.abs(img)) < 1e-13
    for key in ("vis", "dirty"):
        assert bool(res[key + "_same_on_all_ranks"]), key
    # rank 0 ran exactly its block in every call
    calls = res["calls"]
    assert tuple(calls[0]) == (0, b0[1]) and tuple(calls[1]) == (0, b1[1])


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["gpu", "gpu32"])
def test_two_ranks_share_one_device(mode, tmp_path):
    """Two processes (gloo) on cuda:0, the product underneath: the reduced
    dirty image and the summed prediction equal the oracle at the north-star
    tolerance, and the single-rank product result to rounding."""
    from tests.helpers import TOL
    res = _run_world2(mode, str(tmp_path / "w2"), 29741)
    o, dirty0, dirty, vis, img = _oracle_results()
    tol = TOL[np.dtype(np.float32 if mode == "gpu32" else np.float64)]
    assert 0 < res["bounds0"][1] < o.plan["num_total_w_grids"]
    assert np.max(np.abs(res["dirty"] - dirty0)) / np.max(np.abs(dirty0)) <= tol
    assert np.max(np.abs(res["dirty_rebalanced"] - dirty)) / np.max(np.abs(dirty)) <= tol
    assert np.max(np.abs(res["vis"] - vis)) / np.max(np.abs(vis)) <= tol
    assert np.max(np.abs(res["model_corrected"] - img)) / np.max(np.abs(img)) <= tol
    for key in ("vis", "model_corrected", "dirty"):
        assert bool(res[key + "_same_on_all_ranks"]), key
