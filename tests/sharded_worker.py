"""One rank of the multi-process tests of ShardedGridderUvwEsFft
(tests/test_sharded.py launches it with torch.distributed.run).

  python sharded_worker.py MODE OUT_PREFIX
    MODE = oracle : gloo on the CPU; the per-rank gridder is a stand-in built
                    on the oracle (tests may use the oracle), so the host
                    logic -- global lattice, plane ownership, rebalancing,
                    reduce, correction once -- runs without a GPU
    MODE = gpu    : gloo between processes that share cuda:0; the per-rank
                    gridder is the product (C-ABI, device arrays)
Rank 0 writes OUT_PREFIX.npz.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "ska-sdp-func_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from tests import synth  # noqa: E402

C_0 = 299792458.0


def test_case():
    """Small w-stacking problem with ~20 planes (same on every rank)."""
    return synth.reference_test_case(128, num_vis=300, num_chan=6, seed=7)


class OracleLocal:
    """Per-rank stand-in with the interface ShardedGridderUvwEsFft needs,
    computing with oracle/es_oracle (fp64)."""

    def __init__(self, uvw, freq_hz, vis, weight, dirty_image, pix_x, pix_y,
                 epsilon, do_w_stacking, w_range=None):
        from oracle import es_oracle as eo
        self.o = eo.OracleGridder(np.asarray(uvw), np.asarray(freq_hz),
                                  np.asarray(vis), np.asarray(weight),
                                  np.asarray(dirty_image), pix_x, pix_y,
                                  epsilon, do_w_stacking)
        if w_range is not None:
            assert abs(w_range[0] - self.o.plan["min_abs_w"]) < 1e-9
            assert abs(w_range[1] - self.o.plan["max_abs_w"]) < 1e-9
        self.range = (0, self.o.plan["num_total_w_grids"])
        self.calls = []

    def info(self):
        return dict(num_total_w_grids=self.o.plan["num_total_w_grids"],
                    support=self.o.plan["support"])

    def plane_histogram(self, uvw, freq_hz):
        pl = self.o.plan
        w = np.abs(np.asarray(uvw)[:, 2:3]) * np.asarray(freq_hz)[None, :] / C_0
        pos_w = (w - pl["min_plane_w"]) * pl["w_scale"]
        p0 = np.ceil(pos_w - 0.5 * pl["support"]).astype(np.int64)
        p0 = np.clip(p0, 0, pl["num_total_w_grids"] - 1)
        return np.bincount(p0.ravel(), minlength=pl["num_total_w_grids"])

    def set_plane_range(self, lo, hi, apply_correction=True):
        self.range = (lo, hi)

    def correct_image(self, image):
        """The oracle corrects inside its calls; the correction is a
        pointwise factor, so corrected partial images still add up."""

    def grid_uvw_es_fft(self, uvw, freq_hz, vis, weight, dirty_image):
        self.calls.append(("grid",) + self.range)
        if self.range[0] >= self.range[1]:
            return
        out, _ = self.o.grid_uvw_es_fft(
            np.asarray(uvw), np.asarray(freq_hz), np.asarray(vis),
            np.asarray(weight), planes=self.range)
        dirty_image += torch.from_numpy(out) if torch.is_tensor(dirty_image) else out

    def ifft_grid_uvw_es(self, uvw, freq_hz, vis, weight, dirty_image):
        self.calls.append(("degrid",) + self.range)
        lo, hi = self.range
        out, img, _ = self.o.ifft_grid_uvw_es(
            np.asarray(uvw), np.asarray(freq_hz), np.asarray(dirty_image),
            planes=(lo, max(hi, lo)) if hi > lo else (0, 0))
        vis += torch.from_numpy(out) if torch.is_tensor(vis) else out
        dirty_image[...] = torch.from_numpy(img) if torch.is_tensor(dirty_image) else img

    def batch_times(self):
        """Pretend the planes of the upper half of the lattice cost three
        times as much as the others (drives the rebalancing test)."""
        lo, hi = self.range
        first = np.arange(lo, hi, dtype=np.int32)
        count = np.ones(hi - lo, dtype=np.int32)
        half = self.o.plan["num_total_w_grids"] // 2
        ms = np.where(first >= half, 3.0, 1.0)
        return first, count, ms

    def enable_timing(self, enable=True):
        pass


def main():
    mode, out_prefix = sys.argv[1], sys.argv[2]
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    from ska_sdp_func.grid_data import ShardedGridderUvwEsFft
    case = test_case()
    single = mode == "gpu32"
    if single:
        case = synth.reference_test_case(128, num_vis=300, num_chan=6, seed=7,
                                         single=True)
    n = case["image"].shape[0]
    result = {}
    if mode == "oracle":
        arrays = [case[k].copy() for k in ("uvw", "freq_hz", "vis", "weight")]
        image = np.zeros((n, n))
        factory = OracleLocal
    else:
        torch.cuda.set_device(0)
        dev = torch.device("cuda", 0)
        arrays = [torch.from_numpy(case[k]).to(dev)
                  for k in ("uvw", "freq_hz", "vis", "weight")]
        image = torch.zeros((n, n), dtype=arrays[0].dtype, device=dev)
        factory = None
    g = ShardedGridderUvwEsFft(*arrays, image, case["pixel_size"],
                               case["pixel_size"], 1e-5, True,
                               local_factory=factory)
    g.enable_timing(True)
    result["bounds0"] = np.asarray(g.bounds["grid"])
    # gridding: rank 0's initial image content is accumulated into, the other
    # ranks' content must be ignored
    if mode == "oracle":
        image[...] = 0.25 if rank == 0 else 99.0
    else:
        image.fill_(0.25 if rank == 0 else 99.0)
    g.grid_uvw_es_fft(*arrays, image)
    result["dirty"] = np.asarray(image.cpu() if torch.is_tensor(image) else image).copy()
    result["bounds1"] = np.asarray(g.rebalance("grid"))
    if mode == "oracle":
        image[...] = 0.0
    else:
        image.zero_()
    g.grid_uvw_es_fft(*arrays, image, root=0)   # reduce to rank 0 only
    result["dirty_rebalanced"] = np.asarray(
        image.cpu() if torch.is_tensor(image) else image).copy()
    # degridding: model image only on rank 0, vis content of rank 0 is added to
    if mode == "oracle":
        model = case["image"].copy() if rank == 0 else np.zeros((n, n))
        vis = np.full(case["vis"].shape, 1.0 + 2.0j if rank == 0 else 77.0,
                      dtype=case["vis"].dtype)
    else:
        model = torch.from_numpy(case["image"]).to(dev) if rank == 0 else \
            torch.zeros((n, n), dtype=arrays[0].dtype, device=dev)
        vis = torch.full(case["vis"].shape, 1.0 + 2.0j if rank == 0 else 77.0,
                         dtype=arrays[2].dtype, device=dev)
    arrays[2] = vis
    g.ifft_grid_uvw_es(*arrays, model)
    result["vis"] = np.asarray(vis.cpu() if torch.is_tensor(vis) else vis).copy()
    result["model_corrected"] = np.asarray(
        model.cpu() if torch.is_tensor(model) else model).copy()
    # every rank must hold the same prediction and the same corrected image
    # (host arrays: only root's model image is written back)
    same = ("vis", "dirty") if mode == "oracle" else ("vis", "model_corrected", "dirty")
    for key in same:
        t = torch.from_numpy(np.ascontiguousarray(result[key]))
        t = torch.view_as_real(t) if t.is_complex() else t
        lo, hi = t.clone(), t.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        result[key + "_same_on_all_ranks"] = np.array(bool(torch.equal(lo, hi)))
    if mode == "oracle":
        result["calls"] = np.array([c[1:] for c in g.local.calls])
    if rank == 0:
        np.savez(out_prefix + ".npz", world=world, **result)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
