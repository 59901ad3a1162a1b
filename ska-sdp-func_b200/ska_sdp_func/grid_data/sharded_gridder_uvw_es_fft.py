"""Multi-GPU front-end of the w-stacking ES gridder / degridder: one process
per GPU (``torch.distributed``, NCCL over NVLink), w-plane ownership.

The reference has no multi-GPU path (no NCCL / MPI anywhere in it); the loop
that is sharded here is its per-plane loop,
``sdp_gridder_uvw_es_fft.cpp:578-698`` (gridding) and ``:832-952``
(degridding). More than 99 % of the work of a call is per w-plane (scatter /
gather of the plane, its 2-D FFT, its w-screen), so the planes are what is
divided:

* every rank plans the SAME global w lattice (same ``grid_size``, ``support``,
  ``w_scale``, ``min_plane_w``, number of planes) and sees the same
  visibilities (rows x channels);
* rank r owns a contiguous block of planes ``[bounds[r], bounds[r + 1])``
  (``sdp_gridder_uvw_es_fft_set_plane_range``): a visibility is gridded only
  onto the owned planes of its footprint, so the per-rank partial dirty images
  add up to the full one and the per-rank partial predictions add up to the
  full prediction. The blocks are balanced first with a cost model from the
  visibility histogram over planes, then with the device times the library
  measured per plane batch (``rebalance``);
* gridding:   local partial image -> ONE sum-reduce of N^2 reals -> grid
  correction once;
* degridding: model image broadcast -> grid correction (identical on every
  rank) -> local partial predictions -> ONE sum-reduce of the visibilities.

There is no collective on the data path and no fused compute + collective
kernel: the one exchange (1 GiB at 16384^2) follows hundreds of milliseconds
of per-rank work. The collective lives here, in the host layer, and not in
``libska_sdp_func.so``: the C library keeps the reference's dependency set
(CUDA runtime + FFT only, no NCCL / MPI), and the image / visibility arrays
are the caller's anyway.

Arrays may be torch CUDA tensors (processed in place) or numpy arrays (staged
through the device of this rank inside the call; that is what ``bench.py``
times as the end-to-end number).
"""

import numpy as np

from .gridder_uvw_es_fft import GridderUvwEsFft


def partition_planes(cost, num_ranks):
    """Contiguous blocks of planes with (nearly) equal summed cost.

    Returns ``bounds`` (int64 [num_ranks + 1], ``bounds[0] = 0``,
    ``bounds[-1] = len(cost)``); rank r owns ``[bounds[r], bounds[r + 1])``.
    Boundary r is put where the running sum crosses ``r / num_ranks`` of the
    total, rounded to the nearer plane, so no block is further than one
    plane's cost from the ideal share. Blocks may be empty when there are
    fewer planes than ranks."""
    cost = np.maximum(np.asarray(cost, dtype=np.float64), 0.0)
    num_planes = cost.shape[0]
    bounds = np.zeros(num_ranks + 1, dtype=np.int64)
    bounds[-1] = num_planes
    total = float(cost.sum())
    if num_planes == 0 or num_ranks == 1:
        return bounds
    if total <= 0.0:
        cost = np.ones(num_planes)
        total = float(num_planes)
    prefix = np.concatenate([[0.0], np.cumsum(cost)])
    for r in range(1, num_ranks):
        target = total * r / num_ranks
        k = int(np.searchsorted(prefix, target, side="left"))
        k = min(max(k, 1), num_planes)
        # prefix[k - 1] < target <= prefix[k]: take the nearer one
        if target - prefix[k - 1] < prefix[k] - target:
            k -= 1
        bounds[r] = max(k, int(bounds[r - 1]))
    return bounds


def model_plane_cost(first_plane_hist, support, fft_share=0.9):
    """Cost per plane before anything was measured: every plane that
    receives visibilities pays one 2-D FFT + w-screen (``fft_share`` of the
    mean plane cost), the rest is proportional to the number of (visibility,
    plane) visits. ``first_plane_hist[p]`` = visibilities whose footprint
    starts on plane p; a footprint spans ``support`` planes (up to
    ``support + 1`` at exactly-integer positions, ignored here)."""
    hist = np.asarray(first_plane_hist, dtype=np.float64)
    visits = np.convolve(hist, np.ones(int(support)))[:hist.shape[0]]
    occupied = visits > 0
    mean_visits = visits[occupied].mean() if occupied.any() else 1.0
    return np.where(occupied,
                    fft_share + (1.0 - fft_share) * visits / mean_visits, 0.0)


def cost_from_batches(first_plane, num_planes, millis, total_planes):
    """Per-plane cost array (NaN where nothing was measured) from the
    per-batch device times of one rank's last call."""
    cost = np.full(total_planes, np.nan)
    for first, count, ms in zip(first_plane, num_planes, millis):
        if count > 0:
            cost[first:first + count] = ms / count
    return cost


class ShardedGridderUvwEsFft:
    """``GridderUvwEsFft`` over the ranks of a ``torch.distributed`` process
    group, w-planes owned in contiguous blocks. Same constructor and method
    arguments as the single-GPU class (reference:
    ``src/ska_sdp_func/grid_data/gridder_uvw_es_fft.py:16-146``); every rank
    passes the same visibilities.

    ``local_factory`` (tests): builds the per-rank gridder; anything with
    ``info``, ``plane_histogram``, ``set_plane_range``, ``grid_uvw_es_fft``,
    ``ifft_grid_uvw_es``, ``correct_image`` and ``batch_times``.
    """

    def __init__(self, uvw, freq_hz, vis, weight, dirty_image,
                 pixel_size_x_rad, pixel_size_y_rad, epsilon: float,
                 do_w_stacking: bool, group=None, local_factory=None):
        import torch
        import torch.distributed as dist
        self._torch, self._dist = torch, dist
        self.group = group
        self.distributed = dist.is_available() and dist.is_initialized()
        self.rank = dist.get_rank(group) if self.distributed else 0
        self.world = dist.get_world_size(group) if self.distributed else 1
        self.device = None
        if torch.cuda.is_available():
            self.device = torch.device("cuda", torch.cuda.current_device())

        w_range = None
        if do_w_stacking:
            mn, mx = GridderUvwEsFft.get_w_range(uvw, freq_hz)
            rng = torch.tensor([-float(mn), float(mx)], dtype=torch.float64)
            rng = self._reduce_small(rng, "max")
            w_range = (-float(rng[0]), float(rng[1]))
        factory = local_factory or GridderUvwEsFft
        self.local = factory(uvw, freq_hz, vis, weight, dirty_image,
                             pixel_size_x_rad, pixel_size_y_rad, epsilon,
                             do_w_stacking, w_range=w_range)
        info = self.local.info()
        self.num_planes = int(info["num_total_w_grids"])
        self.support = int(info["support"])
        if self.world > 1:
            hist = np.asarray(self.local.plane_histogram(uvw, freq_hz))
            cost = model_plane_cost(hist, self.support)
        else:
            cost = np.ones(self.num_planes)
        # one ownership table per direction: their per-plane costs differ
        self.cost = {"grid": cost.copy(), "degrid": cost.copy()}
        self.bounds = {k: partition_planes(c, self.world)
                       for k, c in self.cost.items()}

    # ------------------------------------------------------------ plumbing

    def _reduce_small(self, t, op):
        """All-reduce of a small host tensor (through the device for NCCL)."""
        if not self.distributed or self.world == 1:
            return t
        dist = self._dist
        rop = {"max": dist.ReduceOp.MAX, "sum": dist.ReduceOp.SUM}[op]
        if dist.get_backend(self.group) == "nccl":
            d = t.to(self.device)
            dist.all_reduce(d, op=rop, group=self.group)
            return d.cpu()
        dist.all_reduce(t, op=rop, group=self.group)
        return t

    def _collective(self, fn, t, **kw):
        """fn(tensor) for a large array: NCCL takes the device tensor as is;
        gloo (tests, one device shared by several ranks) goes through the
        host."""
        dist = self._dist
        if dist.get_backend(self.group) != "nccl" and t.is_cuda:
            h = t.cpu()
            fn(h, group=self.group, **kw)
            t.copy_(h)
        else:
            fn(t, group=self.group, **kw)

    def _own(self, direction):
        b = self.bounds[direction]
        return int(b[self.rank]), int(b[self.rank + 1])

    def _stage(self, arrays):
        """numpy -> device tensors of this rank (None stays None)."""
        torch = self._torch
        out = []
        for a in arrays:
            if isinstance(a, np.ndarray):
                t = torch.from_numpy(a)
                out.append(t.to(self.device, non_blocking=True)
                           if self.device is not None else t)
            else:
                out.append(a)
        return out

    # ---------------------------------------------------------- operations

    def grid_uvw_es_fft(self, uvw, freq_hz, vis, weight, dirty_image,
                        root=None):
        """Dirty image of all visibilities. Rank 0's (``root``'s) dirty_image
        content is accumulated into, the other ranks' content is ignored. On
        return the corrected image is in every rank's dirty_image
        (``root=None``: all-reduce) or only in ``root``'s (reduce)."""
        if self.world == 1:
            self.local.grid_uvw_es_fft(uvw, freq_hz, vis, weight, dirty_image)
            return
        torch, dist = self._torch, self._dist
        host_image = isinstance(dirty_image, np.ndarray)
        d_uvw, d_freq, d_vis, d_wgt, d_img = self._stage(
            [uvw, freq_hz, vis, weight, dirty_image])
        if self.rank != (root or 0):
            d_img.zero_()
        lo, hi = self._own("grid")
        self.local.set_plane_range(lo, hi, apply_correction=False)
        self.local.grid_uvw_es_fft(d_uvw, d_freq, d_vis, d_wgt, d_img)
        if root is None:
            self._collective(dist.all_reduce, d_img, op=dist.ReduceOp.SUM)
        else:
            self._collective(dist.reduce, d_img, dst=root,
                             op=dist.ReduceOp.SUM)
        if root is None or self.rank == root:
            self.local.correct_image(d_img)
            if host_image:
                torch.from_numpy(dirty_image).copy_(d_img)
        self._last = "grid"

    def ifft_grid_uvw_es(self, uvw, freq_hz, vis, weight, dirty_image,
                         root=0, reduce_to_root=False):
        """vis += prediction from ``root``'s dirty_image (broadcast to every
        rank, grid-corrected in place everywhere, like the single-GPU call
        corrects the caller's image). ``root``'s vis content is added to, the
        other ranks' content is ignored. On return the full prediction is in
        every rank's vis (default) or only in ``root``'s."""
        if self.world == 1:
            self.local.ifft_grid_uvw_es(uvw, freq_hz, vis, weight, dirty_image)
            return
        torch, dist = self._torch, self._dist
        host_vis = isinstance(vis, np.ndarray)
        host_image = isinstance(dirty_image, np.ndarray)
        if host_image and self.rank != root:
            # only root's image is an input: the others receive it on the device
            d_uvw, d_freq, d_vis, d_wgt = self._stage([uvw, freq_hz, vis, weight])
            d_img = torch.empty(dirty_image.shape, device=self.device,
                                dtype=torch.from_numpy(dirty_image).dtype)
        else:
            d_uvw, d_freq, d_vis, d_wgt, d_img = self._stage(
                [uvw, freq_hz, vis, weight, dirty_image])
        self._collective(dist.broadcast, d_img, src=root)
        if self.rank != root:
            d_vis.zero_()
        lo, hi = self._own("degrid")
        self.local.set_plane_range(lo, hi, apply_correction=False)
        self.local.correct_image(d_img)
        self.local.ifft_grid_uvw_es(d_uvw, d_freq, d_vis, d_wgt, d_img)
        as_real = torch.view_as_real(d_vis)
        if reduce_to_root:
            self._collective(dist.reduce, as_real, dst=root,
                             op=dist.ReduceOp.SUM)
        else:
            self._collective(dist.all_reduce, as_real, op=dist.ReduceOp.SUM)
        if host_vis and (not reduce_to_root or self.rank == root):
            torch.from_numpy(vis).copy_(d_vis)
        if host_image and self.rank == root:
            # host arrays: only root's image is the caller's model image
            torch.from_numpy(dirty_image).copy_(d_img)
        self._last = "degrid"

    # ----------------------------------------------------------- balancing

    def rebalance(self, direction=None):
        """Re-partitions plane ownership from the device times the library
        measured per plane batch in the last call of ``direction`` (timing
        must be enabled on the local plan). Collective: every rank calls it.
        Returns the new bounds."""
        direction = direction or getattr(self, "_last", "grid")
        if self.world == 1:
            return self.bounds[direction]
        torch = self._torch
        first, count, ms = self.local.batch_times()
        mine = cost_from_batches(first, count, ms, self.num_planes)
        known = torch.from_numpy(np.nan_to_num(mine, nan=0.0))
        seen = torch.from_numpy((~np.isnan(mine)).astype(np.float64))
        known = self._reduce_small(known, "sum").numpy()
        seen = self._reduce_small(seen, "sum").numpy()
        cost = self.cost[direction].copy()
        measured = seen > 0
        if measured.any():
            # planes nobody ran (empty blocks) keep the model value, scaled
            # to the measured ones
            scale = (known[measured].sum() /
                     max(cost[measured].sum(), 1e-30))
            cost = cost * scale
            cost[measured] = known[measured] / seen[measured]
        self.cost[direction] = cost
        self.bounds[direction] = partition_planes(cost, self.world)
        return self.bounds[direction]

    # ------------------------------------------------------------- queries

    def info(self):
        """Planning scalars (identical on every rank)."""
        return self.local.info()

    def stats(self):
        """Statistics of this rank's last local call."""
        return self.local.stats()

    def enable_timing(self, enable=True):
        self.local.enable_timing(enable)
