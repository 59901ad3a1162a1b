#!/usr/bin/env python
"""Benchmark of the B200-native w-stacking ES gridder / degridder
(sdp_grid_uvw_es_fft / sdp_ifft_degrid_uvw_es).

  python bench.py [--gpus N] [--steps K] [--warmup W]
                  [--impl ours|reference|reference-gpu]
                  [--workload C4|C2|C3|C5|C1|C2s|C4s] [--direction both|grid|degrid]

Default workload: C4, the configuration BASELINE.json's north_star names (SKA-Low
snapshot, 130816 baselines x 1024 channels, 16384^2 image, fp32, eps 1e-5), at
every N. One "step" = one gridding call over ALL visibilities of that ONE fixed
snapshot (zeroed image in, corrected dirty image out). N > 1 (torchrun, one
rank per GPU) is STRONG scaling: every rank plans the same global w lattice,
owns a contiguous block of w-planes balanced by measured per-plane cost
(ska_sdp_func.grid_data.ShardedGridderUvwEsFft), grids all visibilities onto
its planes only, ONE NCCL sum-reduce of the image, grid correction once.
Rank 0 prints ONE JSON line:

  value      gridded Mvis/s = N_vis / max-over-ranks device time per step, inputs
             resident in HBM, CUDA events on the stream the kernels run on
  e2e        the same through the public call with HOST (pinned) arrays:
             host->device copies of all inputs on every rank and the
             device->host read of the image inside the timed region
  degrid     the degridding half of the metric (same fields), image broadcast ->
             per-rank partial predictions -> one sum-reduce of the visibilities
  parity     max|out - oracle| / max|oracle| of the product's output on this
             workload for a bounded block of w-planes (all planes when cheap)
  roofline   dominant own kernel: algorithmic bytes / its device time (CUDA
             events recorded around every stage of the last timed step)
  roofline_call  whole call: SURVEY.md 8(d)'s byte count (contract) and the bytes
             the implemented algorithm must move (frac_required)
  cpu_baseline   the reference's kernels compiled for the host (oracle/_ref) on a
             bounded sample of the same workload, on this box's cores
  ref_gpu    the UNMODIFIED reference's own CUDA implementation (baseline/_ref,
             built for sm_100) on the same device-resident inputs, same GPU

--impl reference times the reference's CPU-compiled kernels (oracle/_ref) on a
bounded sample per step; --impl reference-gpu times the unmodified reference's
CUDA path. Neither loads the product library.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "ska-sdp-func_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

from tests import synth  # noqa: E402

METRIC = "gridded Mvis/s (sdp_grid_uvw_es_fft)"
METRIC_DEGRID = "degridded Mvis/s (sdp_ifft_degrid_uvw_es)"
UNIT = "Mvis/s"
HBM_FALLBACK_GBS = 6650.0  # B200_PROFILING.md fallback, used only without MEASURED_PEAKS.json
TOL = {4: 1e-5, 8: 1e-10}   # BASELINE.json north_star: max|err| / max|ref|

WORKLOADS = {
    # BASELINE.json configs[3]: the north_star's target snapshot (default)
    "C4": dict(desc="SKA-Low AA*-like snapshot: 130816 baselines x 1024 channels, "
                    "16384^2 image, fp32, eps=1e-5 (BASELINE.json configs[3])",
               make=lambda seed: synth.ska_low_snapshot(512, 1024, 16384,
                                                        seed=43 + seed, single=True),
               eps=1e-5, wstack=True),
    # BASELINE.json configs[1]
    "C2": dict(desc="dirty image: 1000 rows x 1024 channels, 4096^2 image, fp32, "
                    "eps=1e-5, w-stacking (BASELINE.json configs[1])",
               make=lambda seed: synth.uniform_box(1000, 1024, 4096, seed=41 + seed,
                                                   single=True),
               eps=1e-5, wstack=True),
    # BASELINE.json configs[0]
    "C1": dict(desc="1000 rows x 10 channels, 256^2 image, fp64, eps=1e-5 "
                    "(BASELINE.json configs[0])",
               make=lambda seed: synth.reference_test_case(256, seed=40 + seed),
               eps=1e-5, wstack=True),
    # BASELINE.json configs[2]
    "C3": dict(desc="degrid (predict) from 8192^2 model image, 15500 rows x 1024 "
                    "channels (1.6e7 vis), Mid-like uvw, fp64, eps=1e-10 "
                    "(BASELINE.json configs[2])",
               make=lambda seed: synth.mid_like_predict(seed=42 + seed),
               eps=1e-10, wstack=True),
    # BASELINE.json configs[4]
    "C5": dict(desc="wide-field large-|w|: 31250 rows x 1024 channels, 8192^2 image, "
                    "fp32, eps=1e-4 (BASELINE.json configs[4])",
               make=lambda seed: synth.wide_field(31250, 1024, 8192, seed=44 + seed,
                                                  single=True),
               eps=1e-4, wstack=True),
    # scaled-down shapes for quick checks
    "C2s": dict(desc="C2 geometry at 1/4 linear size (250 x 256, 1024^2)",
                make=lambda seed: synth.uniform_box(250, 256, 1024, seed=41 + seed,
                                                    single=True),
                eps=1e-5, wstack=True),
    "C4s": dict(desc="C4 geometry scaled down (64 stations x 64 channels, 2048^2)",
                make=lambda seed: synth.ska_low_snapshot(64, 64, 2048,
                                                         seed=43 + seed, single=True),
                eps=1e-5, wstack=True),
}


def hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except (OSError, KeyError, ValueError):
        return HBM_FALLBACK_GBS, "fallback (B200_PROFILING.md)"


def contract_bytes(info, n_vis, degrid=False):
    """minimum HBM traffic of w-stacking with a 2-pass FFT over
    full planes, per call -> (total, total without the FFT term)."""
    c = 16 if info["is_double"] else 8
    r = c // 2
    g2 = float(info["grid_size"]) ** 2
    n2 = float(info["image_size"]) ** 2
    p = info["num_total_w_grids"]
    rows, chans = info["num_rows"], info["num_chan"]
    if not degrid:
        vis_stream = n_vis * (c + r) + 3 * r * rows + r * chans
        per_plane = c * g2 + 4 * c * g2 + c * n2 + 2 * r * n2
        total = vis_stream + p * per_plane + 2 * r * n2
    else:
        vis_stream = c * n_vis + 3 * r * rows + r * chans
        per_plane = r * n2 + c * g2 + 4 * c * g2 + c * g2
        total = vis_stream + 2 * r * n2 + p * per_plane
    return total, total - p * 4 * c * g2


def required_bytes(info, stats, n_vis, batches, degrid=False):
    """Bytes the IMPLEMENTED algorithm has to move through HBM per call, by
    stage: presorted records, only the occupied 64x64 tiles of every plane,
    pruned line FFTs through the G x N strip (rows of occupied tile rows
    only), image traffic once per plane batch. Occupancy = the (tile, plane) /
    (tile row, plane) totals the library counted during the call."""
    c = 16 if info["is_double"] else 8
    r = c // 2
    n = float(info["image_size"])
    tile_cells = 64.0 * 64.0
    occ_t = float(stats.get("occupied_tiles") or 0)
    occ_r = float(stats.get("occupied_tile_rows") or 0)
    if not stats.get("pruned_fft") or occ_t == 0:
        g2 = float(info["grid_size"]) ** 2
        ntile = -(-info["grid_size"] // 64)
        occ_t = stats["planes_processed"] * ntile * ntile
        occ_r = stats["planes_processed"] * ntile
        tile_cells = g2 / (ntile * ntile)
    rec = 64 if info["is_double"] else 32
    tiles_b = occ_t * tile_cells * c             # occupied tiles of all planes
    strip_b = occ_r * 64.0 * n * c               # strip rows of occupied tile rows
    image_b = batches * r * n * n
    records_b = float(stats["num_entries"]) * rec
    presort = n_vis * (c + r if not degrid else 0) + records_b + n_vis * 13.0
    if not degrid:
        stages = {"scatter": records_b + tiles_b,
                  "fft_rows": tiles_b + strip_b,
                  "fft_cols_screen": strip_b + 2 * image_b}
        tail = 3 * r * n * n + 2 * r * n * n     # transpose-add, correction
    else:
        stages = {"fft_cols_screen": image_b + strip_b,
                  "fft_rows": strip_b + tiles_b,
                  "gather": tiles_b + records_b + float(stats["num_entries"]) * c}
        tail = 2 * r * n * n + 2 * r * n * n + 2 * c * n_vis
    return stages, sum(stages.values()) + presort + tail


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region."""

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,"
             "clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index):
        self.device_index = device_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.device_index}",
                 f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [],
                    "samples": 0}
        time.sleep(0.15)
        self.proc.terminate()
        sm, sm_max, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                 "sw_power_cap"]
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                sm_max.append(float(f[2]))
            except ValueError:
                continue
            for name, val in zip(names, f[4:8]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(sm_max) if sm_max else None,
                "reasons": sorted(reasons), "samples": len(sm)}


_JSON_FD = None


def emit_json(line):
    text = json.dumps(line) + "\n"
    sys.stdout.flush()
    if _JSON_FD is None:
        sys.stdout.write(text)
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, text.encode())


def host_cores(omp_default):
    """Threads for the CPU arms: every core this process may run on. torchrun
    exports OMP_NUM_THREADS=1 to its workers, which would otherwise make the
    reference arm single-threaded."""
    try:
        avail = len(os.sched_getaffinity(0))
    except AttributeError:
        avail = os.cpu_count() or 1
    return max(int(omp_default), avail)


# --------------------------------------------------------------- CPU arms

def decimate_channels(case, factor):
    """Every factor-th channel of a case (same rows, same band edges kept)."""
    chans = case["vis"].shape[1]
    idx = np.unique(np.concatenate([np.arange(0, chans, factor), [chans - 1]]))
    out = dict(case)
    out["freq_hz"] = np.ascontiguousarray(case["freq_hz"][idx])
    out["vis"] = np.ascontiguousarray(case["vis"][:, idx])
    out["weight"] = np.ascontiguousarray(case["weight"][:, idx])
    return out, len(idx)


class CpuReference:
    """The reference's own kernel source compiled for the host (oracle/_ref,
    kind "reference"; falls back to the C restatement oracle/es_oracle.c, kind
    "port", where /root/reference was not available at build time), timed on
    a bounded sample: a block of w-planes from the middle of the lattice and,
    when one plane of the full visibility set alone exceeds the budget, every
    16th channel. Uses its own plan; does not load the product library."""

    def __init__(self, case, eps, wstack):
        from oracle import es_oracle as eo, ref_oracle as ro
        self.case_full = case
        self.single = case["uvw"].dtype == np.float32
        self.rows, self.chans = case["vis"].shape
        self.n = case["image_size"]
        self.eps, self.wstack = eps, wstack
        self.mn, self.mx = ro.get_w_range(case["uvw"], case["freq_hz"])
        self.ro, self.eo = ro, eo
        self.kind = "reference" if ro.available() else "port"
        self.decim = 1
        self._bind(case)

    def _bind(self, case):
        self.case = case
        rows, chans = case["vis"].shape
        prec = 4 if self.single else 8
        if self.kind == "reference":
            self.g = self.ro.RefHostGridder(rows, chans, self.n, case["pixel_size"],
                                            self.eps, self.wstack, self.mn, self.mx,
                                            prec, prec)
            self.info = dict(self.g.info)
            self.cores = host_cores(self.ro.lib().oref_max_threads())
        else:
            img = np.broadcast_to(np.zeros(1, dtype=case["uvw"].dtype), (self.n, self.n))
            self.g = self.eo.OracleGridder(case["uvw"], case["freq_hz"], case["vis"],
                                           case["weight"], img, case["pixel_size"],
                                           case["pixel_size"], self.eps, self.wstack)
            self.info = dict(self.g.plan)
            self.cores = host_cores(self.eo.lib().eso_max_threads())
        self.planes = int(self.info["num_total_w_grids"])

    def run(self, planes, fast=False, degrid=False):
        """(seconds wall, timings[4] = scatter/gather, fft, screen, correction)."""
        c = self.case
        t0 = time.perf_counter()
        if self.kind == "reference":
            if degrid:
                img = c.get("image")
                if img is None:
                    img = np.zeros((self.n, self.n), dtype=c["uvw"].dtype)
                _, _, t = self.g.degrid(c["uvw"], c["freq_hz"], img, planes=planes,
                                        fast=fast, nthreads=self.cores)
            else:
                _, t = self.g.grid(c["uvw"], c["freq_hz"], c["vis"], c["weight"],
                                   planes=planes, fast=fast, nthreads=self.cores)
        else:
            _, t = self.g.grid_uvw_es_fft(c["uvw"], c["freq_hz"], c["vis"],
                                          c["weight"], planes=planes,
                                          nthreads=self.cores)
        return time.perf_counter() - t0, np.asarray(t, dtype=np.float64)

    def size_sample(self, target_s):
        """Chooses the sample: returns (lo, hi) planes; may switch to every
        16th channel when a single plane is over budget."""
        mid = self.planes // 2
        t1, _ = self.run((mid, mid + 1))
        if t1 > 2.0 * target_s and self.chans >= 64 and self.decim == 1:
            dec, _ = decimate_channels(self.case_full, 16)
            self.decim = self.chans / dec["vis"].shape[1]
            self._bind(dec)
            t1, _ = self.run((mid, mid + 1))
        count = int(max(1, min(self.planes, target_s / max(t1, 1e-3))))
        if t1 * self.planes <= 2.5 * target_s:
            count = self.planes      # the whole call is affordable: no scaling
        lo = max(0, mid - count // 2)
        return lo, min(self.planes, lo + count)

    def extrapolate(self, timings, count):
        """Full-call seconds from the stage times of a sample of `count`
        planes: the FFT and the w-screen pass scale with the planes, the
        visibility pass with planes x visibilities (literal schedule: every
        visibility is visited for every plane, es_fft.cpp:578-698)."""
        per_plane = (timings[0] * self.decim + timings[1] + timings[2]) / count
        return per_plane * self.planes + timings[3]

    def describe(self, lo, hi):
        what = (f"w-planes [{lo},{hi}) of {self.planes}, "
                + ("all channels" if self.decim == 1 else
                   f"every {self.decim:.0f}th channel ({self.case['vis'].shape[1]} of "
                   f"{self.chans}; visibility pass scaled by {self.decim:.1f})")
                + f"; per-plane time scaled by {self.planes}/{hi - lo} to the full call")
        how = ("oracle/_ref: reference kernels.cu + PocketFFT compiled for the host, "
               "literal schedule (every visibility visited per plane), OpenMP over "
               "visibilities / pixels, PocketFFT nthreads=cores"
               if self.kind == "reference" else
               "oracle/es_oracle.c: fp64 C restatement, plane-sorted schedule, OpenMP")
        return what, how


def cpu_baseline_object(case, eps, wstack, n_vis, target_s):
    """cpu_baseline of the product line: literal reference schedule plus the
    plane-sorted "best-effort CPU" figure on the same sample."""
    ref = CpuReference(case, eps, wstack)
    lo, hi = ref.size_sample(target_s)
    secs, t = ref.run((lo, hi))
    t_full = ref.extrapolate(t, hi - lo)
    what, how = ref.describe(lo, hi)
    out = {"value": n_vis / t_full / 1e6, "unit": UNIT, "cores": ref.cores,
           "kind": ref.kind, "sample": f"{what}; sample took {secs:.1f} s",
           "how": how, "stage_seconds_of_sample": [float(x) for x in t]}
    if ref.kind == "reference":
        secs_f, tf = ref.run((lo, hi), fast=True)
        # plane-sorted: the visibility pass touches only the footprints on
        # the plane (it still scales with the channels of the sample)
        t_fast = ref.extrapolate(tf, hi - lo)
        out["fast_value"] = n_vis / t_fast / 1e6
        out["fast_how"] = ("same sample, visibilities pre-filtered per plane "
                           "(plane-sorted schedule, identical results): the "
                           "best-effort CPU figure of BASELINE.md section 2")
    return out


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = WORKLOADS[args.workload]
    case = wl["make"](0)
    single = case["uvw"].dtype == np.float32
    n_vis = case["vis"].size
    degrid = args.direction == "degrid"
    ref = CpuReference(case, wl["eps"], wl["wstack"])
    # one step ~4 s so that warm-up + steps end within minutes
    lo, hi = ref.size_sample(4.0)
    for _ in range(args.warmup):
        ref.run((lo, hi), degrid=degrid)
    walls, stage = [], np.zeros(4)
    for _ in range(args.steps):
        s, t = ref.run((lo, hi), degrid=degrid)
        walls.append(s)
        stage += t
    stage /= max(args.steps, 1)
    t_full = ref.extrapolate(stage, hi - lo)
    value = n_vis / t_full / 1e6
    what, how = ref.describe(lo, hi)
    full = (lo == 0 and hi == ref.planes and ref.decim == 1)
    line = {
        "impl": "reference", "metric": METRIC_DEGRID if degrid else METRIC,
        "value": value, "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": float(np.mean(walls)) * 1e3,
        "ms_per_call_extrapolated": t_full * 1e3,
        "whole_call_timed": full,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32" if single else "f64", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {wl['desc']}",
                   "grid_size": int(ref.info["grid_size"]),
                   "support": int(ref.info["support"]),
                   "w_planes": ref.planes, "n_vis": n_vis,
                   "note": "CPU arm: one host, independent of --gpus"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": ref.cores,
                         "kind": ref.kind, "sample": what, "how": how},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit_json(line)


def run_reference_gpu_arm(args):
    """The unmodified reference's CUDA path (baseline/_ref) in this process."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import ref_gpu
    wl = WORKLOADS[args.workload]
    case = wl["make"](0)
    direction = "degrid" if args.direction == "degrid" else "grid"
    res = ref_gpu.run(case, wl["eps"], wl["wstack"], direction, args.steps,
                      max(args.warmup, 1))
    if not res.get("available"):
        emit_json({"impl": "reference-gpu", "unavailable": res.get("why", "?")})
        return
    single = case["uvw"].dtype == np.float32
    emit_json({
        "impl": "reference-gpu",
        "metric": METRIC_DEGRID if direction == "degrid" else METRIC,
        "value": res["Mvis_per_s"], "unit": UNIT, "n_gpus": 1,
        "steps": args.steps, "warmup": max(args.warmup, 1),
        "ms_per_step": res["ms_per_call"], "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None,
        "dtype": "f32" if single else "f64", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {wl['desc']}",
                   "n_vis": int(case["vis"].size)},
        "how": res["how"]})


# ------------------------------------------------------------ product arm

def parity_block(planes, grid_size, budget_s, cores):
    """Block of w-planes [lo, hi) from the middle of the lattice that the
    fp64 CPU oracle can afford in about budget_s seconds (one G x G fp64 FFT
    per plane dominates: ~6e-8 s per cell on 8 threads, measured)."""
    per_plane = 6e-8 * float(grid_size) ** 2 * 8.0 / max(cores, 1) + 0.02
    count = int(max(1, min(planes, budget_s / per_plane)))
    if count >= planes:
        return 0, planes
    lo = max(0, planes // 2 - count // 2)
    return lo, min(planes, lo + count)


def parity_check(torch, local, case, wl, dev, d_in, d_img, d_img0, d_vis_out,
                 directions, budget_s):
    """Product output vs oracle/es_oracle (fp64 restatement of the reference
    algorithm on the same plan) for one block of w-planes, on THIS workload's
    arrays at its real grid and image size. Returns the parity object."""
    from oracle import es_oracle as eo
    info = local.info()
    n = info["image_size"]
    single = not info["is_double"]
    shape_only = np.broadcast_to(np.zeros(1, dtype=case["uvw"].dtype), (n, n))
    o = eo.OracleGridder(case["uvw"], case["freq_hz"], case["vis"], case["weight"],
                         shape_only, case["pixel_size"], case["pixel_size"],
                         wl["eps"], wl["wstack"])
    cores = host_cores(eo.lib().eso_max_threads())
    planes = info["num_total_w_grids"]
    assert o.plan["num_total_w_grids"] == planes and \
        o.plan["grid_size"] == info["grid_size"] and \
        o.plan["w_scale"] == info["w_scale"] and \
        o.plan["min_plane_w"] == info["min_plane_w"], "oracle and product plans differ"
    lo, hi = parity_block(planes, info["grid_size"], budget_s, cores)
    tol = TOL[4 if single else 8]
    out = {"tol": tol, "planes": [lo, hi], "of_planes": planes,
           "oracle": "oracle/es_oracle.c (fp64, plane-sorted), all visibilities of "
                     "the workload, w-planes [lo, hi) only on both sides",
           "metric": "max|out - oracle| / max|oracle|"}
    d_uvw, d_freq, d_vis, d_wgt = d_in
    t0 = time.perf_counter()
    local.set_plane_range(lo, hi, apply_correction=True)
    try:
        if "grid" in directions:
            d_img.zero_()
            local.grid_uvw_es_fft(d_uvw, d_freq, d_vis, d_wgt, d_img)
            got = d_img.cpu().numpy().astype(np.float64)
            want, _ = o.grid_uvw_es_fft(case["uvw"], case["freq_hz"], case["vis"],
                                        case["weight"], planes=(lo, hi),
                                        nthreads=cores)
            out["grid"] = float(np.max(np.abs(got - want)) / np.max(np.abs(want)))
            del got, want
        if "degrid" in directions:
            d_img.copy_(d_img0)
            d_vis_out.zero_()
            local.ifft_grid_uvw_es(d_uvw, d_freq, d_vis_out, d_wgt, d_img)
            got = d_vis_out.cpu().numpy().astype(np.complex128)
            want, _, _ = o.ifft_grid_uvw_es(case["uvw"], case["freq_hz"],
                                            d_img0.cpu().numpy(), planes=(lo, hi),
                                            nthreads=cores)
            out["degrid"] = float(np.max(np.abs(got - want)) / np.max(np.abs(want)))
            del got, want
    finally:
        local.set_plane_range(0, planes, apply_correction=True)
    out["seconds"] = time.perf_counter() - t0
    out["ok"] = all(out[k] <= tol for k in ("grid", "degrid") if k in out)
    return out


def run_product_arm(args):
    import torch
    import torch.distributed as dist
    from ska_sdp_func.grid_data import ShardedGridderUvwEsFft
    from ska_sdp_func.utility import Lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    Lib.handle()  # fails loudly if the CUDA library is missing

    wl = WORKLOADS[args.workload]
    directions = ["grid", "degrid"] if args.direction == "both" else [args.direction]
    case = wl["make"](0)              # ONE fixed snapshot, the same on every rank
    single = case["uvw"].dtype == np.float32
    rows, chans = case["vis"].shape
    n_vis = rows * chans
    n = case["image_size"]
    real_t = torch.float32 if single else torch.float64

    def pinned(a):
        return torch.from_numpy(np.ascontiguousarray(a)).pin_memory()

    h_uvw, h_freq = pinned(case["uvw"]), pinned(case["freq_hz"])
    h_vis, h_wgt = pinned(case["vis"]), pinned(case["weight"])
    h_img = torch.zeros((n, n), dtype=real_t).pin_memory()
    h_img0 = torch.zeros((n, n), dtype=real_t).pin_memory()   # model image
    if case.get("image") is not None:
        h_img0.copy_(torch.from_numpy(case["image"]))
    else:
        h_img0.uniform_(-0.5, 0.5, generator=torch.Generator().manual_seed(5))
    d_uvw, d_freq = h_uvw.to(dev), h_freq.to(dev)
    d_vis, d_wgt = h_vis.to(dev), h_wgt.to(dev)
    d_img = torch.zeros((n, n), dtype=real_t, device=dev)
    d_img0 = h_img0.to(dev)
    d_vis_out = torch.zeros_like(d_vis) if "degrid" in directions else None

    sg = ShardedGridderUvwEsFft(d_uvw, d_freq, d_vis, d_wgt, d_img,
                                case["pixel_size"], case["pixel_size"], wl["eps"],
                                wl["wstack"])
    sg.enable_timing(True)
    local = sg.local
    info = sg.info()
    planes = info["num_total_w_grids"]

    def fence():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return float(x)
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    # ---- parity gate of this run (rank 0; the others wait at the fence)
    parity = None
    if rank == 0 and not args.no_parity:
        parity = parity_check(torch, local, case, wl, dev,
                              (d_uvw, d_freq, d_vis, d_wgt), d_img, d_img0,
                              d_vis_out, directions, args.parity_seconds)
    fence()

    def device_step(direction):
        if direction == "grid":
            d_img.zero_()
            sg.grid_uvw_es_fft(d_uvw, d_freq, d_vis, d_wgt, d_img,
                               root=0 if world > 1 else None)
        else:
            if rank == 0:
                d_img.copy_(d_img0)
            d_vis_out.zero_()
            sg.ifft_grid_uvw_es(d_uvw, d_freq, d_vis_out, d_wgt, d_img, root=0,
                                reduce_to_root=True)

    def measure(direction, steps, warmup, sample_clocks):
        res = {}
        bounds_history = [sg.bounds[direction].tolist()]
        for i in range(warmup):
            device_step(direction)
            if world > 1 and i + 1 < warmup:
                fence()
                # ownership from the per-plane device times just measured
                bounds_history.append(sg.rebalance(direction).tolist())
        fence()
        sampler = ClockSampler(local_rank) if (sample_clocks and rank == 0) else None
        if sampler:
            sampler.start()
        start = torch.cuda.Event(enable_timing=True)
        stop = torch.cuda.Event(enable_timing=True)
        start.record()
        for _ in range(steps):
            device_step(direction)
        stop.record()
        fence()
        elapsed_ms = max_over_ranks(start.elapsed_time(stop))
        stats = local.stats()         # stage times of this rank's last step
        extra = 0
        if sampler and world == 1:
            # a short timed region can end before nvidia-smi's first sample:
            # keep the same load running until there are two samples
            deadline = time.time() + 6.0
            while len(sampler.lines) < 2 and time.time() < deadline:
                device_step(direction)
                torch.cuda.synchronize()
                extra += 1
        if sampler:
            res["clocks"] = sampler.stop()
            if extra:
                res["clocks"]["note"] = (f"{extra} extra untimed steps of the same "
                                         "load were run after the timed region until "
                                         "nvidia-smi had sampled")
        res["ms_per_step"] = elapsed_ms / steps
        res["value"] = n_vis / (res["ms_per_step"] * 1e-3) / 1e6
        res["stats"] = stats
        res["steps"], res["warmup"] = steps, warmup
        lo, hi = sg._own(direction)
        mine = {"rank": rank, "planes": [lo, hi], "local_call_ms": stats["ms_total"],
                "records": stats["num_entries"]}
        if world > 1:
            everyone = [None] * world
            dist.all_gather_object(everyone, mine)
        else:
            everyone = [mine]
        res["ranks"] = everyone
        res["bounds_history"] = bounds_history
        return res

    def measure_e2e(direction, steps):
        a = [h_uvw.numpy(), h_freq.numpy(), h_vis.numpy(), h_wgt.numpy()]
        item = 4 if single else 8
        in_bytes = (h_uvw.numel() + h_freq.numel() + 2 * h_vis.numel()
                    + h_wgt.numel()) * item
        img_bytes = h_img.numel() * item
        if direction == "grid":
            a_img = h_img.numpy()

            def host_step():
                # gridding accumulates into the caller's image (+=, part of the
                # contract): the same host image is simply reused
                sg.grid_uvw_es_fft(a[0], a[1], a[2], a[3], a_img,
                                   root=0 if world > 1 else None)
            h2d = world * (in_bytes + img_bytes)
            d2h = img_bytes
        else:
            # degridding corrects the caller's image in place: every step gets
            # its own pinned copy of the model image, prepared before the clock
            imgs = [h_img0.clone().pin_memory().numpy() for _ in range(steps + 1)]
            h_out = torch.zeros_like(h_vis).pin_memory()
            a_out = h_out.numpy()
            k = [0]

            def host_step():
                img_k = imgs[k[0] % len(imgs)]
                k[0] += 1
                sg.ifft_grid_uvw_es(a[0], a[1], a_out, a[3], img_k, root=0,
                                    reduce_to_root=True)
            h2d = world * in_bytes + img_bytes
            d2h = 2 * h_vis.numel() * item + img_bytes
        host_step()
        fence()
        t0 = time.perf_counter()
        for _ in range(steps):
            host_step()
        fence()
        secs = max_over_ranks((time.perf_counter() - t0) / steps)
        if world == 1:
            st = local.stats()
            h2d, d2h = int(st["h2d_bytes"]), int(st["d2h_bytes"])
        return {"value": n_vis / secs / 1e6, "unit": UNIT,
                "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": secs * 1e3, "steps": steps,
                "how": "host (pinned) arrays through the public call; "
                       + ("C-ABI staging inside the library" if world == 1 else
                          "every rank stages all inputs, root reads the result back")}

    warm = max(args.warmup, 3)
    results, e2e = {}, {}
    for direction in directions:
        head = direction == directions[0]
        steps = args.steps if head else max(1, min(args.steps, args.second_steps))
        results[direction] = measure(direction, steps, warm if head else
                                     max(2, min(warm, 3)), head)
        e2e[direction] = measure_e2e(direction, max(1, min(steps, args.e2e_steps)))

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    report(args, wl, case, info, world, directions, results, e2e, parity, n_vis,
           single, local, sg, torch)
    if world > 1:
        dist.destroy_process_group()


KERNEL_OF = {
    ("grid", "scatter"): "es_scatter2_kernel",
    ("grid", "fft_rows"): "es_fft_grid_rows_kernel",
    ("grid", "fft_cols_screen"): "es_fft_grid_cols_screen_kernel",
    ("degrid", "fft_cols_screen"): "es_fft_degrid_cols_kernel",
    ("degrid", "fft_rows"): "es_fft_degrid_rows_kernel",
    ("degrid", "gather"): "es_gather2_kernel",
}


def stage_table(direction, info, stats, n_vis, peak):
    """Per-stage device time of rank 0's last step with the bytes each stage
    must move (occupancy applied) and, for the line-FFT passes, the
    algorithmic bytes of DESIGN.md section 5."""
    degrid = direction == "degrid"
    c = 16 if info["is_double"] else 8
    r = c // 2
    g = float(info["grid_size"])
    n = float(info["image_size"])
    planes = max(int(stats["planes_processed"]), 1)
    batch = max(int(stats["plane_batch"]), 1)
    nbatch_min = -(-planes // batch)
    batches = nbatch_min
    req, req_total = required_bytes(info, stats, n_vis, batches, degrid)
    ms = {"fft_rows": stats["ms_fft"], "fft_cols_screen": stats["ms_screen"],
          "gather" if degrid else "scatter": stats["ms_scatter_gather"]}
    # DESIGN.md section 5 (dense model, no occupancy): per plane c G N for the
    # column pass (+ image once per batch), c (G^2 + G N) for the row pass
    algo = {"fft_cols_screen": planes * c * g * n + batches * (1 if degrid else 2) * r * n * n,
            "fft_rows": planes * c * (g * g + g * n)}
    stages = {}
    for name, b in req.items():
        t = ms[name]
        gbs = b / (t * 1e-3) / 1e9 if t > 0 else None
        stages[name] = {"ms": t, "required_GB": b / 1e9, "GB/s": gbs,
                        "frac": gbs / peak if gbs else None,
                        "share_of_call": t / stats["ms_total"] if stats["ms_total"] else None}
        if name in algo:
            stages[name]["algorithmic_GB_dense"] = algo[name] / 1e9
    stages["presort"] = {"ms": stats["ms_presort"],
                         "share_of_call": stats["ms_presort"] / stats["ms_total"]
                         if stats["ms_total"] else None}
    return stages, algo, batches, req_total


def roofline_objects(direction, args, info, res, n_vis, world):
    peak, peak_src = hbm_peak()
    stats = res["stats"]
    stages, algo, batches, req_total = stage_table(direction, info, stats, n_vis, peak)
    own = [k for k in stages if k != "presort"]
    dominant = max(own, key=lambda k: stages[k]["ms"])
    dom = stages[dominant]
    kernel = KERNEL_OF[(direction, dominant)]
    if dominant in algo:
        bytes_launch = algo[dominant] / batches
        model = ("DESIGN.md section 5: c G N per plane + image once per batch "
                 "(column pass) / c (G^2 + G N) per plane (row pass), dense")
    else:
        bytes_launch = dom["required_GB"] * 1e9 / batches
        model = "occupied tiles + records (DESIGN.md section 5)"
    launch_ms = dom["ms"] / batches
    achieved = bytes_launch / (launch_ms * 1e-3) / 1e9 if launch_ms > 0 else None
    roofline = {
        "kernel": kernel, "bound": "hbm", "achieved": achieved, "peak": peak,
        "unit": "GB/s", "frac": achieved / peak if achieved else None,
        "frac_required": dom["frac"], "traffic": None, "peak_source": peak_src,
        "bound_note": "HBM is the contract's roofline; ncu shows the line-FFT "
                      "kernels limited by instruction issue / latency, not by "
                      "DRAM (profiles/README.md)",
        "launches_per_step": batches, "algorithmic_bytes_per_launch": bytes_launch,
        "algorithmic_model": model, "avg_launch_ms": launch_ms,
        "measured_on": "rank 0, its own planes" if world > 1 else "the whole call",
    }
    # measured DRAM bytes per launch of this kernel on this workload, from the
    # committed `ncu --set full` captures
    for name in ("r02_traffic.json", "r01_traffic.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                tr = json.load(f).get(args.workload, {}).get(kernel)
        except (OSError, ValueError):
            tr = None
        if tr:
            roofline["traffic"] = tr["dram_bytes_per_launch"]
            roofline["traffic_source"] = tr["source"]
            roofline["traffic_planes_per_launch"] = tr.get("planes_per_launch")
            roofline["ncu_issue_active_pct"] = tr.get("issue_active_pct")
            break
    total_b, no_fft_b = contract_bytes(info, n_vis, direction == "degrid")
    call_s = res["ms_per_step"] * 1e-3
    roofline_call = {
        "bound": "hbm", "algorithmic_GB": total_b / 1e9,
        "achieved": total_b / call_s / 1e9 / world, "peak": peak, "unit": "GB/s per GPU",
        "frac": total_b / call_s / 1e9 / peak / world,
        "model": "SURVEY 8(d) contract: every plane cell written once, FFT = 4 "
                 "plane traversals, w-screen reads the image footprint; the "
                 "implemented algorithm moves far less (frac_required)",
        "required_GB_rank0": req_total / 1e9,
        "frac_required": req_total / (stats["ms_total"] * 1e-3) / 1e9 / peak
        if stats["ms_total"] else None,
        "required_model": "presorted records + occupied 64x64 tiles + strip rows of "
                          "occupied tile rows (2 G^2 + 2 G N per plane when dense) + "
                          "image once per batch; rank 0's planes / rank 0's call time",
        "roofline_Mvis_per_s": n_vis / (total_b / (peak * 1e9)) / 1e6 * world,
    }
    return roofline, roofline_call, stages


def report(args, wl, case, info, world, directions, results, e2e, parity, n_vis,
           single, local, sg, torch):
    head = directions[0]
    res = results[head]
    stats = res["stats"]
    c = 16 if info["is_double"] else 8
    g2 = float(info["grid_size"]) ** 2
    roofline, roofline_call, stages = roofline_objects(head, args, info, res, n_vis,
                                                       world)
    ntile = -(-info["grid_size"] // 64)
    pp = max(int(stats["planes_processed"]), 1)
    line = {
        "metric": METRIC if head == "grid" else METRIC_DEGRID,
        "value": res["value"], "unit": UNIT, "n_gpus": world, "steps": res["steps"],
        "warmup": res["warmup"], "ms_per_step": res["ms_per_step"],
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32" if single else "f64", "data": "synthetic",
        "config": {
            "workload": f"{args.workload}: {wl['desc']}",
            "n_vis": n_vis, "image_size": info["image_size"],
            "grid_size": info["grid_size"], "support": info["support"],
            "w_planes": info["num_total_w_grids"],
            "plane_batch": stats["plane_batch"],
            "fft": "own pruned in-shared-memory line FFTs, w-screen fused"
                   if stats.get("pruned_fft") else "cuFFT 2-D C2C/Z2Z",
            "records_after_presort_rank0": stats["num_entries"],
            "occupied_tile_fraction": stats["occupied_tiles"] / (pp * ntile * ntile),
            "occupied_tile_row_fraction": stats["occupied_tile_rows"] / (pp * ntile),
            "parallelism": "1 GPU" if world == 1 else
                           f"{world} ranks, one fixed snapshot: contiguous w-plane "
                           "blocks per rank (balanced by measured per-plane device "
                           "time) -> partial dirty images -> 1 NCCL sum-reduce -> "
                           "grid correction",
            "l2": "per-step working set (resident w-plane batch "
                  f"{stats['plane_batch'] * c * g2 / 1e6:.0f} MB) exceeds the 126 MB "
                  "L2; no explicit flush",
        },
        "clocks": res.get("clocks"),
        "e2e": e2e[head],
        "gpu_launches": int(stats["kernel_launches"]) * res["steps"],
        "gpu_launches_per_step_rank0": int(stats["kernel_launches"]),
        "cufft_execs_per_step": int(stats["fft_execs"]),
        "ranks": res["ranks"],
        "plane_bounds_history": res["bounds_history"],
        "parity": parity,
        "roofline": roofline, "roofline_call": roofline_call, "stages": stages,
    }
    if len(directions) > 1:
        other = directions[1]
        ro = results[other]
        r2, rc2, st2 = roofline_objects(other, args, info, ro, n_vis, world)
        line[other] = {
            "metric": METRIC_DEGRID if other == "degrid" else METRIC,
            "value": ro["value"], "unit": UNIT, "ms_per_step": ro["ms_per_step"],
            "steps": ro["steps"], "warmup": ro["warmup"], "e2e": e2e[other],
            "ranks": ro["ranks"], "roofline": r2, "roofline_call": rc2,
            "stages": st2,
            "gpu_launches_per_step_rank0": int(ro["stats"]["kernel_launches"]),
        }

    # ---- CPU baseline on this box's cores (bounded sample), N = 1 only
    line["cpu_baseline"] = None
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_object(case, wl["eps"], wl["wstack"],
                                                   n_vis, args.cpu_seconds)

    # ---- the reference's own CUDA implementation on this GPU (own process:
    # the two libraries share a file name and cannot share a process)
    line["ref_gpu"] = None
    if world == 1 and not args.no_ref_gpu:
        del sg, local
        import gc
        gc.collect()
        torch.cuda.empty_cache()
        line["ref_gpu"] = ref_gpu_subprocess(args, directions)
        for d in directions:
            r = (line["ref_gpu"] or {}).get(d)
            if r and r.get("available"):
                ours = results[d]["value"]
                r["ours_over_ref_gpu"] = ours / r["Mvis_per_s"]
    emit_json(line)


def ref_gpu_subprocess(args, directions):
    out = {}
    script = os.path.join(ROOT, "baseline", "ref_gpu.py")
    for d in directions:
        try:
            p = subprocess.run(
                [sys.executable, script, "--workload", args.workload,
                 "--direction", d, "--steps", str(args.ref_gpu_steps),
                 "--warmup", "1"], capture_output=True, text=True,
                timeout=args.ref_gpu_timeout)
            last = [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
            out[d] = json.loads(last[-1]) if last else {
                "available": False, "why": (p.stderr or p.stdout)[-300:]}
        except subprocess.TimeoutExpired:
            out[d] = {"available": False,
                      "why": f"did not finish within {args.ref_gpu_timeout} s"}
        except (OSError, ValueError) as err:
            out[d] = {"available": False, "why": str(err)}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours",
                    choices=["ours", "reference", "reference-gpu"])
    ap.add_argument("--workload", default="C4", choices=sorted(WORKLOADS))
    ap.add_argument("--direction", default="both",
                    choices=["both", "grid", "degrid"])
    ap.add_argument("--second-steps", type=int, default=2,
                    help="timed steps of the second direction (degridding)")
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--parity-seconds", type=float, default=25.0)
    ap.add_argument("--ref-gpu-steps", type=int, default=1)
    ap.add_argument("--ref-gpu-timeout", type=float, default=600.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-ref-gpu", action="store_true")
    args = ap.parse_args()
    # stdout carries exactly ONE line, the JSON: anything libraries print there
    # (NCCL's version banner under torchrun) is sent to stderr instead
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference_arm(args)
    elif args.impl == "reference-gpu":
        run_reference_gpu_arm(args)
    else:
        run_product_arm(args)


if __name__ == "__main__":
    main()
