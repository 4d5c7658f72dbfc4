#!/usr/bin/env python
"""bench.py — fwd+bwd rasterization throughput of the B200-native Gaussian rasterizer (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--scaling weak|strong]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
           bench.py --gpus N --steps K --warmup W

Workload (config.workload): Replica-shape 1200x680 camera, 1M synthetic Gaussians, SH degree 0, a mapping window
of 10 keyframe views per GPU; one "step" = every view of the rank's window forward + backward with pose
gradients through the public API (`GaussianRasterizer(...)` + autograd), the Gaussian gradients accumulating in
one flat buffer, then (N > 1) ONE NCCL all-reduce of that buffer.  One "it" of the metric = one view fwd+bwd.

  value : views/s over all ranks, inputs resident in HBM, CUDA-event timed, max over ranks.
  e2e   : the same step with each view's camera matrices and upstream-gradient images copied host->device
          from pinned memory inside the timed region and the step's scalar result read back.
  roofline : dominant kernel (render_backward), algorithmic bytes (84*R + 24*N) / its live
          CUDA-event duration, against MEASURED_PEAKS.json hbm_gbs.
  cpu_baseline : the CPU oracle (pure PyTorch port) on BASELINE config 1 (100k Gaussians, 640x480, 1 view).
  --impl reference : the UNMODIFIED reference CUDA extension (oracle/_ref/ref_dgr, sm_100 build) on the same
          workload through its own public API; falls back to the CPU oracle if that build is absent.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "splat-slam_b200"), os.path.join(ROOT, "oracle", "_ref")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

CAMERA = "replica"
P_GAUSS = 1_000_000
VIEWS_PER_RANK = 10
METRIC = "fwd+bwd rasterize it/s @1200x680, 1M Gaussians"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--gaussians", type=int, default=P_GAUSS)
    ap.add_argument("--views", type=int, default=VIEWS_PER_RANK)
    ap.add_argument("--camera", default=CAMERA)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--priority-lanes", type=int, default=1, help="1: short kernels on a high-priority side stream")
    ap.add_argument("--streams", type=int, default=5,
                    help="views of a step are issued round-robin on this many CUDA streams (ours only; the reference "
                         "extension launches on the legacy default stream and always runs with 1)")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------------------
def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks/throttle reasons sampled every 200 ms DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "200"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1]))
                mx.append(float(c[2]))
            except ValueError:
                continue
            for n, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.f.name)
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# ---------------------------------------------------------------------------------------------------------
class Workload:
    """Gaussians (resident on the device, like model weights), the rank's views and upstream gradients."""

    def __init__(self, mod, args, device, rank, world):
        from diff_gaussian_rasterization import synthetic as syn
        from diff_gaussian_rasterization.window import FlatGrads, ViewStreams, partition_views
        self.mod, self.device, self.syn = mod, device, syn
        g = syn.make_gaussians(args.camera, args.gaussians, seed=0, sh_coeffs=1)
        self.params = {k: g[k].to(device).requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
        self.flat = FlatGrads(self.params)
        if args.scaling == "weak":
            self.view_ids = [rank * args.views + i for i in range(args.views)]
            self.global_views = args.views * world
        else:
            self.view_ids = partition_views(args.views, world, rank)
            self.global_views = args.views
        W, H = syn.CAMERAS[args.camera][:2]
        self.W, self.H = W, H
        self.host_views = [syn.make_view(args.camera, syn.window_pose(v)) for v in self.view_ids]
        Wc, Wd = syn.upstream_grads(args.camera)
        # host copies in pinned memory (e2e leg) and device-resident copies (value leg)
        self.h_mats = [torch.stack([v.viewmatrix.reshape(-1), v.projmatrix.reshape(-1), v.projmatrix_raw.reshape(-1),
                                    torch.cat([v.campos, torch.zeros(13)])]).pin_memory() for v in self.host_views]
        self.h_up = torch.cat([Wc, Wd]).contiguous().pin_memory()           # (4,H,W)
        self.d_mats = [m.to(device) for m in self.h_mats]
        self.d_up = self.h_up.to(device)
        self.bg = torch.zeros(3, device=device)
        self.theta = [torch.zeros(3, device=device, requires_grad=True) for _ in self.view_ids]
        self.rho = [torch.zeros(3, device=device, requires_grad=True) for _ in self.view_ids]
        self.means2D = torch.zeros_like(self.params["means3D"], requires_grad=True)
        # e2e leg: double-buffered device staging filled by a copy stream, so the next view's H2D overlaps compute
        self.streams = ViewStreams(device, args.streams if mod.__name__ == "diff_gaussian_rasterization" else 1,
                                   priority_lanes=bool(args.priority_lanes))
        self.nbuf = max(2, len(self.streams) + 1)      # one staging buffer per view in flight, plus the one being filled
        self.stage_mats = [torch.empty_like(self.d_mats[0]) for _ in range(self.nbuf)] if self.d_mats else None
        self.stage_up = [torch.empty_like(self.d_up) for _ in range(self.nbuf)]
        self.copy_stream = torch.cuda.Stream(device=device)
        self.ev_ready = [torch.cuda.Event() for _ in range(self.nbuf)]
        self.ev_mats = [torch.cuda.Event() for _ in range(self.nbuf)]
        self.ev_free = [torch.cuda.Event() for _ in range(self.nbuf)]
        self.last = {}

    def settings(self, i, mats):
        v = self.host_views[i]
        return self.mod.GaussianRasterizationSettings(
            image_height=v.image_height, image_width=v.image_width, tanfovx=v.tanfovx, tanfovy=v.tanfovy, bg=self.bg,
            scale_modifier=1.0, viewmatrix=mats[0].view(4, 4), projmatrix=mats[1].view(4, 4),
            projmatrix_raw=mats[2].view(4, 4), sh_degree=0, campos=mats[3][:3], prefiltered=False, debug=False)

    def render_backward(self, i, mats, up, up_ready=None):
        p = self.params
        rast = self.mod.GaussianRasterizer(raster_settings=self.settings(i, mats))
        color, radii, depth, opacity, n_touched = rast(
            means3D=p["means3D"], means2D=self.means2D, opacities=p["opacities"], shs=p["shs"], scales=p["scales"],
            rotations=p["rotations"], theta=self.theta[i], rho=self.rho[i])
        if up_ready is not None:      # e2e leg: the upstream images are only needed from here on
            torch.cuda.current_stream(self.device).wait_event(up_ready)
        torch.autograd.backward([color, depth], [up[:3], up[3:4]])
        self.last = dict(radii=radii, color=color, depth=depth)
        return color, depth

    def step_resident(self):
        self.flat.zero_()
        # views are independent until their gradients are accumulated: window.ViewStreams issues them round-robin on a
        # few streams (one view's latency-bound sort passes hide under another view's render kernels)
        self.streams.fork()
        for i in range(len(self.view_ids)):
            with self.streams.view(i):
                self.render_backward(i, self.d_mats[i], self.d_up)
        self.streams.join()
        self.flat.all_reduce()

    def step_e2e(self):
        """Same step, host buffers: H2D of each view's matrices + upstream images, D2H of the scalar result."""
        main = torch.cuda.current_stream(self.device)
        n = len(self.view_ids)

        def prefetch(i):
            b = i % self.nbuf
            with torch.cuda.stream(self.copy_stream):
                self.copy_stream.wait_event(self.ev_free[b])       # the view that used this buffer is done
                self.stage_mats[b].copy_(self.h_mats[i], non_blocking=True)
                self.ev_mats[b].record(self.copy_stream)           # 256 B: the forward can start right away
                self.stage_up[b].copy_(self.h_up, non_blocking=True)
                self.ev_ready[b].record(self.copy_stream)          # 13 MB: needed by the backward

        self.flat.zero_()
        lanes = max(1, len(self.streams))
        acc = [torch.zeros((), device=self.device) for _ in range(lanes)]      # one running result per stream
        for b in range(self.nbuf):
            self.ev_free[b].record(main)
        self.streams.fork()
        if n:
            prefetch(0)
        for i in range(n):
            b = i % self.nbuf
            if i + 1 < n:
                prefetch(i + 1)
            with self.streams.view(i):
                st = self.streams.stream(i)
                st.wait_event(self.ev_mats[b])
                color, depth = self.render_backward(i, self.stage_mats[b], self.stage_up[b], up_ready=self.ev_ready[b])
                up = self.stage_up[b]
                acc[i % lanes] = (acc[i % lanes] + torch.dot(color.detach().reshape(-1), up[:3].reshape(-1))
                                  + torch.dot(depth.detach().reshape(-1), up[3:4].reshape(-1)))
                self.ev_free[b].record(st)
        self.streams.join()
        self.flat.all_reduce()
        return float(torch.stack(acc).sum().item())

    def h2d_bytes_per_step(self):
        return len(self.view_ids) * (self.h_mats[0].numel() * 4 + self.h_up.numel() * 4) if self.view_ids else 0

    def d2h_bytes_per_step(self):
        return 4 + 4 * len(self.view_ids)      # scalar result + the per-view num_rendered hand-off


def timed(fn, steps, warmup, device, world):
    for _ in range(warmup):
        fn()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(device)
    ms = torch.tensor([e0.elapsed_time(e1)], device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def cpu_baseline_port(threads=None):
    """CPU oracle (pure-PyTorch port) on BASELINE.json config 1: 100k Gaussians, 640x480, one pose, fwd+bwd."""
    from diff_gaussian_rasterization import synthetic as syn
    from oracle import torch_rasterizer as orc
    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    cam, P = "tum", 100_000
    g = syn.make_gaussians(cam, P, seed=0, sh_coeffs=1)
    v = syn.make_view(cam)
    Wc, Wd = syn.upstream_grads(cam)
    S = orc.Settings(v.image_height, v.image_width, v.tanfovx, v.tanfovy, torch.zeros(3), 1.0, v.viewmatrix,
                     v.projmatrix, v.projmatrix_raw, 0, v.campos)
    t0 = time.perf_counter()
    orc.rasterize(g["means3D"], g["opacities"], S, shs=g["shs"], scales=g["scales"], rotations=g["rotations"],
                  dL_dcolor=Wc, dL_ddepth=Wd)
    dt = time.perf_counter() - t0
    return {"value": 1.0 / dt, "unit": "it/s", "cores": cores, "kind": "port",
            "sample": "BASELINE config 1: 1 view fwd+bwd, 100k Gaussians, 640x480 (TUM intrinsics), CPU oracle "
                      f"oracle/torch_rasterizer.py, {dt:.1f} s"}


def main():
    args = parse()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    n_gpus = world if world > 1 else 1

    if args.impl == "reference" and rank != 0:
        return 0                       # reference arm: rank 0 alone runs and prints
    ref_world = 1 if args.impl == "reference" else world

    if not torch.cuda.is_available():
        if args.impl == "reference":
            cb = cpu_baseline_port()
            print(json.dumps({"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": "it/s",
                              "n_gpus": args.gpus, "steps": 1, "warmup": 0, "higher_is_better": True,
                              "cpu_baseline": cb, "e2e": {"value": cb["value"], "unit": "it/s",
                                                          "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
            return 0
        raise SystemExit("bench.py needs a CUDA device: the rasterizer has no CPU path")

    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    if ref_world > 1:
        dist.init_process_group("nccl", device_id=device)

    kind = "ours"
    if args.impl == "reference":
        try:
            import ref_dgr as mod
            kind = "reference-cuda"
        except Exception as ex:  # reference build absent: time the CPU port instead (tier framing, item 4)
            cb = cpu_baseline_port()
            print(json.dumps({"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": "it/s",
                              "n_gpus": n_gpus, "steps": 1, "warmup": 0, "ms_per_step": 1e3 / cb["value"],
                              "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32",
                              "data": "synthetic", "config": {"workload": "BASELINE config 1 (CPU port)"},
                              "cpu_baseline": cb, "note": f"oracle/_ref/ref_dgr not importable: {ex}",
                              "e2e": {"value": cb["value"], "unit": "it/s", "h2d_bytes_per_step": 0,
                                      "d2h_bytes_per_step": 0}}))
            return 0
    else:
        import diff_gaussian_rasterization as mod

    wl = Workload(mod, args, device, rank if ref_world > 1 else 0, ref_world)
    views_per_step_global = wl.global_views

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    launches0 = 0
    if kind == "ours":
        from diff_gaussian_rasterization import _cabi
        _cabi.priority_lanes(bool(args.priority_lanes) and args.streams > 1)
    ms_value = timed(wl.step_resident, args.steps, args.warmup, device, ref_world)
    if kind == "ours":      # kernels of this library per step, counted by the library itself, x timed steps
        _cabi.launch_count(reset=True)
        wl.step_resident()
        launches0 = _cabi.launch_count(reset=True) * args.steps
    ms_e2e = timed(wl.step_e2e, args.steps, max(1, args.warmup // 2), device, ref_world)
    clocks = sampler.stop() if sampler else None

    n_streams = max(1, len(wl.streams))
    value = views_per_step_global * args.steps / (ms_value * 1e-3)
    e2e = views_per_step_global * args.steps / (ms_e2e * 1e-3)

    # ---- per-kernel durations (live CUDA events on the launching stream) and roofline of the dominant kernel
    roofline, stages, step_roof = None, None, None
    radii = wl.last["radii"]
    Pn = args.gaussians
    V = int((radii > 0).sum())
    N = wl.W * wl.H
    peak, peak_src = peaks()
    if kind == "ours":
        R = int(mod._PinnedInt.get(device).item())
        from diff_gaussian_rasterization.window import ViewStreams
        wl.streams = ViewStreams(device, 1)     # per-stage times are taken with the views back to back on one stream
        _cabi.profile(True)
        _cabi.profile_read(reset=True)
        for _ in range(2):
            wl.step_resident()
        torch.cuda.synchronize(device)
        prof = _cabi.profile_read(reset=True)
        _cabi.profile(False)
        stages = {k: (ms / max(c, 1)) for k, (ms, c) in prof.items()}
        total = sum(stages.values())
        dom = max(stages, key=stages.get)
        alg = {"render_backward": 84 * R + 24 * N, "render_forward": 48 * R + 28 * N,
               "radix_sort": 24 * R, "preprocess_forward": 44 * Pn + 79 * V, "preprocess_backward": 140 * Pn + 351 * V,
               "duplicate_with_keys": 8 * Pn + 12 * V + 12 * R, "ranges_and_records": 12 * R + 44 * R + 48 * R}
        a_bytes = alg.get(dom, 0)
        ach = a_bytes / (stages[dom] * 1e-3) / 1e9
        # dram__bytes_read.sum + dram__bytes_write.sum per launch from the ncu --set full capture of this workload
        # (profiles/r01_render_final_summary.md); only valid for the default scene
        default_scene = (args.camera == CAMERA and args.gaussians == P_GAUSS)
        traffic = {"render_backward": 248.1e6, "render_forward": 174.8e6}.get(dom) if default_scene else None
        roofline = {"kernel": dom, "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                    "traffic": traffic, "algorithmic_bytes_per_launch": a_bytes, "ms_per_launch": stages[dom],
                    "share_of_kernel_time": stages[dom] / total, "peak_source": peak_src,
                    "note": "render kernels reuse each 48-B record 256x; their limiter is FP32/SFU issue, shuffles and "
                            "L2 atomics, not HBM (SURVEY.md §8d) — see profiles/"}
    else:
        R = None
    if R is not None:
        M = 1
        b_alg = (244 + 12 * M) * Pn + (406 + 36 * M) * V + 172 * R + 52 * N
        t_view = ms_value * 1e-3 / (args.steps * len(wl.view_ids))
        step_roof = {"algorithmic_bytes_per_view": b_alg, "achieved_gbs": b_alg / t_view / 1e9,
                     "frac": b_alg / t_view / 1e9 / peak, "P": Pn, "V": V, "R": R, "N": N}

    cb = None
    if rank == 0 and n_gpus == 1 and args.impl == "ours" and not args.no_cpu_baseline:
        cb = cpu_baseline_port()

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "it/s", "n_gpus": n_gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_value / args.steps, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.camera}_{wl.W}x{wl.H}_P{args.gaussians}_sh0_window{args.views}"
                                   + ("_per_gpu" if args.scaling == "weak" else "_total"),
                       "views_per_step_global": views_per_step_global, "views_per_rank": len(wl.view_ids),
                       "it": "one view forward+backward with pose gradients",
                       "collective": "all_reduce(sum,fp32) of the flat Gaussian-gradient buffer, "
                                     f"{wl.flat.flat.numel() * 4} B/step" if ref_world > 1 else "none (1 GPU)",
                       "l2": "inputs larger than L2: ~0.4 GB of state per view, 10 distinct views per step",
                       "resident": "Gaussian parameters stay in HBM (model state)",
                       "streams": n_streams,
                       "priority_lanes": bool(args.priority_lanes) and n_streams > 1 and kind == "ours"},
            "e2e": {"value": e2e, "unit": "it/s", "h2d_bytes_per_step": wl.h2d_bytes_per_step(),
                    "d2h_bytes_per_step": wl.d2h_bytes_per_step(), "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": int(launches0) if kind == "ours" else 0,
            "clocks": clocks,
        }
        if roofline:
            line["roofline"] = roofline
        if stages:
            line["stage_ms"] = stages
        if step_roof:
            line["roofline_step"] = step_roof
        if cb:
            line["cpu_baseline"] = cb
        if args.impl == "reference":
            line["impl"] = "reference"
            line["reference_kind"] = "unmodified reference CUDA extension, sm_100 build (oracle/_ref/ref_dgr)"
            line["cpu_baseline"] = {"value": value, "unit": "it/s", "cores": 0, "kind": "reference",
                                    "sample": "reference CUDA extension on the same GPU workload (the reference has "
                                              "no CPU rasterizer, DGR/rasterize_points.cu:75)"}
        print(json.dumps(line))
    if ref_world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
