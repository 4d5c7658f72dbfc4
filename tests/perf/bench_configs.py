"""BASELINE.json configs 1 and 5 (the bench covers config 3/4's shape, tests/perf/bench_tracking.py config 2):
  config 1: Replica 1200x680, 500 k Gaussians, SH degree 0, fwd+bwd with pose gradients, one view per iteration
  config 5: ScanNet 1296x968 final colour refinement, 3 M Gaussians, SH degree 3, one view per iteration
Times this library and the compiled reference extension on the same inputs (single stream, one view at a time).
    gpurun -- 'python tests/perf/bench_configs.py'
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "splat-slam_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
import diff_gaussian_rasterization as ours  # noqa: E402
from diff_gaussian_rasterization import synthetic as syn  # noqa: E402

try:
    import ref_dgr as ref
except Exception:
    ref = None

dev = torch.device("cuda:0")
for label, cam, P, M, D, iters in (("config 1", "replica", 500_000, 1, 0, 30), ("config 5", "scannet", 3_000_000, 16, 3, 10)):
    g = {k: v.to(dev) for k, v in syn.make_gaussians(cam, P, seed=0, sh_coeffs=M).items() if torch.is_tensor(v)}
    leaves = {k: g[k].requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
    views = [syn.make_view(cam, syn.window_pose(v)) for v in range(4)]
    mats = [[m.to(dev) for m in (v.viewmatrix, v.projmatrix, v.projmatrix_raw, v.campos)] for v in views]
    Wc, Wd = (t.to(dev) for t in syn.upstream_grads(cam))
    bg = torch.zeros(3, device=dev)
    line = f"{label}: {cam} {views[0].image_width}x{views[0].image_height}, P={P}, SH degree {D}:"
    for name, mod in (("reference", ref), ("ours", ours)):
        if mod is None:
            continue

        def step(i):
            v, m = views[i % 4], mats[i % 4]
            S = mod.GaussianRasterizationSettings(
                image_height=v.image_height, image_width=v.image_width, tanfovx=v.tanfovx, tanfovy=v.tanfovy, bg=bg,
                scale_modifier=1.0, viewmatrix=m[0], projmatrix=m[1], projmatrix_raw=m[2], sh_degree=D, campos=m[3],
                prefiltered=False, debug=False)
            theta = torch.zeros(3, device=dev, requires_grad=True)
            rho = torch.zeros(3, device=dev, requires_grad=True)
            color, radii, depth, opacity, n_touched = mod.GaussianRasterizer(S)(
                means3D=leaves["means3D"], means2D=torch.zeros_like(leaves["means3D"], requires_grad=True),
                opacities=leaves["opacities"], shs=leaves["shs"], scales=leaves["scales"], rotations=leaves["rotations"],
                theta=theta, rho=rho)
            torch.autograd.backward([color, depth], [Wc, Wd])
            for t in leaves.values():
                t.grad = None

        for i in range(3):
            step(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(iters):
            step(i)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        line += f"  {name} {ms:.3f} ms/it ({1e3 / ms:.0f} it/s)"
    print(line)
    del g, leaves
    torch.cuda.empty_cache()
