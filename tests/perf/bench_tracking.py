"""BASELINE.json config 3: TUM-RGBD-shape 640x480 tracking, pose-only (theta, rho) gradients, 200k Gaussians,
100 iterations. Times this library and the compiled reference extension on the same inputs.
    gpurun -- 'python tests/perf/bench_tracking.py'
"""
import os
import sys
import time

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

CAM, P, ITERS = "tum", int(os.environ.get("P", 200_000)), 100
dev = torch.device("cuda:0")
g = {k: v.to(dev) for k, v in syn.make_gaussians(CAM, P, seed=0, sh_coeffs=1).items() if torch.is_tensor(v)}
view = syn.make_view(CAM, syn.window_pose(1))
mats = [m.to(dev) for m in (view.viewmatrix, view.projmatrix, view.projmatrix_raw, view.campos)]
Wc, Wd = (t.to(dev) for t in syn.upstream_grads(CAM))
bg = torch.zeros(3, device=dev)


def run(mod, iters):
    theta = torch.zeros(3, device=dev, requires_grad=True)
    rho = torch.zeros(3, device=dev, requires_grad=True)
    S = mod.GaussianRasterizationSettings(
        image_height=view.image_height, image_width=view.image_width, tanfovx=view.tanfovx, tanfovy=view.tanfovy, bg=bg,
        scale_modifier=1.0, viewmatrix=mats[0], projmatrix=mats[1], projmatrix_raw=mats[2], sh_degree=0, campos=mats[3],
        prefiltered=False, debug=False)
    rast = mod.GaussianRasterizer(S)
    for _ in range(iters):
        color, radii, depth, opacity, n_touched = rast(
            means3D=g["means3D"], means2D=torch.zeros_like(g["means3D"]), opacities=g["opacities"], shs=g["shs"],
            scales=g["scales"], rotations=g["rotations"], theta=theta, rho=rho)
        torch.autograd.backward([color, depth], [Wc, Wd])
    return theta.grad, rho.grad


for name, mod in (("reference", ref), ("ours", ours)):
    if mod is None:
        continue
    run(mod, 10)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    th, rh = run(mod, ITERS)
    e1.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    print(f"{name}: {e0.elapsed_time(e1) / ITERS:.3f} ms/it (events), {wall / ITERS * 1e3:.3f} ms/it (wall), "
          f"{ITERS / (e0.elapsed_time(e1) * 1e-3):.0f} it/s; theta.grad {th.tolist()}")
