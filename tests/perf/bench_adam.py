"""Times the fused Adam step against torch.optim.Adam (what the reference steps, gaussian_model.py:302) on the six
Gaussian parameter groups at 1 M Gaussians, SH degree 0 (14 M parameters).
    gpurun -- 'python tests/perf/bench_adam.py'
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "splat-slam_b200"))
from diff_gaussian_rasterization.adam import GaussianAdam  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
WIDTHS = (("xyz", 3, 1e-3), ("f_dc", 3, 2.5e-3), ("opacity", 1, 0.05), ("scaling", 3, 6e-3), ("rotation", 4, 1e-3))
n_param = P * sum(w for _, w, _ in WIDTHS)

for name, cls in (("torch.optim.Adam", torch.optim.Adam), ("fused", GaussianAdam)):
    torch.manual_seed(0)
    groups = [{"params": [torch.nn.Parameter(torch.randn(P, w, device="cuda"))], "lr": lr, "name": n} for n, w, lr in WIDTHS]
    opt = cls(groups, lr=0.0, eps=1e-15)
    for g in groups:
        g["params"][0].grad = torch.randn_like(g["params"][0])
    for _ in range(5):
        opt.step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        opt.step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 50
    print(f"{name}: {ms * 1e3:.1f} us per step, {n_param / 1e6:.0f} M params, "
          f"{32 * n_param / ms / 1e6:.0f} GB/s at the fused kernel's 32 B/param")
