"""Times the fused densification bookkeeping against the reference's per-view torch indexing loop
(src/mapper.py:519-529 + gaussian_model.py:738-742, run here on the GPU as the reference runs it) for 12 views.
    gpurun -- 'python tests/perf/bench_densify.py'
"""
import os
import sys
import types

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "splat-slam_b200"))
from diff_gaussian_rasterization.densify import add_densification_stats_window  # noqa: E402
from oracle.densify import densification_stats_window, synthetic_densify_case  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
K = 12
c = synthetic_densify_case(P, K, 0)
radii = [r.cuda() for r in c["radii"]]
grads = [g.cuda() for g in c["grads"]]
st = types.SimpleNamespace(max_radii2D=c["max_radii2D"].cuda(), xyz_gradient_accum=c["xyz_gradient_accum"].cuda(),
                           denom=c["denom"].cuda())


def torch_loop():
    densification_stats_window(st.max_radii2D, st.xyz_gradient_accum, st.denom, radii, grads)


def fused():
    add_densification_stats_window(st, grads, radii)


for name, fn in (("torch indexing loop", torch_loop), ("fused", fused)):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"{name}: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us per iteration ({K} views, {P} Gaussians)")
