"""Times distCUDA2 (ours) against the compiled reference simple-knn (oracle/_ref/ref_simple_knn) on a B200.
    gpurun -- 'python tests/perf/bench_knn.py'
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "splat-slam_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
from diff_gaussian_rasterization.knn import distCUDA2  # noqa: E402
from oracle.knn import synthetic_points  # noqa: E402

try:
    from ref_simple_knn._C import distCUDA2 as ref_fn
except Exception:
    ref_fn = None

for P in (25_000, 100_000, 1_000_000):
    pts = torch.from_numpy(synthetic_points(P, 0, "surface")).cuda()
    line = f"P={P}:"
    for name, fn in (("reference", ref_fn), ("ours", distCUDA2)):
        if fn is None:
            continue
        for _ in range(3):
            fn(pts)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 20 if P <= 100_000 else 5
        e0.record()
        for _ in range(n):
            fn(pts)
        e1.record()
        torch.cuda.synchronize()
        line += f"  {name} {e0.elapsed_time(e1) / n * 1e3:.0f} us"
    print(line)
