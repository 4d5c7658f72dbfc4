"""Times the fused mapping-loss kernel (forward + backward) against the reference's chain of torch ops
(oracle.mapping_loss on the GPU = the same ops slam_utils.py:71-105 issues) at the Replica frame size.
    gpurun -- 'python tests/perf/bench_loss.py'
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "splat-slam_b200"))
from diff_gaussian_rasterization.mapping_loss import fused_mapping_loss  # noqa: E402
from oracle.mapping_loss import mapping_loss, synthetic_loss_case  # noqa: E402

H, W = 680, 1200
c = {k: v.cuda() for k, v in synthetic_loss_case(H, W, 0).items()}


def step(fn):
    leaves = [c[k].detach().requires_grad_(True) for k in ("image", "depth", "exposure_a", "exposure_b")]
    loss = fn(leaves[0], leaves[1], leaves[2], leaves[3], c["gt_image"], c["gt_depth"], 0.95, 0.01)
    loss.backward()


for name, fn in (("torch_chain", mapping_loss), ("fused", fused_mapping_loss)):
    for _ in range(20):
        step(fn)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(200):
        step(fn)
    e1.record()
    torch.cuda.synchronize()
    print(f"{name}: {e0.elapsed_time(e1) / 200 * 1e3:.1f} us per loss fwd+bwd ({H}x{W})")
