"""Runs each of the kernels around the rasterizer once at bench size, for an ncu capture:
    ncu --set full --clock-control none --import-source on -k regex:'adam_step|mapping_loss|densify_stats|knn_search|deform_points|activations_' \
        -o gpurun_out/r01_f python tests/perf/ncu_f_kernels.py
"""
import os
import sys
import types

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "splat-slam_b200"))
from diff_gaussian_rasterization.activations import fused_activations  # noqa: E402
from diff_gaussian_rasterization.adam import GaussianAdam  # noqa: E402
from diff_gaussian_rasterization.deform import deform_points_  # noqa: E402
from diff_gaussian_rasterization.densify import add_densification_stats_window  # noqa: E402
from diff_gaussian_rasterization.knn import distCUDA2  # noqa: E402
from diff_gaussian_rasterization.mapping_loss import fused_mapping_loss  # noqa: E402
import oracle.activations as oa  # noqa: E402
import oracle.deform as od  # noqa: E402
import oracle.densify as odn  # noqa: E402
import oracle.knn as ok  # noqa: E402
import oracle.mapping_loss as oml  # noqa: E402

P = 1_000_000
dev = "cuda"
# mapping loss 1200x680
c = {k: v.to(dev) for k, v in oml.synthetic_loss_case(680, 1200, 0).items()}
leaves = [c[k].requires_grad_(True) for k in ("image", "depth", "exposure_a", "exposure_b")]
fused_mapping_loss(*leaves, c["gt_image"], c["gt_depth"], 0.95, 0.01).backward()
# activations
raw = [t.to(dev).requires_grad_(True) for t in oa.synthetic_raw(P, 0)]
out = fused_activations(*raw)
torch.autograd.backward(list(out), [torch.ones_like(o) for o in out])
# adam, 14 M parameters
groups = [{"params": [torch.nn.Parameter(torch.randn(P, w, device=dev))], "lr": 1e-3} for w in (3, 3, 1, 3, 4)]
opt = GaussianAdam(groups, lr=0.0, eps=1e-15)
for g in groups:
    g["params"][0].grad = torch.randn_like(g["params"][0])
opt.step()
# densification stats, 12 views
d = odn.synthetic_densify_case(P, 12, 0)
st = types.SimpleNamespace(max_radii2D=d["max_radii2D"].to(dev), xyz_gradient_accum=d["xyz_gradient_accum"].to(dev),
                           denom=d["denom"].to(dev))
add_densification_stats_window(st, [g.to(dev) for g in d["grads"]], [r.to(dev) for r in d["radii"]])
# knn, 100 k points
distCUDA2(torch.from_numpy(ok.synthetic_points(100_000, 0, "surface")).to(dev))
# deformation, 100 keyframes
dc = od.synthetic_deform_case(P=P, n_kf=99, H=340, W=600, seed=0)
ups = [(5 * k, dc["w2c_new"][k], dc["w2c_old"][k], dc["depth_new"][k], dc["depth_old"][k], None if k % 2 else "rigid")
       for k in range(100)]
deform_points_(dc["xyz"].to(dev), dc["rotation"].to(dev), dc["scaling"].to(dev), dc["kf_ids"].to(dev), ups, dc["intrinsics"])
torch.cuda.synchronize()
print("done")
