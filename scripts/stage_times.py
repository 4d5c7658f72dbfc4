"""Per-stage CUDA-event timings (library profile mode) of one view forward + backward for any camera / size / SH degree.
    gpurun -- 'python scripts/stage_times.py scannet 3000000 3'
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "splat-slam_b200"))
import diff_gaussian_rasterization as ours  # noqa: E402
from diff_gaussian_rasterization import _cabi, synthetic as syn  # noqa: E402

cam = sys.argv[1] if len(sys.argv) > 1 else "replica"
P = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
D = int(sys.argv[3]) if len(sys.argv) > 3 else 0
M = (D + 1) ** 2
dev = torch.device("cuda:0")
g = {k: v.to(dev) for k, v in syn.make_gaussians(cam, P, seed=0, sh_coeffs=M).items() if torch.is_tensor(v)}
leaves = {k: g[k].requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
v = syn.make_view(cam, syn.window_pose(1))
m = [x.to(dev) for x in (v.viewmatrix, v.projmatrix, v.projmatrix_raw, v.campos)]
Wc, Wd = (t.to(dev) for t in syn.upstream_grads(cam))
bg = torch.zeros(3, device=dev)


def step():
    S = ours.GaussianRasterizationSettings(
        image_height=v.image_height, image_width=v.image_width, tanfovx=v.tanfovx, tanfovy=v.tanfovy, bg=bg,
        scale_modifier=1.0, viewmatrix=m[0], projmatrix=m[1], projmatrix_raw=m[2], sh_degree=D, campos=m[3],
        prefiltered=False, debug=False)
    theta = torch.zeros(3, device=dev, requires_grad=True)
    rho = torch.zeros(3, device=dev, requires_grad=True)
    c, r, d, o, n = ours.GaussianRasterizer(S)(
        means3D=leaves["means3D"], means2D=torch.zeros_like(leaves["means3D"], requires_grad=True),
        opacities=leaves["opacities"], shs=leaves["shs"], scales=leaves["scales"], rotations=leaves["rotations"],
        theta=theta, rho=rho)
    torch.autograd.backward([c, d], [Wc, Wd])
    for t in leaves.values():
        t.grad = None
    return int((r > 0).sum())


for _ in range(3):
    V = step()
_cabi.profile(True)
_cabi.profile_read(reset=True)
for _ in range(5):
    step()
torch.cuda.synchronize()
prof = _cabi.profile_read(reset=True)
_cabi.profile(False)
st = {k: round(ms / max(c, 1), 4) for k, (ms, c) in prof.items()}
print(f"{cam} {v.image_width}x{v.image_height} P={P} SH{D}: V={V} R={int(ours._PinnedInt.get(dev).item())} sum={sum(st.values()):.3f} ms", st)
