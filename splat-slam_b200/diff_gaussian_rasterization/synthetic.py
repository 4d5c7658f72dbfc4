"""Synthetic Gaussian clouds, cameras and upstream gradients for parity tests and bench.py.

Everything is generated on the CPU from fixed seeds so that the reference
extension, the CPU oracle and the CUDA path all see identical bits.

Camera conventions follow the reference's callers:
  * viewmatrix      = W2C^T                      (thirdparty/monogs/utils/camera_utils.py:94-96)
  * projmatrix_raw  = getProjectionMatrix2(...)^T (thirdparty/gaussian_splatting/utils/graphics_utils.py:72-93,
                                                  src/mapper.py:841-850, znear=0.01, zfar=100)
  * projmatrix      = viewmatrix @ projmatrix_raw (camera_utils.py:98-104)
  * campos          = inverse(viewmatrix)[3, :3]  (camera_utils.py:106-108)
  * tanfov          = dim / (2 f)                 (src/utils/datasets.py:106-107 + focal2fov)
"""
from __future__ import annotations

import math
from typing import Dict, NamedTuple, Optional

import torch

# name -> (W, H, fx, fy, cx, cy); intrinsics from the reference's dataset configs
CAMERAS: Dict[str, tuple] = {
    # configs/Replica/replica.yaml:43-48
    "replica": (1200, 680, 600.0, 600.0, 599.5, 339.5),
    # configs/TUM_RGBD/freiburg1_desk.yaml:7-12
    "tum": (640, 480, 517.3, 516.5, 318.6, 255.3),
    # configs/Scannet/scannet.yaml:38-43 (640x480 intrinsics) scaled to the sensor's 1296x968
    "scannet": (1296, 968, 577.590698 * 1296 / 640, 578.729797 * 968 / 480,
                318.905426 * 1296 / 640, 242.683609 * 968 / 480),
    # small shapes for fast tests (W not a multiple of 16 on purpose for "ragged")
    "tiny": (64, 48, 60.0, 60.0, 31.5, 23.5),
    "ragged": (100, 70, 80.0, 82.0, 49.0, 36.0),
    "small": (320, 240, 260.0, 260.0, 159.5, 119.5),
}


class View(NamedTuple):
    image_width: int
    image_height: int
    tanfovx: float
    tanfovy: float
    viewmatrix: torch.Tensor      # (4,4) = W2C^T
    projmatrix: torch.Tensor      # (4,4)
    projmatrix_raw: torch.Tensor  # (4,4)
    campos: torch.Tensor          # (3,)


def projection_matrix2(znear, zfar, cx, cy, fx, fy, W, H) -> torch.Tensor:
    """OpenCV-intrinsics perspective matrix, same closed form as graphics_utils.py:72-93."""
    left = ((2 * cx - W) / W - 1.0) * W / 2.0
    right = ((2 * cx - W) / W + 1.0) * W / 2.0
    top = ((2 * cy - H) / H + 1.0) * H / 2.0
    bottom = ((2 * cy - H) / H - 1.0) * H / 2.0
    left, right = znear / fx * left, znear / fx * right
    top, bottom = znear / fy * top, znear / fy * bottom
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def so3_exp(theta: torch.Tensor) -> torch.Tensor:
    angle = float(theta.norm())
    W = torch.tensor([[0.0, -theta[2], theta[1]], [theta[2], 0.0, -theta[0]], [-theta[1], theta[0], 0.0]],
                     dtype=theta.dtype)
    I = torch.eye(3, dtype=theta.dtype)
    if angle < 1e-5:
        return I + W + 0.5 * (W @ W)
    return I + (math.sin(angle) / angle) * W + ((1 - math.cos(angle)) / angle ** 2) * (W @ W)


def se3_exp(tau: torch.Tensor) -> torch.Tensor:
    """Left-multiplied retraction used by the reference (thirdparty/monogs/utils/pose_utils.py:66-78)."""
    rho, theta = tau[:3], tau[3:]
    angle = float(theta.norm())
    W = torch.tensor([[0.0, -theta[2], theta[1]], [theta[2], 0.0, -theta[0]], [-theta[1], theta[0], 0.0]],
                     dtype=tau.dtype)
    I = torch.eye(3, dtype=tau.dtype)
    if angle < 1e-5:
        V = I + 0.5 * W + (1.0 / 6.0) * (W @ W)
    else:
        V = I + W * ((1.0 - math.cos(angle)) / angle ** 2) + (W @ W) * ((angle - math.sin(angle)) / angle ** 3)
    T = torch.eye(4, dtype=tau.dtype)
    T[:3, :3] = so3_exp(theta)
    T[:3, 3] = V @ rho
    return T


def window_pose(k: int) -> torch.Tensor:
    """W2C of window view k: identity for k=0, else SE3_exp(tau_k) with seed 1000+k."""
    if k == 0:
        return torch.eye(4)
    g = torch.Generator().manual_seed(1000 + k)
    u = torch.rand(6, generator=g) * 2 - 1
    tau = u * torch.tensor([0.15, 0.15, 0.15, 0.1, 0.1, 0.1])
    return se3_exp(tau.double()).float()


def make_view(camera: str, w2c: Optional[torch.Tensor] = None) -> View:
    W, H, fx, fy, cx, cy = CAMERAS[camera]
    if w2c is None:
        w2c = torch.eye(4)
    viewmatrix = w2c.float().t().contiguous()
    proj_raw = projection_matrix2(0.01, 100.0, cx, cy, fx, fy, W, H).t().contiguous()
    projmatrix = (viewmatrix.unsqueeze(0).bmm(proj_raw.unsqueeze(0))).squeeze(0).contiguous()
    campos = viewmatrix.inverse()[3, :3].contiguous()
    return View(W, H, W / (2.0 * fx), H / (2.0 * fy), viewmatrix, projmatrix, proj_raw, campos)


def make_gaussians(camera: str, P: int, seed: int = 0, sh_coeffs: int = 1, sigma_px: float = 2.0,
                   near_frac: float = 0.01) -> Dict[str, torch.Tensor]:
    """Gaussian cloud in the frame of window view 0.

    ~31 % of the centres project outside the image (rect clipping, 1.3*FoV guard band), `near_frac`
    of them sit behind/at the near plane (cull path), colours partly clamp (dc ~ U(-1.5, 1.5)).
    """
    W, H, fx, fy, cx, cy = CAMERAS[camera]
    g = torch.Generator().manual_seed(seed)
    z = torch.rand(P, generator=g) * 5.5 + 0.5
    n_near = int(P * near_frac)
    if n_near > 0:
        z[:n_near] = torch.rand(n_near, generator=g) * 1.001 - 1.0   # U(-1, 0.001)
    u = (torch.rand(P, generator=g) * 1.2 - 0.1) * W
    v = (torch.rand(P, generator=g) * 1.2 - 0.1) * H
    xyz = torch.stack([(u - cx) * z / fx, (v - cy) * z / fy, z], dim=1)
    base = torch.log(sigma_px * z.abs().clamp_min(0.05) / fx)
    scales = torch.exp(base[:, None] + 0.5 * torch.randn(P, 3, generator=g))
    q = torch.randn(P, 4, generator=g)
    rotations = q / q.norm(dim=1, keepdim=True)
    opacities = torch.sigmoid(1.5 * torch.randn(P, 1, generator=g))
    shs = 0.1 * torch.randn(P, sh_coeffs, 3, generator=g)
    shs[:, 0, :] = torch.rand(P, 3, generator=g) * 3.0 - 1.5
    colors = torch.rand(P, 3, generator=g)
    # a permutation so that the near-culled ones are not all at the front of the array
    perm = torch.randperm(P, generator=g)
    out = dict(means3D=xyz[perm], scales=scales[perm], rotations=rotations[perm], opacities=opacities[perm],
               shs=shs[perm], colors_precomp=colors[perm])
    return {k: t.contiguous().float() for k, t in out.items()}


def cov3d_from_scale_rot(scales: torch.Tensor, rotations: torch.Tensor, mod: float = 1.0) -> torch.Tensor:
    """Upper-triangular 3D covariance (P,6) for the `cov3D_precomp` input path."""
    r, x, y, z = rotations.unbind(1)
    R = torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=1).view(-1, 3, 3)
    S = torch.diag_embed((mod * scales) ** 2)
    Sig = R @ S @ R.transpose(1, 2)
    return torch.stack([Sig[:, 0, 0], Sig[:, 0, 1], Sig[:, 0, 2], Sig[:, 1, 1], Sig[:, 1, 2], Sig[:, 2, 2]],
                       dim=1).contiguous()


def upstream_grads(camera: str, seed: int = 7):
    """Fixed random weights Wc (3,H,W), Wd (1,H,W): loss = (color*Wc).sum() + (depth*Wd).sum()."""
    W, H = CAMERAS[camera][:2]
    g = torch.Generator().manual_seed(seed)
    Wc = torch.rand(3, H, W, generator=g) * 2 - 1
    Wd = torch.rand(1, H, W, generator=g) * 2 - 1
    return Wc, Wd
