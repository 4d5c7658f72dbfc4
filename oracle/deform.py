"""TEST INFRASTRUCTURE — CPU oracle of the map deformation after a keyframe pose / depth update.

Plain-PyTorch restatement of Mapper.update_mapping_points (src/mapper.py:154-255) with its helpers
rotation_matrix_to_quaternion / quaternion_multiply (thirdparty/gaussian_splatting/utils/general_utils.py:138-176),
operating on plain tensors instead of the GaussianModel. Pinned by tests/golden/make_golden_deform.py, which executes
the reference's own method (its source is taken from the mounted reference at run time, bound to a stand-in Mapper
holding a real reference GaussianModel) on the same seeded inputs. Used only by tests.
"""
import torch


def rotation_matrix_to_quaternion(R):                      # general_utils.py:138-162
    q = torch.zeros((R.size(0), 4), dtype=R.dtype)
    zero = torch.tensor(0.0, dtype=R.dtype)
    q[:, 0] = torch.sqrt(torch.max(zero, 1 + R[:, 0, 0] + R[:, 1, 1] + R[:, 2, 2])) / 2
    q[:, 1] = torch.sqrt(torch.max(zero, 1 + R[:, 0, 0] - R[:, 1, 1] - R[:, 2, 2])) / 2
    q[:, 2] = torch.sqrt(torch.max(zero, 1 - R[:, 0, 0] + R[:, 1, 1] - R[:, 2, 2])) / 2
    q[:, 3] = torch.sqrt(torch.max(zero, 1 - R[:, 0, 0] - R[:, 1, 1] + R[:, 2, 2])) / 2
    q[:, 1] *= torch.sign(q[:, 1] * (R[:, 2, 1] - R[:, 1, 2]))
    q[:, 2] *= torch.sign(q[:, 2] * (R[:, 0, 2] - R[:, 2, 0]))
    q[:, 3] *= torch.sign(q[:, 3] * (R[:, 1, 0] - R[:, 0, 1]))
    return q


def quaternion_multiply(q1, q2):                           # general_utils.py:164-176
    w1, x1, y1, z1 = q1[..., 0], q1[..., 1], q1[..., 2], q1[..., 3]
    w2, x2, y2, z2 = q2[..., 0], q2[..., 1], q2[..., 2], q2[..., 3]
    return torch.stack((w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                        w1 * y2 + y1 * w2 + z1 * x2 - x1 * z2, w1 * z2 + z1 * w2 + x1 * y2 - y1 * x2), dim=-1)


def update_mapping_points(xyz, rotation, scaling, kf_ids, frame_idx, w2c, w2c_old, depth, depth_old, intrinsics,
                          method=None):
    """Returns (xyz', rotation', scaling') or None when no Gaussian is anchored to frame_idx (mapper.py:159-160).
    `rotation` / `scaling` are the raw parameters (_rotation, _scaling); rotation' is normalised for every Gaussian."""
    frame_mask = kf_ids == frame_idx
    if frame_mask.sum() == 0:
        return None
    xyz, scaling = xyz.clone(), scaling.clone()
    rots = torch.nn.functional.normalize(rotation)                                   # get_rotation
    transformation = torch.linalg.inv(torch.linalg.inv(w2c_old) @ w2c)
    ones = torch.ones(int(frame_mask.sum()), 1)
    if method == "rigid":
        pts4 = torch.cat((xyz[frame_mask], ones), dim=1)
        xyz[frame_mask] = (transformation @ pts4.T).T[:, :3]
    else:
        means = xyz[frame_mask]
        pts4 = torch.cat((means, ones), dim=1)
        pix = (intrinsics @ (w2c_old @ pts4.T)[:3, :]).T
        pix[:, 0] /= pix[:, 2]
        pix[:, 1] /= pix[:, 2]
        pix = pix[:, :2].long()
        height, width = depth.shape
        pix[:, 0] = torch.clamp(pix[:, 0], min=0, max=width - 1)
        pix[:, 1] = torch.clamp(pix[:, 1], min=0, max=height - 1)
        d_new = depth[pix[:, 1], pix[:, 0]]
        d_old = depth_old[pix[:, 1], pix[:, 0]]
        means_cam = (w2c_old @ pts4.T).T[:, :3]
        rescale = (1 + 1 / (means_cam[:, 2]) * (d_new - d_old)).unsqueeze(-1)
        rescale[torch.logical_or(d_new == 0, d_old == 0)] = 1
        rescale[rescale <= 0.0] = 1
        means_cam = rescale.repeat(1, 3) * means_cam
        pts4 = torch.cat((means_cam, ones), dim=1)
        means = (torch.linalg.inv(w2c_old) @ pts4.T).T[:, :3]
        pts4 = torch.cat((means, ones), dim=1)
        xyz[frame_mask] = (transformation @ pts4.T).T[:, :3]
        scaling[frame_mask] = scaling[frame_mask] + torch.log(rescale)
    q = rotation_matrix_to_quaternion(transformation.unsqueeze(0))
    rots[frame_mask] = quaternion_multiply(q.expand_as(rots[frame_mask]), rots[frame_mask])
    return xyz, rots, scaling


def _pose(g, scale):
    """A rigid world-to-camera matrix near identity."""
    w = torch.randn(3, generator=g) * scale
    th = float(w.norm())
    Kx = torch.tensor([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]]) / max(th, 1e-9)
    R = torch.eye(3) + torch.sin(torch.tensor(th)) * Kx + (1 - torch.cos(torch.tensor(th))) * (Kx @ Kx)
    M = torch.eye(4)
    M[:3, :3] = R
    M[:3, 3] = torch.randn(3, generator=g) * scale
    return M


def synthetic_deform_case(P=2000, n_kf=4, H=48, W=64, seed=0):
    g = torch.Generator().manual_seed(seed)
    fx = fy = 0.9 * W
    K = torch.tensor([[fx, 0, W / 2], [0, fy, H / 2], [0, 0, 1.0]])
    kf_ids = torch.randint(0, n_kf + 1, (P,), generator=g).int() * 5           # ids 0,5,10,...; one id may stay unused
    w2c_old = [_pose(g, 0.2) for _ in range(n_kf + 1)]
    w2c_new = [_pose(g, 0.02) @ m for m in w2c_old]
    depth_old = [1.0 + 3.0 * torch.rand(H, W, generator=g) for _ in range(n_kf + 1)]
    depth_new = [d * (1.0 + 0.1 * torch.randn(H, W, generator=g)) for d in depth_old]
    for d in depth_new + depth_old:
        d[torch.rand(H, W, generator=g) < 0.05] = 0.0                          # invalid depth -> rigid for that point
    depth_new[0][:4] = -5.0                                                     # drives rescale <= 0 -> 1
    # points: back-project random pixels of their keyframe at the old depth (+ some far outside the frustum)
    xyz = torch.zeros(P, 3)
    for k in range(n_kf + 1):
        m = kf_ids == 5 * k
        n = int(m.sum())
        u = torch.rand(n, generator=g) * (W * 1.3) - 0.15 * W
        v = torch.rand(n, generator=g) * (H * 1.3) - 0.15 * H
        z = 1.0 + 3.0 * torch.rand(n, generator=g)
        cam = torch.stack([(u - W / 2) / fx * z, (v - H / 2) / fy * z, z, torch.ones(n)], 1)
        xyz[m] = (torch.linalg.inv(w2c_old[k]) @ cam.T).T[:, :3]
    rotation = torch.randn(P, 4, generator=g) * (0.5 + torch.rand(P, 1, generator=g))
    scaling = torch.log(0.01 + 0.05 * torch.rand(P, 3, generator=g))
    return dict(xyz=xyz, rotation=rotation, scaling=scaling, kf_ids=kf_ids, w2c_old=w2c_old, w2c_new=w2c_new,
                depth_old=depth_old, depth_new=depth_new, intrinsics=K, H=H, W=W, n_kf=n_kf)
