"""TEST INFRASTRUCTURE — CPU oracle of the mapping loss that follows the rasterizer on every view.

Plain-PyTorch restatement of `get_loss_mapping` / `get_loss_mapping_rgbd` with ssim_loss = False
(reference thirdparty/monogs/utils/slam_utils.py:71-105). Pinned against the reference's own function, imported in
the build container, by tests/golden/make_golden_loss.py -> tests/golden/mapping_loss.pt. Used only by tests.
"""
import torch


def mapping_loss(image, depth, exposure_a, exposure_b, gt_image, gt_depth, alpha=0.95, rgb_boundary_threshold=0.01,
                 initialization=False):
    # slam_utils.py:72-76
    image_ab = image if initialization else torch.exp(exposure_a) * image + exposure_b
    # slam_utils.py:83-105 (ssim_loss False)
    _, h, w = gt_image.shape
    rgb_pixel_mask = (gt_image.sum(dim=0) > rgb_boundary_threshold).view(1, h, w)
    l1_rgb = torch.abs(image_ab * rgb_pixel_mask - gt_image * rgb_pixel_mask)
    depth_pixel_mask = (gt_depth > 0.01).view(*depth.shape)
    l1_depth = torch.abs(depth * depth_pixel_mask - gt_depth.view(*depth.shape) * depth_pixel_mask)
    return alpha * l1_rgb.mean() + (1 - alpha) * l1_depth.mean()


def synthetic_loss_case(H=48, W=64, seed=0):
    g = torch.Generator().manual_seed(seed)
    image = torch.rand(3, H, W, generator=g)
    depth = torch.rand(1, H, W, generator=g) * 4
    gt_image = torch.rand(3, H, W, generator=g)
    gt_image[:, : H // 6] = 0.0                      # black boundary -> rgb mask off
    gt_image[:, H // 2, W // 2] = image[:, H // 2, W // 2]  # exact ties -> sign 0
    gt_depth = torch.rand(1, H, W, generator=g) * 4
    gt_depth[:, :, : W // 5] = 0.0                   # invalid depth -> depth mask off
    a = torch.tensor([0.13])
    b = torch.tensor([-0.05])
    return dict(image=image, depth=depth, gt_image=gt_image, gt_depth=gt_depth, exposure_a=a, exposure_b=b)
