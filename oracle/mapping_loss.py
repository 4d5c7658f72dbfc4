"""TEST INFRASTRUCTURE — CPU oracle of the mapping loss that follows the rasterizer on every view.

Plain-PyTorch restatement of `get_loss_mapping` / `get_loss_mapping_rgbd` with ssim_loss = False
(reference thirdparty/monogs/utils/slam_utils.py:71-105). Pinned against the reference's own function, imported in
the build container, by tests/golden/make_golden_loss.py -> tests/golden/mapping_loss.pt. Used only by tests.
"""
import math

import torch


def ssim(img1, img2, window_size=11):
    """thirdparty/gaussian_splatting/utils/loss_utils.py:37-99: 11x11 Gaussian window (sigma 1.5), zero padding, mean."""
    import torch.nn.functional as F
    channel = img1.size(-3)
    g = torch.tensor([math.exp(-((x - window_size // 2) ** 2) / float(2 * 1.5 ** 2)) for x in range(window_size)])
    g = (g / g.sum()).unsqueeze(1)
    window = (g.mm(g.t()).float().unsqueeze(0).unsqueeze(0)).expand(channel, 1, window_size, window_size).contiguous()
    window = window.to(img1)
    pad = window_size // 2
    mu1 = F.conv2d(img1, window, padding=pad, groups=channel)
    mu2 = F.conv2d(img2, window, padding=pad, groups=channel)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    sigma1_sq = F.conv2d(img1 * img1, window, padding=pad, groups=channel) - mu1_sq
    sigma2_sq = F.conv2d(img2 * img2, window, padding=pad, groups=channel) - mu2_sq
    sigma12 = F.conv2d(img1 * img2, window, padding=pad, groups=channel) - mu1_mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim_map = ((2 * mu1_mu2 + C1) * (2 * sigma12 + C2)) / ((mu1_sq + mu2_sq + C1) * (sigma1_sq + sigma2_sq + C2))
    return ssim_map.mean()


def mapping_loss(image, depth, exposure_a, exposure_b, gt_image, gt_depth, alpha=0.95, rgb_boundary_threshold=0.01,
                 initialization=False, lambda_dssim=None):
    # slam_utils.py:72-76
    image_ab = image if initialization else torch.exp(exposure_a) * image + exposure_b
    # slam_utils.py:83-105 (ssim_loss False)
    _, h, w = gt_image.shape
    rgb_pixel_mask = (gt_image.sum(dim=0) > rgb_boundary_threshold).view(1, h, w)
    l1_rgb = torch.abs(image_ab * rgb_pixel_mask - gt_image * rgb_pixel_mask)
    if lambda_dssim is not None:    # slam_utils.py:91-98 (ssim_loss True): (1 - h) * l1_rgb + h * (1 - ssim), broadcast
        l1_rgb = (1.0 - lambda_dssim) * l1_rgb + lambda_dssim * (1.0 - ssim(image_ab, gt_image))
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
