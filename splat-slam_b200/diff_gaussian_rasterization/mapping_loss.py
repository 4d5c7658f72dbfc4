"""Fused mapping loss (forward + backward in one CUDA pass) for the views a mapping step renders.

Mirrors `get_loss_mapping(config, image, depth, viewpoint, opacity, initialization=False)` of the reference
(thirdparty/monogs/utils/slam_utils.py:71-105; the `ssim_loss: False` path that configs/splat_slam.yaml:36 selects is
the fused kernel, the SSIM term of `ssim_loss: True` is added with torch ops):
same arguments, same value, same gradients w.r.t. `image`, `depth`, `viewpoint.exposure_a`, `viewpoint.exposure_b`,
but one kernel instead of ~10 elementwise/reduction kernels, and the ground-truth depth is uploaded once per
`viewpoint.depth` array instead of on every call (the reference does `torch.from_numpy(viewpoint.depth).to(device)`
each time, slam_utils.py:87-89; see `_device_depth` for when it is uploaded again). No CPU path.
"""
from __future__ import annotations

import ctypes
from ctypes import c_float, c_int, c_void_p

import torch

from . import _cabi

_fn = None
_workspaces = {}       # (device index, stream handle) -> zeroed ticket + per-block partials, reused call after call


def _entry():
    global _fn
    if _fn is None:
        L = _cabi.lib()
        f = L.splat_raster_mapping_loss
        P = c_void_p
        f.restype = c_int
        f.argtypes = [c_int, c_int, P, P, P, P, P, P, c_float, c_float, c_int, P, P, P, P, P]
        L.splat_raster_mapping_loss_blocks.restype = c_int
        L.splat_raster_mapping_loss_blocks.argtypes = [c_int, c_int]
        _fn = (f, L.splat_raster_mapping_loss_blocks)
    return _fn


class _FusedMappingLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, depth, exposure_a, exposure_b, gt_image, gt_depth, alpha, rgb_boundary_threshold,
                initialization):
        fn, nblocks = _entry()
        image = _cabi.require_cuda("image", image)
        depth = _cabi.require_cuda("depth", depth)
        gt_image = _cabi.require_cuda("gt_image", gt_image)
        gt_depth = _cabi.require_cuda("gt_depth", gt_depth)
        if image.dim() != 3 or image.size(0) != 3:
            raise RuntimeError("image must have shape (3, H, W)")
        H, W = image.size(1), image.size(2)
        if depth.numel() != H * W or gt_depth.numel() != H * W or gt_image.shape != image.shape:
            raise RuntimeError("depth / gt_image / gt_depth do not match the image shape")
        dev = image.device
        ea = eb = None
        if not initialization:
            ea = _cabi.require_cuda("exposure_a", exposure_a.reshape(-1))
            eb = _cabi.require_cuda("exposure_b", exposure_b.reshape(-1))
        with torch.cuda.device(dev):
            stream = _cabi.stream_ptr(dev)
            words = 4 + 4 * nblocks(W, H)
            ws = _workspaces.get((dev.index, stream))
            if ws is None or ws.numel() < words:
                # zeroed once; the kernel hands the ticket word back as 0, and calls on one stream are ordered
                ws = _workspaces[(dev.index, stream)] = torch.zeros(words, dtype=torch.float32, device=dev)
            out = torch.empty(5, dtype=torch.float32, device=dev)
            dimg = torch.empty_like(image)
            ddepth = torch.empty_like(depth)
            _cabi.check(fn(W, H, image.data_ptr(), depth.data_ptr(), gt_image.data_ptr(), gt_depth.data_ptr(),
                           _cabi.ptr(ea), _cabi.ptr(eb), float(alpha), float(rgb_boundary_threshold),
                           int(bool(initialization)), ws.data_ptr(), out.data_ptr(), dimg.data_ptr(),
                           ddepth.data_ptr(), stream), "mapping_loss")
            loss = out[0]
        ctx.save_for_backward(dimg, ddepth, out)
        ctx.initialization = bool(initialization)
        ctx.exposure_shapes = (None if exposure_a is None else exposure_a.shape,
                               None if exposure_b is None else exposure_b.shape)
        return loss

    @staticmethod
    def backward(ctx, grad_loss):
        dimg, ddepth, sums = ctx.saved_tensors
        ga = gb = None
        if not ctx.initialization:
            ga = (grad_loss * sums[3]).reshape(ctx.exposure_shapes[0])
            gb = (grad_loss * sums[4]).reshape(ctx.exposure_shapes[1])
        return grad_loss * dimg, grad_loss * ddepth, ga, gb, None, None, None, None, None


def fused_mapping_loss(image, depth, exposure_a, exposure_b, gt_image, gt_depth, alpha=0.95,
                       rgb_boundary_threshold=0.01, initialization=False):
    """loss = alpha * mean|image_ab*m - gt_image*m| + (1-alpha) * mean|depth*md - gt_depth*md| (see module docstring)."""
    return _FusedMappingLoss.apply(image, depth, exposure_a, exposure_b, gt_image, gt_depth, alpha,
                                   rgb_boundary_threshold, initialization)


def _device_depth(viewpoint, device):
    """Ground-truth depth of a viewpoint on the device, shape (1, H, W).

    The reference uploads `viewpoint.depth` (a numpy array) on every call (slam_utils.py:87-89). Here the upload is
    kept on the viewpoint and reused for as long as `viewpoint.depth` is THE SAME OBJECT: Splat-SLAM replaces the array
    when it re-aligns a keyframe's depth (src/mapper.py:633-639 after bundle adjustment, :1040-1046 in final_refine), and
    a replaced array is uploaded again. The cache holds a reference to the source array, so its identity cannot be
    recycled. Code that edits the array in place after the first loss call must call `invalidate_depth_cache`."""
    d = viewpoint.depth
    cached = getattr(viewpoint, "_depth_cuda", None)
    if cached is not None and cached[0] is d and cached[1].device == device:
        return cached[1]
    t = (torch.from_numpy(d) if not torch.is_tensor(d) else d).to(dtype=torch.float32, device=device)[None]
    try:
        viewpoint._depth_cuda = (d, t)
    except Exception:
        pass
    return t


def invalidate_depth_cache(viewpoint):
    """Forget the uploaded ground-truth depth of `viewpoint` (needed only after an in-place edit of `viewpoint.depth`)."""
    if getattr(viewpoint, "_depth_cuda", None) is not None:
        viewpoint._depth_cuda = None


class attach_loss:
    """Evaluate the mapping loss of a view INSIDE its render instead of after it.

        with attach_loss(config, viewpoint, initialization=False) as fused:
            render_pkg = render(viewpoint, gaussians, pipe, background)     # any code that ends up in GaussianRasterizer
        loss_mapping += fused.loss          # == get_loss_mapping(config, image, depth, viewpoint, opacity, ...)

    The next `GaussianRasterizer` call of this thread computes the loss in the epilogue of its forward kernel, from the
    pixel values still in registers, and its backward uses the loss gradients directly: one launch, one pass over the
    rendered images and the Python glue of a second autograd node less per view. `image` and `depth` of the render are
    still returned and still differentiable, so other terms can use them. Same value and gradients as
    `get_loss_mapping` (ssim_loss: False path, thirdparty/monogs/utils/slam_utils.py:71-105)."""

    def __init__(self, config, viewpoint, initialization=False):
        training = config["Training"]
        if training.get("ssim_loss", False):
            raise NotImplementedError("the fused loss covers the ssim_loss: False path only (configs/splat_slam.yaml:36)")
        self._args = (training["alpha"] if "alpha" in training else 0.95, training["rgb_boundary_threshold"])
        self._viewpoint, self._initialization = viewpoint, initialization
        self._request = None

    def __enter__(self):
        from . import LossRequest, _pending_loss
        vp = self._viewpoint
        gt_image = vp.original_image if vp.original_image.is_cuda else vp.original_image.cuda()
        self._request = LossRequest(gt_image, _device_depth(vp, gt_image.device), vp.exposure_a, vp.exposure_b,
                                    self._args[0], self._args[1], self._initialization)
        self._previous = getattr(_pending_loss, "request", None)
        _pending_loss.request = self._request
        return self

    def __exit__(self, *exc):
        from . import _pending_loss
        _pending_loss.request = self._previous
        return False

    @property
    def loss(self):
        if self._request is None or self._request.loss is None:
            raise RuntimeError("attach_loss: no render ran inside the with-block")
        return self._request.loss


def _ssim(img1, img2, window_size=11):
    """SSIM exactly as thirdparty/gaussian_splatting/utils/loss_utils.py:37-99 (11x11 Gaussian window, sigma 1.5, zero
    padding, mean over the map), on torch ops: this branch is off in the shipped configuration."""
    import math

    import torch.nn.functional as F
    channel = img1.size(-3)
    g = torch.tensor([math.exp(-((x - window_size // 2) ** 2) / float(2 * 1.5 ** 2)) for x in range(window_size)])
    g = (g / g.sum()).unsqueeze(1)
    window = g.mm(g.t()).float().unsqueeze(0).unsqueeze(0).expand(channel, 1, window_size, window_size).contiguous()
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


def get_loss_mapping(config, image, depth, viewpoint, opacity, initialization=False):
    """Drop-in for thirdparty/monogs/utils/slam_utils.py:71-78 (+ get_loss_mapping_rgbd :81-105).

    `ssim_loss: False` (configs/splat_slam.yaml:36) is one fused kernel. With `ssim_loss: True` the reference's
    loss is alpha * ((1-h) * mean l1_rgb + h * (1 - ssim(image_ab, gt))) + (1-alpha) * mean l1_depth with
    h = config["opt_params"]["lambda_dssim"] (slam_utils.py:91-98): the two L1 terms still come from the fused kernel
    (called with the equivalent mixing weight and rescaled), the SSIM term is added with torch ops."""
    training = config["Training"]
    alpha = training["alpha"] if "alpha" in training else 0.95
    thr = training["rgb_boundary_threshold"]
    gt_image = viewpoint.original_image
    if not gt_image.is_cuda:
        gt_image = gt_image.cuda()
    gt_depth = _device_depth(viewpoint, image.device)
    if not training.get("ssim_loss", False):
        return fused_mapping_loss(image, depth, viewpoint.exposure_a, viewpoint.exposure_b, gt_image, gt_depth, alpha,
                                  thr, initialization)
    h = config["opt_params"]["lambda_dssim"]
    # alpha (1-h) A + (1-alpha) D  =  s * (alpha' A + (1-alpha') D)
    s = alpha * (1.0 - h) + (1.0 - alpha)
    l1 = fused_mapping_loss(image, depth, viewpoint.exposure_a, viewpoint.exposure_b, gt_image, gt_depth,
                            alpha * (1.0 - h) / s, thr, initialization)
    image_ab = image if initialization else torch.exp(viewpoint.exposure_a) * image + viewpoint.exposure_b
    return s * l1 + alpha * h * (1.0 - _ssim(image_ab, gt_image))
