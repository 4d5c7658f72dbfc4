"""Drop-in `diff_gaussian_rasterization` for Splat-SLAM, backed by hand-written sm_100a CUDA.

Public surface and semantics mirror the reference package
(thirdparty/diff-gaussian-rasterization-w-pose/diff_gaussian_rasterization/__init__.py, "DGR" below):

  * `GaussianRasterizationSettings`  — the 13-field NamedTuple (DGR __init__.py:173-186)
  * `GaussianRasterizer(raster_settings)(means3D, means2D, opacities, shs, colors_precomp, scales, rotations,
     cov3D_precomp, theta, rho)` -> `(color, radii, depth, opacity, n_touched)` (DGR __init__.py:188-244)
  * `GaussianRasterizer.markVisible(positions)` (DGR __init__.py:193-202)
  * `rasterize_gaussians(...)` / `_RasterizeGaussians` (DGR __init__.py:21-171)

Backward returns gradients for every Gaussian attribute plus the camera pose (`theta`, `rho`).  `means2D`,
`theta` and `rho` are pure gradient sinks exactly as in the reference (never read by a kernel).

Host code stays Python/PyTorch: it allocates tensors, selects device + current stream and calls the C ABI of
`libsplat_raster.so` (include/splat_raster.h) through ctypes.  No CPU path exists.
"""
from __future__ import annotations

import threading
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _cabi

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians"]


def cpu_deep_copy_tuple(input_tuple):
    copied_tensors = [item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple]
    return tuple(copied_tensors)


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    projmatrix_raw: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, theta, rho,
                        raster_settings):
    req = getattr(_pending_loss, "request", None)
    if req is not None and req.loss is None:
        # a mapping loss was attached to this render (mapping_loss.attach_loss): it is evaluated in the epilogue of
        # the forward renderer and comes back as a sixth, differentiable output that is handed to the requester
        color, radii, depth, opacity, n_touched, loss = _RasterizeGaussiansLoss.apply(
            means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, theta, rho,
            req.exposure_a, req.exposure_b, raster_settings, req)
        req.loss = loss
        return color, radii, depth, opacity, n_touched
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, theta, rho, raster_settings)


_pending_loss = threading.local()      # .request: the LossRequest the next render of this thread should evaluate


class LossRequest:
    """What `mapping_loss.attach_loss` leaves for the next render of the calling thread."""

    def __init__(self, gt_image, gt_depth, exposure_a, exposure_b, alpha, rgb_boundary_threshold, initialization):
        self.gt_image, self.gt_depth = gt_image, gt_depth
        self.exposure_a, self.exposure_b = exposure_a, exposure_b
        self.alpha, self.rgb_boundary_threshold = float(alpha), float(rgb_boundary_threshold)
        self.initialization = bool(initialization)
        self.loss = None            # set by the render: 0-dim tensor, differentiable


class _PinnedInt:
    """One pinned int32 per device and host thread for the num_rendered hand-off (the reference's blocking 4-byte
    cudaMemcpy, DGR/cuda_rasterizer/rasterizer_impl.cu:284). A thread reads its slot before it enqueues the next
    forward, so views issued on several streams by one thread can share it; threads must not."""
    _local = threading.local()

    @classmethod
    def get(cls, device):
        slots = getattr(cls._local, "slots", None)
        if slots is None:
            slots = cls._local.slots = {}
        key = (device.type, device.index)
        t = slots.get(key)
        if t is None:
            t = slots[key] = torch.zeros(1, dtype=torch.int32).pin_memory()
        return t


_loss_workspaces = {}     # (device index, stream handle) -> zeroed ticket + per-tile partials, reused call after call


def _forward_impl(S, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, loss=None):
    """Equivalent of `_C.rasterize_gaussians` (DGR/rasterize_points.cu:35-122): returns
    (num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, depth, opacity, n_touched).
    With `loss` (a LossRequest) the mapping loss is evaluated in the render epilogue and the tuple grows by
    (loss sums[5], dL/dimage, dL/ddepth)."""
    if means3D.ndimension() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    L = _cabi.lib()
    means3D = _cabi.require_cuda("means3D", means3D)
    dev = means3D.device
    P, H, W = means3D.size(0), int(S.image_height), int(S.image_width)
    f32 = dict(dtype=torch.float32, device=dev)
    i32 = dict(dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        stream = _cabi.stream_ptr(dev)
        out_color = torch.empty((3, H, W), **f32)
        out_depth = torch.empty((1, H, W), **f32)
        out_opacity = torch.empty((1, H, W), **f32)
        radii = torch.empty((P,), **i32)
        n_touched = torch.empty((P,), **i32)
        geom = torch.empty((L.splat_raster_geom_bytes(P),), dtype=torch.uint8, device=dev)
        image = torch.empty((L.splat_raster_image_bytes(W, H),), dtype=torch.uint8, device=dev)
        bg = _cabi.require_cuda("bg", S.bg)
        M = 0
        if sh.numel() != 0:
            sh = _cabi.require_cuda("sh", sh)
            M = sh.size(1)
        if colors_precomp.numel() != 0:
            colors_precomp = _cabi.require_cuda("colors_precomp", colors_precomp)
        if scales.numel() != 0:
            scales = _cabi.require_cuda("scales", scales)
        if rotations.numel() != 0:
            rotations = _cabi.require_cuda("rotations", rotations)
        if cov3Ds_precomp.numel() != 0:
            cov3Ds_precomp = _cabi.require_cuda("cov3Ds_precomp", cov3Ds_precomp)
        opacities = _cabi.require_cuda("opacities", opacities) if P else opacities
        view = _cabi.require_cuda("viewmatrix", S.viewmatrix)
        proj = _cabi.require_cuda("projmatrix", S.projmatrix)
        campos = _cabi.require_cuda("campos", S.campos)
        host_R = _PinnedInt.get(dev)
        rendered = 0
        # debug: wait and check after EVERY stage inside the library (the reference's CHECK_CUDA(..., debug),
        # DGR/cuda_rasterizer/auxiliary.h:166-173); a failure names its stage in the exception
        was_debug = _cabi.set_debug(True, _debug_fail_stage) if S.debug else None
        try:
            if P != 0:
                _cabi.check(L.splat_raster_forward_preprocess(
                    P, int(S.sh_degree), M, W, H, _cabi.ptr(means3D), _cabi.ptr(scales), float(S.scale_modifier),
                    _cabi.ptr(rotations), _cabi.ptr(opacities), _cabi.ptr(sh), _cabi.ptr(cov3Ds_precomp),
                    _cabi.ptr(colors_precomp), _cabi.ptr(view), _cabi.ptr(proj), _cabi.ptr(campos), float(S.tanfovx),
                    float(S.tanfovy), int(bool(S.prefiltered)), radii.data_ptr(), n_touched.data_ptr(),
                    geom.data_ptr(), host_R.data_ptr(), stream), "forward_preprocess")
                _cabi.check(L.splat_raster_forward_wait(), "forward_wait")   # returns while the depth sort still runs
                rendered = int(host_R.item())
            binning = torch.empty((L.splat_raster_binning_bytes(rendered),), dtype=torch.uint8, device=dev)
            common = (P, rendered, W, H, _cabi.ptr(bg), _cabi.ptr(colors_precomp), _cabi.ptr(radii), geom.data_ptr(),
                      binning.data_ptr(), image.data_ptr(), out_color.data_ptr(), out_depth.data_ptr(),
                      out_opacity.data_ptr(), _cabi.ptr(n_touched))
            extra = ()
            if loss is None:
                _cabi.check(L.splat_raster_forward_render(*common, stream), "forward_render")
            else:
                gt_image = _cabi.require_cuda("gt_image", loss.gt_image)
                gt_depth = _cabi.require_cuda("gt_depth", loss.gt_depth)
                if gt_image.numel() != 3 * H * W or gt_depth.numel() != H * W:
                    raise RuntimeError("gt_image / gt_depth do not match the rendered image shape")
                ea = eb = None
                if not loss.initialization:
                    ea = _cabi.require_cuda("exposure_a", loss.exposure_a.detach().reshape(-1))
                    eb = _cabi.require_cuda("exposure_b", loss.exposure_b.detach().reshape(-1))
                words = 4 + 4 * L.splat_raster_loss_partials(W, H)
                ws = _loss_workspaces.get((dev.index, stream))
                if ws is None or ws.numel() < words:
                    # zeroed once; the kernel hands the ticket word back as 0, and calls on one stream are ordered
                    ws = _loss_workspaces[(dev.index, stream)] = torch.zeros(words, **f32)
                sums = torch.empty(5, **f32)
                dimg = torch.empty((3, H, W), **f32)
                ddepth = torch.empty((1, H, W), **f32)
                desc = _cabi._LossDesc(gt_image.data_ptr(), gt_depth.data_ptr(), _cabi.ptr(ea), _cabi.ptr(eb),
                                       loss.alpha, loss.rgb_boundary_threshold, int(loss.initialization),
                                       ws.data_ptr(), sums.data_ptr(), dimg.data_ptr(), ddepth.data_ptr())
                import ctypes
                _cabi.check(L.splat_raster_forward_render_loss(*common, ctypes.byref(desc), stream),
                            "forward_render_loss")
                extra = (sums, dimg, ddepth)
        finally:
            if was_debug is not None:
                _cabi.set_debug(was_debug)
    return (rendered, out_color, radii, geom, binning, image, out_depth, out_opacity, n_touched) + extra


_debug_fail_stage = -1      # tests set this to a stage index to exercise the error path of debug=True


ALL_GRADS = dict(means2D=True, colors=True, opacity=True, means3D=True, cov3D=True, sh=True, scale_rot=True)


def _backward_impl(S, means3D, radii, colors_precomp, scales, rotations, cov3Ds_precomp, grad_out_color, grad_out_depth,
                   sh, geom, num_rendered, binning, image, needs=None):
    """Equivalent of `_C.rasterize_gaussians_backward` (DGR/rasterize_points.cu:124-211) followed by the pose
    reduction of DGR __init__.py:152-154. Returns (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D,
    dL_dsh, dL_dscales, dL_drotations, grad_tau[6]).

    `needs` (keys of ALL_GRADS) says which per-Gaussian gradients the caller wants; the others are neither allocated
    nor computed and come back as None (e.g. pose-only tracking needs none of them). The reference always computes
    everything."""
    L = _cabi.lib()
    dev = means3D.device
    P = means3D.size(0)
    H, W = int(S.image_height), int(S.image_width)
    M = sh.size(1) if sh.numel() != 0 else 0
    f32 = dict(dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        stream = _cabi.stream_ptr(dev)
        needs = ALL_GRADS if needs is None else needs
        has_scale = scales.numel() != 0
        # the SH direction term also feeds the pose gradient, so dL_dsh is always computed when degree > 0
        want_sh = M > 0 and (needs["sh"] or int(S.sh_degree) > 0)
        dL_dmeans3D = torch.empty((P, 3), **f32) if needs["means3D"] else None
        dL_dmeans2D = torch.empty((P, 3), **f32) if needs["means2D"] else None
        dL_dcolors = torch.empty((P, 3), **f32) if needs["colors"] else None
        dL_dopacity = torch.empty((P, 1), **f32) if needs["opacity"] else None
        dL_dcov3D = torch.empty((P, 6), **f32) if needs["cov3D"] else None
        dL_dsh = torch.empty((P, M, 3), **f32) if want_sh else None
        if needs["scale_rot"]:
            dL_dscales = torch.empty((P, 3), **f32) if has_scale else torch.zeros((P, 3), **f32)
            dL_drotations = torch.empty((P, 4), **f32) if has_scale else torch.zeros((P, 4), **f32)
        else:
            dL_dscales = dL_drotations = None
        nblk = (P + 255) // 256
        tau_partial = torch.empty((nblk, 6), **f32)
        scratch = torch.empty((P * 12,), **f32)
        if P != 0:
            gc = _cabi.require_cuda("grad_out_color", grad_out_color)
            gd = _cabi.require_cuda("grad_out_depth", grad_out_depth)
            was_debug = _cabi.set_debug(True, _debug_fail_stage) if S.debug else None
            try:
                _cabi.check(L.splat_raster_backward(
                    P, int(S.sh_degree), M, int(num_rendered), W, H, _cabi.ptr(_cabi.require_cuda("bg", S.bg)),
                    _cabi.ptr(means3D), _cabi.ptr(sh), _cabi.ptr(colors_precomp), _cabi.ptr(scales),
                    float(S.scale_modifier), _cabi.ptr(rotations), _cabi.ptr(cov3Ds_precomp),
                    _cabi.ptr(_cabi.require_cuda("viewmatrix", S.viewmatrix)),
                    _cabi.ptr(_cabi.require_cuda("projmatrix", S.projmatrix)),
                    _cabi.ptr(_cabi.require_cuda("projmatrix_raw", S.projmatrix_raw)),
                    _cabi.ptr(_cabi.require_cuda("campos", S.campos)), float(S.tanfovx), float(S.tanfovy),
                    radii.data_ptr(), geom.data_ptr(), binning.data_ptr(), image.data_ptr(), gc.data_ptr(), gd.data_ptr(),
                    scratch.data_ptr(), _cabi.ptr(dL_dmeans2D), _cabi.ptr(dL_dcolors), _cabi.ptr(dL_dopacity),
                    _cabi.ptr(dL_dmeans3D), _cabi.ptr(dL_dcov3D), _cabi.ptr(dL_dsh),
                    _cabi.ptr(dL_dscales) if has_scale else None, _cabi.ptr(dL_drotations) if has_scale else None,
                    tau_partial.data_ptr(), stream), "backward")
            finally:
                if was_debug is not None:
                    _cabi.set_debug(was_debug)
            grad_tau = tau_partial.sum(dim=0)
        else:
            grad_tau = torch.zeros(6, **f32)
    return dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations, grad_tau


def _pose_grads(grad_tau, shapes):
    """(grad_rho, grad_theta) from tau = [rho, theta]. The reference hands back (1, 3) tensors and lets autograd sum them
    down to the parameters' shape (DGR __init__.py:152-171); a 3-element parameter gets its own shape directly, which is
    the same value without the two reduction kernels."""
    theta_shape, rho_shape = shapes
    g_rho, g_theta = grad_tau[:3], grad_tau[3:]
    g_rho = g_rho.view(rho_shape) if len(rho_shape) and rho_shape.count(1) == len(rho_shape) - 1 and 3 in rho_shape \
        else g_rho.view(1, -1)
    g_theta = g_theta.view(theta_shape) if len(theta_shape) and theta_shape.count(1) == len(theta_shape) - 1 \
        and 3 in theta_shape else g_theta.view(1, -1)
    return g_rho, g_theta


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, theta, rho,
                raster_settings):
        # Dense fp32 copies are made HERE (the reference calls .contiguous().data<float>() in both directions,
        # DGR/rasterize_points.cu) and the same tensors are saved for backward: an expanded / sliced / fp16 input is
        # then read with the right strides by both passes.
        def dense(name, t):
            return _cabi.require_cuda(name, t) if t.numel() != 0 else t
        if means3D.ndimension() != 2 or means3D.size(1) != 3:
            raise RuntimeError("means3D must have dimensions (num_points, 3)")
        means3D = _cabi.require_cuda("means3D", means3D)
        sh, colors_precomp = dense("sh", sh), dense("colors_precomp", colors_precomp)
        opacities, scales, rotations = dense("opacities", opacities), dense("scales", scales), dense("rotations", rotations)
        cov3Ds_precomp = dense("cov3Ds_precomp", cov3Ds_precomp)
        args = (raster_settings, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp)
        if raster_settings.debug:
            cpu_args = cpu_deep_copy_tuple(args[1:])  # copy them before they can be corrupted
            try:
                res = _forward_impl(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise ex
        else:
            res = _forward_impl(*args)
        num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, depth, opacity, n_touched = res

        ctx.raster_settings = raster_settings
        ctx.num_rendered = num_rendered
        ctx.pose_shapes = (tuple(theta.shape), tuple(rho.shape))
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer,
                              binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii, n_touched)
        # outputs nobody differentiates through (opacity, radii, n_touched; often one of color / depth) arrive as None in
        # backward instead of as freshly filled zero tensors: three fill kernels and their allocations less per view
        ctx.set_materialize_grads(False)
        return color, radii, depth, opacity, n_touched

    @staticmethod
    def backward(ctx, grad_out_color, grad_out_radii, grad_out_depth, grad_out_opacity, grad_n_touched):
        # grad_out_opacity / radii / n_touched are ignored, as in the reference (DGR __init__.py:116-138)
        S = ctx.raster_settings
        colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer, binningBuffer, imgBuffer = \
            ctx.saved_tensors
        dev = means3D.device
        H, W = int(S.image_height), int(S.image_width)
        if grad_out_color is None:
            grad_out_color = torch.zeros((3, H, W), dtype=torch.float32, device=dev)
        if grad_out_depth is None:
            grad_out_depth = torch.zeros((1, H, W), dtype=torch.float32, device=dev)
        # input order: means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, theta, rho
        ng = ctx.needs_input_grad
        needs = dict(means3D=ng[0], means2D=ng[1], sh=ng[2], colors=ng[3], opacity=ng[4],
                     scale_rot=ng[5] or ng[6], cov3D=ng[7])
        args = (S, _cabi.require_cuda("means3D", means3D), radii, colors_precomp, scales, rotations, cov3Ds_precomp,
                grad_out_color, grad_out_depth, sh, geomBuffer, ctx.num_rendered, binningBuffer, imgBuffer, needs)
        if S.debug:
            cpu_args = cpu_deep_copy_tuple(args[1:])
            try:
                res = _backward_impl(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise ex
        else:
            res = _backward_impl(*args)
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh, grad_scales,
         grad_rotations, grad_tau) = res
        grad_rho, grad_theta = _pose_grads(grad_tau, ctx.pose_shapes)
        if not ng[2]:
            grad_sh = None          # computed only for its pose term
        grads = (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities, grad_scales, grad_rotations,
                 grad_cov3Ds_precomp, grad_theta, grad_rho, None)
        return grads


class _RasterizeGaussiansLoss(torch.autograd.Function):
    """`_RasterizeGaussians` plus the mapping loss of the view, evaluated in the epilogue of the forward renderer
    (thirdparty/monogs/utils/slam_utils.py:71-105). Extra inputs: exposure_a, exposure_b (differentiable) and the
    LossRequest; extra output: the loss. Backward adds `grad_loss * dL/dimage` (and `* dL/ddepth`) to whatever arrives
    for color / depth directly."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, theta, rho,
                exposure_a, exposure_b, raster_settings, req):
        def dense(name, t):
            return _cabi.require_cuda(name, t) if t.numel() != 0 else t
        if means3D.ndimension() != 2 or means3D.size(1) != 3:
            raise RuntimeError("means3D must have dimensions (num_points, 3)")
        means3D = _cabi.require_cuda("means3D", means3D)
        sh, colors_precomp = dense("sh", sh), dense("colors_precomp", colors_precomp)
        opacities, scales, rotations = dense("opacities", opacities), dense("scales", scales), dense("rotations", rotations)
        cov3Ds_precomp = dense("cov3Ds_precomp", cov3Ds_precomp)
        res = _forward_impl(raster_settings, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                            loss=req)
        num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, depth, opacity, n_touched, sums, dimg, ddepth = res
        ctx.raster_settings = raster_settings
        ctx.num_rendered = num_rendered
        ctx.pose_shapes = (tuple(theta.shape), tuple(rho.shape))
        ctx.initialization = req.initialization
        ctx.exposure_shapes = (None if exposure_a is None else exposure_a.shape,
                               None if exposure_b is None else exposure_b.shape)
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer,
                              binningBuffer, imgBuffer, sums, dimg, ddepth)
        ctx.mark_non_differentiable(radii, n_touched)
        ctx.set_materialize_grads(False)
        return color, radii, depth, opacity, n_touched, sums[0]

    @staticmethod
    def backward(ctx, grad_out_color, grad_out_radii, grad_out_depth, grad_out_opacity, grad_n_touched, grad_loss):
        S = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer, binningBuffer, imgBuffer,
         sums, dimg, ddepth) = ctx.saved_tensors
        dev = means3D.device
        H, W = int(S.image_height), int(S.image_width)
        up_c, up_d = grad_out_color, grad_out_depth
        ga = gb = None
        if grad_loss is not None:
            up_c = dimg * grad_loss if up_c is None else up_c + dimg * grad_loss
            up_d = ddepth * grad_loss if up_d is None else up_d + ddepth * grad_loss
            if not ctx.initialization:
                ga = (grad_loss * sums[3]).reshape(ctx.exposure_shapes[0])
                gb = (grad_loss * sums[4]).reshape(ctx.exposure_shapes[1])
        if up_c is None:
            up_c = torch.zeros((3, H, W), dtype=torch.float32, device=dev)
        if up_d is None:
            up_d = torch.zeros((1, H, W), dtype=torch.float32, device=dev)
        ng = ctx.needs_input_grad
        needs = dict(means3D=ng[0], means2D=ng[1], sh=ng[2], colors=ng[3], opacity=ng[4],
                     scale_rot=ng[5] or ng[6], cov3D=ng[7])
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh, grad_scales,
         grad_rotations, grad_tau) = _backward_impl(S, means3D, radii, colors_precomp, scales, rotations, cov3Ds_precomp,
                                                    up_c, up_d, sh, geomBuffer, ctx.num_rendered, binningBuffer,
                                                    imgBuffer, needs)
        grad_rho, grad_theta = _pose_grads(grad_tau, ctx.pose_shapes)
        if not ng[2]:
            grad_sh = None
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities, grad_scales, grad_rotations,
                grad_cov3Ds_precomp, grad_theta, grad_rho, ga, gb, None, None)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """Boolean mask of points in front of the near plane (DGR __init__.py:193-202)."""
        with torch.no_grad():
            S = self.raster_settings
            positions = _cabi.require_cuda("positions", positions)
            P = positions.size(0)
            present = torch.zeros((P,), dtype=torch.bool, device=positions.device)
            if P != 0:
                with torch.cuda.device(positions.device):
                    _cabi.check(_cabi.lib().splat_raster_mark_visible(
                        P, positions.data_ptr(), _cabi.ptr(_cabi.require_cuda("viewmatrix", S.viewmatrix)),
                        _cabi.ptr(_cabi.require_cuda("projmatrix", S.projmatrix)), present.data_ptr(),
                        _cabi.stream_ptr(positions.device)), "mark_visible")
        return present

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, theta=None, rho=None):
        raster_settings = self.raster_settings

        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')

        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')

        if shs is None:
            shs = torch.Tensor([])
        if colors_precomp is None:
            colors_precomp = torch.Tensor([])
        if scales is None:
            scales = torch.Tensor([])
        if rotations is None:
            rotations = torch.Tensor([])
        if cov3D_precomp is None:
            cov3D_precomp = torch.Tensor([])
        if theta is None:
            theta = torch.Tensor([])
        if rho is None:
            rho = torch.Tensor([])

        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                                   theta, rho, raster_settings)
