"""Densification bookkeeping after backward — one launch for all views of a mapping iteration.

Replaces the per-view loop of src/mapper.py:519-529 with GaussianModel.add_densification_stats
(thirdparty/gaussian_splatting/scene/gaussian_model.py:738-742), and the occlusion-aware visibility of
src/mapper.py:494-499,507-509. The reference's boolean-mask indexing costs a nonzero() with a host sync plus a
gather/scatter per statement and per view; here nothing synchronises. No CPU path.
"""
from __future__ import annotations

import ctypes
from ctypes import POINTER, c_int, c_void_p

import torch

from . import _cabi

_fns = None


def _entry():
    global _fns
    if _fns is None:
        L = _cabi.lib()
        L.splat_raster_densify_stats.restype = c_int
        L.splat_raster_densify_stats.argtypes = [c_int, c_int, POINTER(c_void_p), POINTER(c_void_p), c_int, c_void_p,
                                                 c_void_p, c_void_p, c_void_p]
        L.splat_raster_visibility_count.restype = c_int
        L.splat_raster_visibility_count.argtypes = [c_int, c_int, POINTER(c_void_p), POINTER(c_void_p), c_void_p, c_void_p]
        _fns = (L.splat_raster_densify_stats, L.splat_raster_visibility_count)
    return _fns


def _table(tensors):
    return (c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])


def _stat(name, t, P):
    if not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous() or t.numel() != P:
        raise _cabi.SplatRasterError(f"{name} must be a contiguous float32 CUDA tensor with one entry per Gaussian")
    return t


@torch.no_grad()
def add_densification_stats_window(gaussians, viewspace_point_tensors, radii_list):
    """For every rendered view, in order: gaussians.max_radii2D / xyz_gradient_accum / denom updated in place for the
    Gaussians with radii > 0 (the view's `visibility_filter`). `viewspace_point_tensors[k].grad` is the [P,3] screen-space
    gradient the rasterizer's backward produced (a tensor without `.grad` is taken as the gradient itself)."""
    K = len(radii_list)
    if K != len(viewspace_point_tensors):
        raise ValueError("one viewspace tensor per radii tensor")
    if K == 0:
        return
    P = radii_list[0].numel()
    grads = []
    for vs in viewspace_point_tensors:
        g = vs.grad if getattr(vs, "grad", None) is not None else vs
        g = _cabi.require_cuda("viewspace gradient", g)
        if g.dim() != 2 or g.size(0) != P or g.size(1) < 2:
            raise _cabi.SplatRasterError("viewspace gradient must have shape (P, >=2)")
        grads.append(g)
    stride = grads[0].size(1)
    if any(g.size(1) != stride for g in grads):
        raise _cabi.SplatRasterError("viewspace gradients must share one row width")
    radii = []
    for r in radii_list:
        if not r.is_cuda or r.numel() != P:
            raise _cabi.SplatRasterError("radii must be CUDA tensors with one entry per Gaussian")
        radii.append(r.to(torch.int32).contiguous())
    mr = _stat("max_radii2D", gaussians.max_radii2D, P)
    acc = _stat("xyz_gradient_accum", gaussians.xyz_gradient_accum, P)
    dn = _stat("denom", gaussians.denom, P)
    dev = mr.device
    with torch.cuda.device(dev):
        _cabi.check(_entry()[0](P, K, _table(radii), _table(grads), stride, mr.data_ptr(), acc.data_ptr(),
                                dn.data_ptr(), _cabi.stream_ptr(dev)), "densify_stats")


@torch.no_grad()
def occ_aware_visibility(n_touched_list, want_n_obs=True):
    """[(n_touched > 0).long() for each window view], and n_obs = how many views touched each Gaussian (int32, on the
    device — the reference accumulates it on the CPU, src/mapper.py:507-509)."""
    K = len(n_touched_list)
    if K == 0:
        return [], None
    P = n_touched_list[0].numel()
    nt = []
    for t in n_touched_list:
        if not t.is_cuda or t.numel() != P:
            raise _cabi.SplatRasterError("n_touched must be CUDA tensors with one entry per Gaussian")
        nt.append(t.to(torch.int32).contiguous())
    dev = nt[0].device
    vis = [torch.empty(P, dtype=torch.int64, device=dev) for _ in range(K)]
    n_obs = torch.empty(P, dtype=torch.int32, device=dev) if want_n_obs else None
    with torch.cuda.device(dev):
        _cabi.check(_entry()[1](P, K, _table(nt), _table(vis), _cabi.ptr(n_obs), _cabi.stream_ptr(dev)),
                    "visibility_count")
    return vis, n_obs
