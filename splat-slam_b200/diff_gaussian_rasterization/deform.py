"""Map deformation: move the Gaussians anchored to updated keyframes with their keyframe.

Drop-in for Mapper.update_mapping_points(frame_idx, w2c, w2c_old, depth, depth_old, intrinsics, method=None)
(src/mapper.py:154-255; called per keyframe at mapper.py:640-647 and :1048-1055), taking the GaussianModel instead of
`self`, plus a batched form that handles every updated keyframe of a mapping call in ONE pass over the Gaussians
(two kernel launches, no host synchronisation) instead of ~40 torch kernels and three mask-index syncs per keyframe.
The optimizer bookkeeping stays the reference's own: `gaussians.replace_tensor_to_optimizer(...)` for xyz, rotation
and (non-rigid only) scaling, which also resets their Adam state (gaussian_model.py:488-501). No CPU path.
"""
from __future__ import annotations

import ctypes
from ctypes import POINTER, c_int, c_void_p

import torch

from . import _cabi

TABLE_FLOATS = 2560        # SPLAT_DEFORM_TABLE_FLOATS


class _Keyframe(ctypes.Structure):
    _fields_ = [("w2c_old", c_void_p), ("w2c_new", c_void_p), ("depth_new", c_void_p), ("depth_old", c_void_p),
                ("frame_idx", c_int), ("rigid", c_int)]


_fn = None


def _entry():
    global _fn
    if _fn is None:
        f = _cabi.lib().splat_raster_deform_points
        f.restype = c_int
        f.argtypes = [c_int, c_int, POINTER(_Keyframe), c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p,
                      c_void_p, c_void_p]
        _fn = f
    return _fn


def _dev_f32(t, dev):
    return torch.as_tensor(t).to(device=dev, dtype=torch.float32).contiguous()


@torch.no_grad()
def deform_points_(xyz, rotation, scaling, kf_ids, updates, intrinsics):
    """In place on contiguous float32 CUDA tensors xyz [P,3], rotation [P,4], scaling [P,3]; kf_ids [P] int32 on the same
    device. updates: iterable of (frame_idx, w2c_new, w2c_old, depth_new, depth_old, method) with method "rigid" or None.
    Every rotation is written back normalised (mapper.py:176,245)."""
    for name, t, w in (("xyz", xyz, 3), ("rotation", rotation, 4), ("scaling", scaling, 3)):
        if not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous() or t.dim() != 2 or t.size(1) != w:
            raise _cabi.SplatRasterError(f"{name} must be a contiguous float32 CUDA tensor of shape (P, {w})")
    P, dev = xyz.size(0), xyz.device
    if not kf_ids.is_cuda or kf_ids.dtype != torch.int32 or kf_ids.numel() != P:
        raise _cabi.SplatRasterError("kf_ids must be an int32 CUDA tensor with one entry per Gaussian")
    keep, recs, W, H = [], [], 0, 0
    for frame_idx, w2c, w2c_old, depth, depth_old, method in updates:
        rigid = method == "rigid"
        a, b = _dev_f32(w2c_old, dev), _dev_f32(w2c, dev)
        if a.shape != (4, 4) or b.shape != (4, 4):
            raise _cabi.SplatRasterError("w2c / w2c_old must be 4x4")
        dn = dold = None
        if not rigid:
            dn, dold = _dev_f32(depth, dev), _dev_f32(depth_old, dev)
            if dn.dim() != 2 or dn.shape != dold.shape or (W and (dn.size(1), dn.size(0)) != (W, H)):
                raise _cabi.SplatRasterError("depth maps must be 2-D and share one size")
            H, W = dn.shape
        keep += [a, b, dn, dold]
        recs.append(_Keyframe(a.data_ptr(), b.data_ptr(), _cabi.ptr(dn), _cabi.ptr(dold), int(frame_idx), int(rigid)))
    if not recs or P == 0:
        return
    K = _dev_f32(intrinsics, dev).reshape(-1) if intrinsics is not None else None
    if K is not None and K.numel() != 9:
        raise _cabi.SplatRasterError("intrinsics must be 3x3")
    with torch.cuda.device(dev):
        table = torch.empty(TABLE_FLOATS, dtype=torch.float32, device=dev)
        arr = (_Keyframe * len(recs))(*recs)
        _cabi.check(_entry()(P, len(recs), arr, kf_ids.data_ptr(), _cabi.ptr(K), W, H, table.data_ptr(), xyz.data_ptr(),
                             rotation.data_ptr(), scaling.data_ptr(), _cabi.stream_ptr(dev)), "deform_points")
    del keep


@torch.no_grad()
def update_mapping_points_batch(gaussians, updates, intrinsics):
    """All keyframe updates of one mapping call at once. Returns the frame ids that had Gaussians (the others are
    skipped like the reference's early return, mapper.py:159-160 — nothing is touched if none has any)."""
    ids = gaussians.unique_kfIDs
    present = set(torch.unique(ids).tolist())
    updates = [u for u in updates if int(u[0]) in present]
    if not updates:
        return []
    dev = gaussians._xyz.device
    kf_ids = ids.to(device=dev, dtype=torch.int32)
    xyz, rot, scl = gaussians._xyz.detach(), gaussians._rotation.detach(), gaussians._scaling.detach()
    deform_points_(xyz, rot, scl, kf_ids, updates, intrinsics)
    # the reference re-registers the edited tensors with the optimizer, which zeroes their Adam moments
    gaussians._xyz = gaussians.replace_tensor_to_optimizer(xyz, "xyz")["xyz"]
    gaussians._rotation = gaussians.replace_tensor_to_optimizer(rot, "rotation")["rotation"]
    if any(u[5] != "rigid" for u in updates):
        gaussians._scaling = gaussians.replace_tensor_to_optimizer(scl, "scaling")["scaling"]
    return [int(u[0]) for u in updates]


def update_mapping_points(gaussians, frame_idx, w2c, w2c_old, depth, depth_old, intrinsics, method=None):
    """Mapper.update_mapping_points with the GaussianModel in place of `self` (src/mapper.py:154-255)."""
    update_mapping_points_batch(gaussians, [(frame_idx, w2c, w2c_old, depth, depth_old, method)], intrinsics)
