"""distCUDA2 — mean squared distance of every point to its three nearest neighbours.

Drop-in for `from simple_knn._C import distCUDA2` (thirdparty/simple-knn/spatial.cu:15-26), which
GaussianModel.create_pcd_from_image_and_depth calls at every keyframe insert to set the initial Gaussian scales
(thirdparty/gaussian_splatting/scene/gaussian_model.py:194-200). Same argument (an (N,3) float CUDA tensor), same
(N,) float32 result, but no host synchronisation and no cudaMalloc inside the call. No CPU path.
"""
from __future__ import annotations

from ctypes import c_int, c_size_t, c_void_p

import torch

from . import _cabi

_fns = None


def _entry():
    global _fns
    if _fns is None:
        L = _cabi.lib()
        L.splat_raster_knn_scratch_bytes.restype = c_size_t
        L.splat_raster_knn_scratch_bytes.argtypes = [c_int]
        L.splat_raster_knn_mean_dist2.restype = c_int
        L.splat_raster_knn_mean_dist2.argtypes = [c_int, c_void_p, c_void_p, c_void_p, c_void_p]
        _fns = (L.splat_raster_knn_scratch_bytes, L.splat_raster_knn_mean_dist2)
    return _fns


def distCUDA2(points: torch.Tensor) -> torch.Tensor:
    nbytes, run = _entry()
    points = _cabi.require_cuda("points", points)
    if points.dim() != 2 or points.size(1) != 3:
        raise _cabi.SplatRasterError("points must have shape (N, 3)")
    P = points.size(0)
    dev = points.device
    means = torch.full((P,), 0.0, dtype=torch.float32, device=dev)       # spatial.cu:21
    if P == 0:
        return means
    with torch.cuda.device(dev):
        scratch = torch.empty(nbytes(P), dtype=torch.uint8, device=dev)
        _cabi.check(run(P, points.data_ptr(), scratch.data_ptr(), means.data_ptr(), _cabi.stream_ptr(dev)),
                    "knn_mean_dist2")
    return means
