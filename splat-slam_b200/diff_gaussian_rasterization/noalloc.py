"""Binding of `splat_raster_backward_noalloc` (include/splat_raster.h): the backward pass for callers that cannot
allocate device memory at call time.

The regular backward needs P*12 floats of scratch from its caller. A tracking loop with a fixed frame budget, a
host whose device memory was all handed to fixed pools at start-up, or a thin foreign-language binding has no
allocator to ask at that point, so this variant accumulates in a fixed 12 MiB device array that lives inside the
library. Limits and rules (details in the header):

  * at most MAX_GAUSSIANS (262144) Gaussians, more is refused with `SplatRasterCapacityError`;
  * the first call on a device resolves the array's address; do it before any stream capture;
  * all calls on a device share the array — keep them ordered on one stream, overlapping calls corrupt each other;
  * the outputs are still the caller's tensors; this module allocates nothing on its own.
"""
from __future__ import annotations

import ctypes
from ctypes import c_float, c_int, c_void_p

from . import _cabi

MAX_GAUSSIANS = 262144
ERR_CAPACITY = -3


class SplatRasterCapacityError(_cabi.SplatRasterError):
    pass


_fn = None


def _entry():
    global _fn
    if _fn is None:
        f = _cabi.lib().splat_raster_backward_noalloc
        P = c_void_p
        f.restype = c_int
        f.argtypes = [c_int] * 6 + [P, P, P, P, P, c_float, P, P, P, P, P, P, c_float, c_float] + [P] * 16
        _fn = f
    return _fn


def backward_noalloc(P, D, M, R, W, H, background, means3D, shs, colors_precomp, scales, scale_modifier, rotations,
                     cov3D_precomp, viewmatrix, projmatrix, projmatrix_raw, campos, tan_fovx, tan_fovy, radii, geom,
                     binning, image, dL_dout_color, dL_dout_depth, dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D,
                     dL_dcov3D, dL_dsh, dL_dscales, dL_drotations, dL_dtau_partial, stream) -> None:
    """Raw call: every tensor argument is a CUDA tensor (or None for an absent optional), all outputs preallocated by
    the caller. Raises SplatRasterCapacityError when P exceeds MAX_GAUSSIANS."""
    p = _cabi.ptr
    rc = _entry()(int(P), int(D), int(M), int(R), int(W), int(H), p(background), p(means3D), p(shs), p(colors_precomp),
                  p(scales), float(scale_modifier), p(rotations), p(cov3D_precomp), p(viewmatrix), p(projmatrix),
                  p(projmatrix_raw), p(campos), float(tan_fovx), float(tan_fovy), p(radii), p(geom), p(binning),
                  p(image), p(dL_dout_color), p(dL_dout_depth), p(dL_dmeans2D), p(dL_dcolors), p(dL_dopacity),
                  p(dL_dmeans3D), p(dL_dcov3D), p(dL_dsh), p(dL_dscales), p(dL_drotations), p(dL_dtau_partial),
                  stream)
    if rc == ERR_CAPACITY:
        raise SplatRasterCapacityError(
            f"splat_raster_backward_noalloc: {P} Gaussians do not fit the library's built-in scratch "
            f"(limit {MAX_GAUSSIANS}); call splat_raster_backward with caller-provided scratch instead")
    _cabi.check(rc, "backward_noalloc")
