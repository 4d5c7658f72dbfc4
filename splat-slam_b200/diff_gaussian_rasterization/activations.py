"""Fused Gaussian parameter activations (forward and backward, one kernel each).

Replaces the three properties the renderer reads on every call, get_scaling = exp(_scaling), get_rotation =
normalize(_rotation), get_opacity = sigmoid(_opacity) (thirdparty/gaussian_splatting/scene/gaussian_model.py:53-61,
76-94; read in gaussian_renderer/__init__.py:76-90). In a window step call it once per iteration and hand the three
results to every view's rasterizer call: autograd sums the views' gradients on them and runs this backward once.
No CPU path.
"""
from __future__ import annotations

from ctypes import c_int, c_void_p

import torch

from . import _cabi

_fns = None


def _entry():
    global _fns
    if _fns is None:
        L = _cabi.lib()
        L.splat_raster_activations_forward.restype = c_int
        L.splat_raster_activations_forward.argtypes = [c_int] + [c_void_p] * 7
        L.splat_raster_activations_backward.restype = c_int
        L.splat_raster_activations_backward.argtypes = [c_int] + [c_void_p] * 10
        _fns = (L.splat_raster_activations_forward, L.splat_raster_activations_backward)
    return _fns


class _Activate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, raw_scaling, raw_rotation, raw_opacity):
        fwd, _ = _entry()
        s = _cabi.require_cuda("_scaling", raw_scaling)
        q = _cabi.require_cuda("_rotation", raw_rotation)
        o = _cabi.require_cuda("_opacity", raw_opacity)
        P = s.size(0)
        if s.shape != (P, 3) or q.shape != (P, 4) or o.numel() != P:
            raise _cabi.SplatRasterError("expected _scaling (P,3), _rotation (P,4), _opacity (P,1)")
        scales, rots, opac = torch.empty_like(s), torch.empty_like(q), torch.empty_like(o)
        with torch.cuda.device(s.device):
            _cabi.check(fwd(P, _cabi.ptr(s), _cabi.ptr(q), _cabi.ptr(o), _cabi.ptr(scales), _cabi.ptr(rots),
                            _cabi.ptr(opac), _cabi.stream_ptr(s.device)), "activations_forward")
        ctx.save_for_backward(s, q, o)
        return scales, rots, opac

    @staticmethod
    def backward(ctx, g_scales, g_rots, g_opac):
        _, bwd = _entry()
        s, q, o = ctx.saved_tensors
        P = s.size(0)
        need = ctx.needs_input_grad

        def grad_in(g, like):
            return None if g is None else _cabi.require_cuda("grad", g.expand_as(like))

        gs, gq, go = grad_in(g_scales, s), grad_in(g_rots, q), grad_in(g_opac, o)
        ds = torch.empty_like(s) if need[0] else None
        dq = torch.empty_like(q) if need[1] else None
        do = torch.empty_like(o) if need[2] else None
        if P:
            with torch.cuda.device(s.device):
                _cabi.check(bwd(P, _cabi.ptr(s), _cabi.ptr(q), _cabi.ptr(o), _cabi.ptr(gs), _cabi.ptr(gq), _cabi.ptr(go),
                                _cabi.ptr(ds), _cabi.ptr(dq), _cabi.ptr(do), _cabi.stream_ptr(s.device)),
                            "activations_backward")
        return ds, dq, do


def fused_activations(raw_scaling, raw_rotation, raw_opacity):
    """(exp(_scaling), normalize(_rotation), sigmoid(_opacity)) = (get_scaling, get_rotation, get_opacity)."""
    return _Activate.apply(raw_scaling, raw_rotation, raw_opacity)
