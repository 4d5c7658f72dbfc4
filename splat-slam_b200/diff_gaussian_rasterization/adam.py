"""Fused Adam for the Gaussian parameter groups and the per-keyframe pose/exposure parameters.

Drop-in for the two optimizers the mapping loop steps after every backward (src/mapper.py:557-562,703-708):
`torch.optim.Adam(l, lr=0.0, eps=1e-15)` built in thirdparty/gaussian_splatting/scene/gaussian_model.py:264-302 and
`torch.optim.Adam(opt_params)` of src/mapper.py:1111. Same constructor, same `param_groups`, same per-parameter
state keys (`step`, `exp_avg`, `exp_avg_sq`), so the reference's optimizer surgery on densify/prune
(gaussian_model.py:488-590: it swaps `group["params"][0]` and edits `optimizer.state[...]` by hand) and its
`update_learning_rate` (gaussian_model.py:315-329) work unchanged. `step()` is ONE kernel launch per 32 tensors
(`splat_raster_adam_step`) instead of seven multi-tensor passes. No CPU path.
"""
from __future__ import annotations

import ctypes
from ctypes import c_double, c_int, c_longlong, c_void_p

import torch

from . import _cabi


class _Segment(ctypes.Structure):
    _fields_ = [("param", c_void_p), ("grad", c_void_p), ("exp_avg", c_void_p), ("exp_avg_sq", c_void_p),
                ("n", c_longlong), ("lr", c_double), ("beta1", c_double), ("beta2", c_double), ("eps", c_double),
                ("step", c_longlong)]


_fn = None


def _entry():
    global _fn
    if _fn is None:
        f = _cabi.lib().splat_raster_adam_step
        f.restype = c_int
        f.argtypes = [c_int, ctypes.POINTER(_Segment), c_int, c_void_p]
        _fn = f
    return _fn


class GaussianAdam(torch.optim.Optimizer):
    """torch.optim.Adam(params, lr, betas, eps) with a fused CUDA step.

    zero_grads=True additionally clears every `.grad` inside the same kernel (for callers that keep the gradients in
    a persistent buffer such as `window.FlatGrads`; `zero_grad(set_to_none=True)` keeps working as in the reference).
    """

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False, *,
                 maximize=False, zero_grads=False):
        if weight_decay != 0 or amsgrad or maximize:
            raise NotImplementedError("GaussianAdam covers the reference's configuration only: "
                                      "weight_decay=0, amsgrad=False, maximize=False")
        if lr < 0.0 or eps < 0.0 or not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0:
            raise ValueError("invalid Adam hyper-parameters")
        if not betas[0] >= 0.5:
            raise NotImplementedError("beta1 < 0.5 takes torch.lerp's other rounding branch; not built")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))
        self.zero_grads = bool(zero_grads)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        per_device = {}
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                g = p.grad
                if g.is_sparse:
                    raise RuntimeError("GaussianAdam does not support sparse gradients")
                _cabi.require_cuda("param", p)
                _cabi.require_cuda("grad", g)
                if p.dtype != torch.float32 or not p.is_contiguous() or not g.is_contiguous():
                    raise RuntimeError("GaussianAdam needs contiguous float32 parameters and gradients")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                m, v = st["exp_avg"], st["exp_avg_sq"]
                if m.shape != p.shape or v.shape != p.shape or not m.is_contiguous() or not v.is_contiguous():
                    raise RuntimeError("optimizer state does not match its parameter (shape/contiguity)")
                st["step"] += 1
                per_device.setdefault(p.device, []).append(
                    _Segment(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(), float(group["lr"]),
                             float(beta1), float(beta2), float(group["eps"]), int(st["step"].item())))
        fn = _entry()
        for dev, segs in per_device.items():
            arr = (_Segment * len(segs))(*segs)
            with torch.cuda.device(dev):
                _cabi.check(fn(len(segs), arr, int(self.zero_grads), _cabi.stream_ptr(dev)), "adam_step")
        return loss
