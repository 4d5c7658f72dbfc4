"""TEST INFRASTRUCTURE — CPU oracle of the optimizer step that follows backward in the mapping loop.

The reference builds `torch.optim.Adam(groups, lr=0.0, eps=1e-15)` with one lr per group
(thirdparty/gaussian_splatting/scene/gaussian_model.py:264-302) and `torch.optim.Adam(opt_params)` for the
per-keyframe pose / exposure parameters (src/mapper.py:1111); it steps them at src/mapper.py:557-562,703-708.
The algorithm therefore lives in a third-party dependency: PyTorch (the reference's Dockerfile pins torch 2.1;
this image has 2.11 — same update rule, `torch/optim/adam.py::_multi_tensor_adam`). This file restates that rule in
numpy float32, operation by operation; tests pin it against `torch.optim.Adam` itself (tests/test_adam.py).
Used only by tests.
"""
import numpy as np


def adam_step(param, grad, exp_avg, exp_avg_sq, step, lr, beta1=0.9, beta2=0.999, eps=1e-8):
    """One update, in place on float32 numpy arrays. `step` is the 1-based count of this update."""
    f = np.float32
    # exp_avg.lerp_(grad, 1 - beta1)           (weight < 0.5 branch: self + weight * (end - self))
    exp_avg += f(1 - beta1) * (grad - exp_avg)
    # exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
    exp_avg_sq *= f(beta2)
    exp_avg_sq += f(1 - beta2) * (grad * grad)
    bias_correction1 = 1 - beta1 ** step
    bias_correction2 = 1 - beta2 ** step
    step_size = lr / bias_correction1
    bias_correction2_sqrt = bias_correction2 ** 0.5
    denom = np.sqrt(exp_avg_sq) / f(bias_correction2_sqrt) + f(eps)
    # param.addcdiv_(exp_avg, denom, value=-step_size)
    param += f(-step_size) * (exp_avg / denom)
    return param


class AdamOracle:
    """Minimal optimizer over named float32 arrays with a per-name lr (the reference's param groups)."""

    def __init__(self, params, lrs, betas=(0.9, 0.999), eps=1e-8):
        self.params = {k: np.array(v, dtype=np.float32) for k, v in params.items()}
        self.lrs = dict(lrs)
        self.betas, self.eps = betas, eps
        self.m = {k: np.zeros_like(v) for k, v in self.params.items()}
        self.v = {k: np.zeros_like(v) for k, v in self.params.items()}
        self.t = {k: 0 for k in self.params}

    def step(self, grads):
        for k, g in grads.items():
            if g is None:
                continue
            self.t[k] += 1
            adam_step(self.params[k], np.asarray(g, dtype=np.float32), self.m[k], self.v[k], self.t[k], self.lrs[k],
                      self.betas[0], self.betas[1], self.eps)
