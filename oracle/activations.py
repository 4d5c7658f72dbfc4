"""TEST INFRASTRUCTURE — CPU oracle of the Gaussian parameter activations.

The reference's GaussianModel sets scaling_activation = torch.exp, opacity_activation = torch.sigmoid,
rotation_activation = torch.nn.functional.normalize (thirdparty/gaussian_splatting/scene/gaussian_model.py:53-61) and
applies them in get_scaling / get_rotation / get_opacity (:76-94). The algorithm is those three PyTorch functions, so
the restatement IS them, differentiated by autograd; tests also evaluate the reference class's own properties when
the reference tree is mounted. Used only by tests.
"""
import torch


def activations(raw_scaling, raw_rotation, raw_opacity):
    return torch.exp(raw_scaling), torch.nn.functional.normalize(raw_rotation), torch.sigmoid(raw_opacity)


def synthetic_raw(P, seed=0):
    g = torch.Generator().manual_seed(seed)
    s = torch.log(0.005 + 0.1 * torch.rand(P, 3, generator=g))
    q = torch.randn(P, 4, generator=g) * (0.2 + 2.0 * torch.rand(P, 1, generator=g))
    o = torch.randn(P, 1, generator=g) * 3.0
    return s, q, o
