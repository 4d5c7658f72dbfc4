"""Fused parameter activations (first half): the algorithm is torch.exp / F.normalize /
torch.sigmoid (gaussian_model.py:53-61), so the oracle is those functions under autograd, checked against the reference
class's own properties when the reference tree is mounted; the CUDA kernels are checked against the oracle."""
import os
import sys
import types

import pytest
import torch

import parity as T  # noqa: F401  (sys.path)
from oracle.activations import activations, synthetic_raw


def test_oracle_is_the_reference_properties():
    if not os.path.isdir("/root/reference/thirdparty/gaussian_splatting"):
        pytest.skip("reference tree not present")
    sys.path.insert(0, "/root/reference")
    for name in ("open3d", "plyfile", "simple_knn", "simple_knn._C"):
        sys.modules.setdefault(name, types.ModuleType(name))
    if not hasattr(sys.modules["plyfile"], "PlyData"):
        sys.modules["plyfile"].PlyData = sys.modules["plyfile"].PlyElement = object
    if not hasattr(sys.modules["simple_knn._C"], "distCUDA2"):
        sys.modules["simple_knn._C"].distCUDA2 = None
    from thirdparty.gaussian_splatting.scene.gaussian_model import GaussianModel
    src = open("/root/reference/thirdparty/gaussian_splatting/scene/gaussian_model.py").read()
    assert "self.scaling_activation = torch.exp" in src and "self.opacity_activation = torch.sigmoid" in src
    assert "self.rotation_activation = torch.nn.functional.normalize" in src
    gm = GaussianModel.__new__(GaussianModel)
    gm.scaling_activation, gm.opacity_activation = torch.exp, torch.sigmoid
    gm.rotation_activation = torch.nn.functional.normalize
    gm._scaling, gm._rotation, gm._opacity = synthetic_raw(257, 1)
    a = activations(gm._scaling, gm._rotation, gm._opacity)
    for x, y in zip(a, (gm.get_scaling, gm.get_rotation, gm.get_opacity)):
        assert torch.equal(x, y)


@pytest.mark.gpu
@pytest.mark.parametrize("P", [1, 1000, 300_001])
def test_kernels_vs_oracle(P):
    from diff_gaussian_rasterization.activations import fused_activations
    raw = synthetic_raw(P, 2)
    g = torch.Generator().manual_seed(5)
    ups = [torch.randn(P, w, generator=g) for w in (3, 4, 1)]
    leaves_o = [t.clone().requires_grad_(True) for t in raw]
    out_o = activations(*leaves_o)
    torch.autograd.backward(list(out_o), ups)
    leaves = [t.cuda().requires_grad_(True) for t in raw]
    out = fused_activations(*leaves)
    torch.autograd.backward(list(out), [u.cuda() for u in ups])
    for a, b in zip(out, out_o):
        assert T.rel_err(a, b) <= 1e-6
    for a, b in zip(leaves, leaves_o):
        assert T.rel_err(a.grad, b.grad) <= 2e-6


@pytest.mark.gpu
def test_kernels_partial_gradients_and_zero_quaternion():
    from diff_gaussian_rasterization.activations import fused_activations
    s, q, o = (t.cuda() for t in synthetic_raw(100, 3))
    q[7] = 0.0                                            # normalize clamps the denominator at 1e-12
    s.requires_grad_(True)
    q.requires_grad_(True)
    scales, rots, opac = fused_activations(s, q, o)       # _opacity needs no gradient; opacities get no upstream one
    assert float(rots[7].detach().abs().max()) == 0.0
    (scales.sum() + (rots * rots).sum()).backward()
    assert torch.allclose(s.grad, torch.exp(s.detach()), rtol=1e-6)
    assert float(q.grad[:7].abs().max()) < 1e-5           # d/dq of ||normalize(q)||^2 is 0
    with pytest.raises(RuntimeError):
        fused_activations(s.detach().cpu(), q, o)
