"""Golden vectors for the densification bookkeeping, produced with the REFERENCE's own
GaussianModel.add_densification_stats (thirdparty/gaussian_splatting/scene/gaussian_model.py:738-742), imported in the
build container. The module's unrelated third-party imports that this image lacks (open3d, plyfile, simple_knn) are
stubbed in sys.modules; the reference file itself is untouched. The max_radii2D line (src/mapper.py:520-523) sits
inline in Mapper.map and is executed here as the same two-tensor torch.max over the visibility filter.

    python tests/golden/make_golden_densify.py        # writes tests/golden/densify_stats.pt
"""
import os
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")
for name in ("open3d", "plyfile", "simple_knn", "simple_knn._C"):
    sys.modules.setdefault(name, types.ModuleType(name))
sys.modules["plyfile"].PlyData = sys.modules["plyfile"].PlyElement = object
sys.modules["simple_knn._C"].distCUDA2 = None
from thirdparty.gaussian_splatting.scene.gaussian_model import GaussianModel  # noqa: E402
from oracle.densify import synthetic_densify_case  # noqa: E402

out = {}
for name, (P, K, seed) in {"a": (500, 5, 0), "b": (1031, 12, 1)}.items():
    c = synthetic_densify_case(P, K, seed)
    model = types.SimpleNamespace(max_radii2D=c["max_radii2D"].clone(), xyz_gradient_accum=c["xyz_gradient_accum"].clone(),
                                  denom=c["denom"].clone())
    for radii, grad in zip(c["radii"], c["grads"]):
        vis = radii > 0
        model.max_radii2D[vis] = torch.max(model.max_radii2D[vis], radii[vis])
        GaussianModel.add_densification_stats(model, types.SimpleNamespace(grad=grad), vis)
    out[name] = dict(P=P, K=K, seed=seed, max_radii2D=model.max_radii2D, xyz_gradient_accum=model.xyz_gradient_accum,
                     denom=model.denom)
    print(name, float(model.xyz_gradient_accum.sum()), float(model.denom.sum()))
path = os.path.join(HERE, "densify_stats.pt")
if os.environ.get("GOLDEN_DRY"):
    old = torch.load(path, weights_only=False)
    for name, g in out.items():
        for k, v in g.items():
            assert torch.equal(v, old[name][k]) if torch.is_tensor(v) else v == old[name][k], (name, k)
else:
    torch.save(out, path)
