"""Golden vectors for the map deformation, produced by the REFERENCE's own Mapper.update_mapping_points
(src/mapper.py:154-255). src/mapper.py cannot be imported here (lietorch, droid_backends, ...), so the method's
source is taken from the mounted reference AT RUN TIME (ast -> compile; nothing is copied into this repo) and bound to a
stand-in Mapper whose `gaussians` is a real reference GaussianModel (built without __init__, which wants a GPU) with a
real torch.optim.Adam, so replace_tensor_to_optimizer is the reference's own too. `.cuda()` is patched to the identity;
open3d / plyfile / simple_knn, which this image lacks and the method never touches, are stubbed in sys.modules.

    python tests/golden/make_golden_deform.py        # writes tests/golden/deform.pt
"""
import ast
import os
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)
for name in ("open3d", "plyfile", "simple_knn", "simple_knn._C"):
    sys.modules.setdefault(name, types.ModuleType(name))
sys.modules["plyfile"].PlyData = sys.modules["plyfile"].PlyElement = object
sys.modules["simple_knn._C"].distCUDA2 = None
torch.Tensor.cuda = lambda self, *a, **k: self
from thirdparty.gaussian_splatting.scene.gaussian_model import GaussianModel  # noqa: E402
from thirdparty.gaussian_splatting.utils.general_utils import quaternion_multiply, rotation_matrix_to_quaternion  # noqa: E402
from oracle.deform import synthetic_deform_case  # noqa: E402

tree = ast.parse(open(os.path.join(REF, "src", "mapper.py")).read())
fn = next(n for c in tree.body if isinstance(c, ast.ClassDef) and c.name == "Mapper"
          for n in c.body if isinstance(n, ast.FunctionDef) and n.name == "update_mapping_points")
ns = {"torch": torch, "rotation_matrix_to_quaternion": rotation_matrix_to_quaternion,
      "quaternion_multiply": quaternion_multiply}
exec(compile(ast.Module(body=[fn], type_ignores=[]), "reference:src/mapper.py", "exec"), ns)
reference_update_mapping_points = ns["update_mapping_points"]


def reference_model(c):
    gm = GaussianModel.__new__(GaussianModel)
    gm.scaling_activation, gm.rotation_activation = torch.exp, torch.nn.functional.normalize
    gm._xyz = torch.nn.Parameter(c["xyz"].clone())
    gm._rotation = torch.nn.Parameter(c["rotation"].clone())
    gm._scaling = torch.nn.Parameter(c["scaling"].clone())
    gm.unique_kfIDs = c["kf_ids"].clone()
    gm.optimizer = torch.optim.Adam([{"params": [gm._xyz], "lr": 1e-3, "name": "xyz"},
                                     {"params": [gm._scaling], "lr": 1e-3, "name": "scaling"},
                                     {"params": [gm._rotation], "lr": 1e-3, "name": "rotation"}], lr=0.0, eps=1e-15)
    for p in (gm._xyz, gm._rotation, gm._scaling):       # one step, so there is optimizer state to reset
        p.grad = torch.ones_like(p)
    gm.optimizer.step()
    gm.optimizer.zero_grad(set_to_none=True)
    return gm


out = {}
for name, (seed, methods) in {"mixed": (0, (None, "rigid", None, "rigid", None)), "all_rigid": (1, ("rigid",) * 5),
                              "all_depth": (2, (None,) * 5)}.items():
    c = synthetic_deform_case(seed=seed)
    gm = reference_model(c)
    start = {k: getattr(gm, k).detach().clone() for k in ("_xyz", "_rotation", "_scaling")}
    mapper = types.SimpleNamespace(gaussians=gm, device="cpu")
    for k, method in enumerate(methods):                   # keyframe ids 0,5,10,15,20 in turn; 99 matches nothing
        reference_update_mapping_points(mapper, 5 * k, c["w2c_new"][k], c["w2c_old"][k], c["depth_new"][k],
                                        c["depth_old"][k], c["intrinsics"], method=method)
    reference_update_mapping_points(mapper, 99, c["w2c_new"][0], c["w2c_old"][0], c["depth_new"][0], c["depth_old"][0],
                                    c["intrinsics"])
    st = gm.optimizer.state[gm._xyz]
    out[name] = dict(seed=seed, methods=methods, start=start, xyz=gm._xyz.detach().clone(),
                     rotation=gm._rotation.detach().clone(), scaling=gm._scaling.detach().clone(),
                     xyz_state_zeroed=bool(st["exp_avg"].abs().max() == 0 and st["exp_avg_sq"].abs().max() == 0))
    print(name, float(gm._xyz.abs().sum()), out[name]["xyz_state_zeroed"])
path = os.path.join(HERE, "deform.pt")
if os.environ.get("GOLDEN_DRY"):
    old = torch.load(path, weights_only=False)
    for name, g in out.items():
        for k in ("xyz", "rotation", "scaling"):
            assert torch.equal(g[k], old[name][k]), (name, k)
else:
    torch.save(out, path)
