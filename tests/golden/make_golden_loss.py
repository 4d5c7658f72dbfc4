"""Golden vectors for the mapping loss, produced by the REFERENCE's own get_loss_mapping imported in the build
container (thirdparty/monogs/utils/slam_utils.py:71-105). The reference moves its inputs with `.cuda()` /
`device=image.device`; here `Tensor.cuda` is patched to the identity so the unmodified function runs on the CPU.

    python tests/golden/make_golden_loss.py        # writes tests/golden/mapping_loss.pt
"""
import os
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")
torch.Tensor.cuda = lambda self, *a, **k: self
from thirdparty.monogs.utils.slam_utils import get_loss_mapping  # noqa: E402
from oracle.mapping_loss import synthetic_loss_case  # noqa: E402

out = {}
for name, (H, W, seed, alpha, init) in {"a": (48, 64, 0, 0.8, False), "b": (70, 100, 1, 0.95, False),
                                        "init": (48, 64, 2, 0.8, True), "ssim": (48, 64, 3, 0.8, False)}.items():
    c = synthetic_loss_case(H, W, seed)
    leaves = {k: c[k].clone().requires_grad_(True) for k in ("image", "depth", "exposure_a", "exposure_b")}
    vp = types.SimpleNamespace(original_image=c["gt_image"], depth=c["gt_depth"][0].numpy(),
                               exposure_a=leaves["exposure_a"], exposure_b=leaves["exposure_b"])
    cfg = {"Training": {"alpha": alpha, "rgb_boundary_threshold": 0.01, "ssim_loss": name == "ssim"},
           "opt_params": {"lambda_dssim": 0.2}}
    loss = get_loss_mapping(cfg, leaves["image"], leaves["depth"], vp, None, initialization=init)
    loss.backward()
    out[name] = dict(H=H, W=W, seed=seed, alpha=alpha, initialization=init, loss=loss.detach(),
                     **{"grad_" + k: (None if v.grad is None else v.grad.clone()) for k, v in leaves.items()})
    print(name, float(loss))
if os.environ.get("GOLDEN_DRY"):      # check the committed fixture against a fresh run instead of rewriting it
    old = torch.load(os.path.join(HERE, "mapping_loss.pt"), weights_only=False)
    for name, g in out.items():
        for k, v in g.items():
            w = old[name][k]
            assert (v is None and w is None) or (torch.equal(v, w) if torch.is_tensor(v) else v == w), (name, k)
else:
    torch.save(out, os.path.join(HERE, "mapping_loss.pt"))
