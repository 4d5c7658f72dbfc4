"""Times the fused map deformation against the reference's per-keyframe torch sequence (oracle.deform on the GPU = the
same ops src/mapper.py:154-255 issues, without the optimizer re-registration) for 100 keyframes over 1 M Gaussians.
    gpurun -- 'python tests/perf/bench_deform.py'
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "splat-slam_b200"))
from diff_gaussian_rasterization.deform import deform_points_  # noqa: E402
import oracle.deform as od  # noqa: E402

P, NKF = 1_000_000, 99
c = od.synthetic_deform_case(P=P, n_kf=NKF, H=340, W=600, seed=0)
dev = "cuda"
g = {k: ([t.to(dev) for t in v] if isinstance(v, list) else (v.to(dev) if torch.is_tensor(v) else v)) for k, v in c.items()}
methods = [None if k % 2 else "rigid" for k in range(NKF + 1)]


def torch_loop():
    xyz, rot, scl = g["xyz"], g["rotation"], g["scaling"]
    for k, m in enumerate(methods):
        r = od.update_mapping_points(xyz, rot, scl, g["kf_ids"], 5 * k, g["w2c_new"][k], g["w2c_old"][k],
                                     g["depth_new"][k], g["depth_old"][k], g["intrinsics"], m)
        if r is not None:
            xyz, rot, scl = r


# the oracle builds its helpers on the CPU; run it on the GPU by moving the few constructors
_zeros, _ones, _tensor = torch.zeros, torch.ones, torch.tensor
torch.zeros = lambda *a, **k: _zeros(*a, **{"device": dev, **k})
torch.ones = lambda *a, **k: _ones(*a, **{"device": dev, **k})
torch.tensor = lambda *a, **k: _tensor(*a, **{"device": dev, **k})
for _ in range(2):
    torch_loop()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
torch_loop()
e1.record()
torch.cuda.synchronize()
print(f"torch sequence, {NKF + 1} keyframes: {e0.elapsed_time(e1):.1f} ms")
torch.zeros, torch.ones, torch.tensor = _zeros, _ones, _tensor

xyz, rot, scl = g["xyz"].clone(), g["rotation"].clone(), g["scaling"].clone()
ups = [(5 * k, g["w2c_new"][k], g["w2c_old"][k], g["depth_new"][k], g["depth_old"][k], m) for k, m in enumerate(methods)]
for _ in range(3):
    deform_points_(xyz, rot, scl, g["kf_ids"], ups, g["intrinsics"])
torch.cuda.synchronize()
e0.record()
for _ in range(10):
    deform_points_(xyz, rot, scl, g["kf_ids"], ups, g["intrinsics"])
e1.record()
torch.cuda.synchronize()
print(f"fused, {NKF + 1} keyframes in one call: {e0.elapsed_time(e1) / 10:.3f} ms")
