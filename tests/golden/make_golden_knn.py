"""Golden vectors for distCUDA2, produced by the UNMODIFIED reference extension (oracle/_ref/ref_simple_knn, built by
oracle/build_ref_knn.sh from thirdparty/simple-knn) on a B200:

    gpurun -- 'python tests/golden/make_golden_knn.py gpurun_out/golden'   # then copy gpurun_out/golden/knn.pt here
"""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
from ref_simple_knn._C import distCUDA2  # noqa: E402
from oracle.knn import synthetic_points  # noqa: E402

CASES = {"surface": (3000, 0), "uniform": (2500, 1), "clusters": (3000, 2), "duplicates": (1500, 3), "line": (1200, 4),
         "three": (3, 5), "four": (4, 6), "one": (1, 7)}
out = {}
for kind, (P, seed) in CASES.items():
    pts = synthetic_points(P, seed, kind if kind in ("surface", "uniform", "clusters", "duplicates", "line") else "uniform")
    out[kind] = dict(P=P, seed=seed, mean_dist2=distCUDA2(torch.from_numpy(pts).cuda()).cpu())
    print(kind, out[kind]["mean_dist2"][:3])
dst = sys.argv[1] if len(sys.argv) > 1 else HERE
os.makedirs(dst, exist_ok=True)
torch.save(out, os.path.join(dst, "knn.pt"))
