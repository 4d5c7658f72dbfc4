"""distCUDA2: CPU oracle against the golden vectors the compiled reference produced on a B200,
CUDA kernels against the compiled reference (bit-exact), the golden vectors (bit-exact) and the oracle."""
import os
import sys

import numpy as np
import pytest
import torch

import parity as T  # noqa: F401  (sys.path)
from oracle.knn import dist_cuda2, synthetic_points

GOLD = os.path.join(os.path.dirname(__file__), "golden", "knn.pt")
KINDS = ("surface", "uniform", "clusters", "duplicates", "line")


def _points(kind, P, seed):
    return synthetic_points(P, seed, kind if kind in KINDS else "uniform")


def _ulps(a, b):
    a = np.ascontiguousarray(a, dtype=np.float32).view(np.int32).astype(np.int64)
    b = np.ascontiguousarray(b, dtype=np.float32).view(np.int32).astype(np.int64)
    return int(np.abs(a - b).max()) if a.size else 0


def load_ref_knn():
    ref_dir = os.path.join(os.path.dirname(os.path.dirname(__file__)), "oracle", "_ref")
    if not os.path.isdir(os.path.join(ref_dir, "ref_simple_knn")):
        return None
    if ref_dir not in sys.path:
        sys.path.insert(0, ref_dir)
    try:
        from ref_simple_knn._C import distCUDA2
        return distCUDA2
    except Exception:
        return None


def test_oracle_known_answers():
    g = np.array([[0, 0, 0], [1, 0, 0], [0, 2, 0], [0, 0, 3], [5, 5, 5]], dtype=np.float32)
    np.testing.assert_allclose(dist_cuda2(g), [14 / 3, 16 / 3, 22 / 3, 32 / 3, 179 / 3], rtol=1e-7)
    assert np.isinf(dist_cuda2(g[:1])[0])                       # no neighbour at all: 3 x FLT_MAX overflows
    assert dist_cuda2(g[:3])[0] > 1e38                          # unfilled slot stays FLT_MAX
    dup = np.zeros((5, 3), dtype=np.float32)
    assert np.all(dist_cuda2(dup) == 0.0)                       # coincident points count, with distance 0


def test_oracle_matches_reference_golden():
    gold = torch.load(GOLD, weights_only=False)
    assert set(KINDS) <= set(gold)
    for kind, g in gold.items():
        o = dist_cuda2(_points(kind, g["P"], g["seed"]))
        r = g["mean_dist2"].numpy()
        assert np.array_equal(np.isinf(o), np.isinf(r)), kind
        fin = np.isfinite(r)
        # float64-then-round vs the GPU's product + two FMAs: at most an ulp per distance, so a couple on the mean
        assert _ulps(o[fin], r[fin]) <= 3, (kind, _ulps(o[fin], r[fin]))


@pytest.mark.gpu
def test_kernel_vs_golden_bit_exact():
    from diff_gaussian_rasterization.knn import distCUDA2
    for kind, g in torch.load(GOLD, weights_only=False).items():
        ours = distCUDA2(torch.from_numpy(_points(kind, g["P"], g["seed"])).cuda()).cpu()
        assert torch.equal(ours, g["mean_dist2"]), kind


@pytest.mark.gpu
@pytest.mark.parametrize("kind,P", [("surface", 40_000), ("uniform", 100_003), ("clusters", 65_536),
                                    ("duplicates", 10_000), ("line", 5_000), ("surface", 1_000_000)])
def test_kernel_vs_compiled_reference_and_oracle(kind, P):
    from diff_gaussian_rasterization.knn import distCUDA2
    pts = torch.from_numpy(synthetic_points(P, 11, kind)).cuda()
    ours = distCUDA2(pts)
    assert ours.shape == (P,) and ours.dtype == torch.float32
    ref_fn = load_ref_knn()
    if ref_fn is not None:
        assert torch.equal(ours, ref_fn(pts)), "differs from the compiled reference"
    # spot-check 64 queries by brute force on the device (separate torch ops: within an ulp or two of the FMA form)
    g = torch.Generator().manual_seed(0)
    q = torch.randint(0, P, (64,), generator=g).cuda()
    d = pts[q][:, None, :].double() - pts[None, :, :].double()
    d2 = (d * d).sum(-1)
    d2[torch.arange(64, device="cuda"), q] = float("inf")
    best = d2.topk(3, dim=1, largest=False).values.sum(1) / 3.0
    assert T.rel_err(ours[q], best.float()) <= 2e-6
    if P <= 40_000 and ref_fn is None:
        assert _ulps(ours.cpu().numpy(), dist_cuda2(pts.cpu().numpy())) <= 3


@pytest.mark.gpu
def test_kernel_edge_cases_and_errors():
    from diff_gaussian_rasterization.knn import distCUDA2
    assert distCUDA2(torch.zeros(0, 3, device="cuda")).shape == (0,)
    assert torch.isinf(distCUDA2(torch.ones(1, 3, device="cuda")))[0]
    three = distCUDA2(torch.eye(3, device="cuda"))
    assert float(three.min()) > 1e38
    assert float(distCUDA2(torch.zeros(100, 3, device="cuda")).abs().max()) == 0.0
    with pytest.raises(RuntimeError):
        distCUDA2(torch.zeros(10, 3))                 # CPU tensor: no CPU path
    with pytest.raises(RuntimeError):
        distCUDA2(torch.zeros(10, 2, device="cuda"))
    # two calls in a row on the same input give the same bits (the scratch carries no state)
    pts = torch.from_numpy(synthetic_points(5000, 1)).cuda()
    assert torch.equal(distCUDA2(pts), distCUDA2(pts))
