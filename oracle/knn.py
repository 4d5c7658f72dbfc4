"""TEST INFRASTRUCTURE — CPU oracle of simple_knn.distCUDA2: mean squared distance to the 3 nearest neighbours.

Restates thirdparty/simple-knn/simple_knn.cu:129-172 (updateKBest<3>, boxMeanDist): for every point the three
smallest squared distances d = dx*dx + dy*dy + dz*dz to the OTHER points (coincident points count with 0), slots that
stay unfilled hold FLT_MAX, result (best0 + best1 + best2) / 3 in float32. The Morton sort and the boxes
(simple_knn.cu:47-127,174-216) only prune the search; the result is the exact 3-NN, which is what this brute-force
restatement computes. Squared distances are formed in float64 from float32 differences and rounded to float32 — the
GPU's product + two FMAs can differ from that by one float32 ulp (double rounding), so the bit-level check is against
the compiled reference (oracle/_ref/ref_simple_knn, oracle/build_ref_knn.sh) and the committed golden vectors it
produced on a B200 (tests/golden/knn_*.pt). Used only by tests.
"""
import numpy as np

FLT_MAX = np.float32(3.4028234663852886e38)


def dist_cuda2(points, chunk=2048):
    pts = np.ascontiguousarray(points, dtype=np.float32)
    P = pts.shape[0]
    out = np.empty(P, dtype=np.float32)
    for s in range(0, P, chunk):
        q = pts[s:s + chunk]
        d = (pts[None, :, :] - q[:, None, :]).astype(np.float64)          # float32 differences, as on the GPU
        d2 = (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1] + d[..., 2] * d[..., 2]).astype(np.float32)
        d2[np.arange(q.shape[0]), np.arange(s, s + q.shape[0])] = np.inf   # i == idx is skipped
        k = min(3, P - 1)
        best = np.full((q.shape[0], 3), FLT_MAX, dtype=np.float32)
        if k > 0:
            part = np.partition(d2, k - 1, axis=1)[:, :k]
            best[:, :k] = np.sort(part, axis=1)
        with np.errstate(over="ignore"):
            out[s:s + chunk] = ((best[:, 0] + best[:, 1]) + best[:, 2]) / np.float32(3.0)
    return out


def synthetic_points(P, seed=0, kind="surface"):
    """Point sets shaped like a back-projected depth map (a noisy surface), plus degenerate variants."""
    rng = np.random.default_rng(seed)
    if kind == "surface":
        u, v = rng.uniform(-1, 1, P), rng.uniform(-1, 1, P)
        z = 2.0 + 0.3 * np.sin(3 * u) * np.cos(2 * v) + 0.01 * rng.standard_normal(P)
        pts = np.stack([u * z, v * z, z], 1)
    elif kind == "uniform":
        pts = rng.uniform(-3, 3, (P, 3))
    elif kind == "clusters":
        c = rng.uniform(-5, 5, (8, 3))
        pts = c[rng.integers(0, 8, P)] + 0.01 * rng.standard_normal((P, 3))
    elif kind == "duplicates":
        base = rng.uniform(-1, 1, (max(P // 4, 1), 3))
        pts = base[rng.integers(0, base.shape[0], P)]
    elif kind == "line":
        pts = np.zeros((P, 3))
        pts[:, 0] = rng.uniform(0, 1, P)
    else:
        raise ValueError(kind)
    return pts.astype(np.float32)
