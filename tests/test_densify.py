"""Densification bookkeeping: oracle pinned on the reference's own add_densification_stats
(tests/golden/make_golden_densify.py), CUDA kernels checked against oracle and golden through the C ABI."""
import os
import subprocess
import sys
import types

import pytest
import torch

import parity as T  # noqa: F401  (sys.path)
from oracle.densify import densification_stats_window, occ_aware_visibility, synthetic_densify_case

GOLD = os.path.join(os.path.dirname(__file__), "golden", "densify_stats.pt")
KEYS = ("max_radii2D", "xyz_gradient_accum", "denom")


def test_oracle_matches_reference_golden():
    for name, g in torch.load(GOLD, weights_only=False).items():
        c = synthetic_densify_case(g["P"], g["K"], g["seed"])
        densification_stats_window(c["max_radii2D"], c["xyz_gradient_accum"], c["denom"], c["radii"], c["grads"])
        for k in KEYS:
            assert torch.equal(c[k], g[k]), (name, k)


def test_reference_function_live_when_available():
    if not os.path.isdir("/root/reference/thirdparty/gaussian_splatting"):
        pytest.skip("reference tree not present")
    r = subprocess.run([sys.executable, os.path.join(os.path.dirname(__file__), "golden", "make_golden_densify.py")],
                       capture_output=True, text=True, cwd="/tmp", env=dict(os.environ, GOLDEN_DRY="1"))
    assert r.returncode == 0, r.stderr[-500:]


def test_visibility_oracle_counts():
    c = synthetic_densify_case(64, 3, 2)
    vis, n_obs = occ_aware_visibility(c["n_touched"])
    assert all(v.dtype == torch.int64 for v in vis) and int(n_obs.max()) <= 3
    assert torch.equal(n_obs.long(), sum(vis))


def _gpu_state(c):
    return types.SimpleNamespace(**{k: c[k].clone().cuda() for k in KEYS})


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["a", "b"])
def test_kernel_vs_golden(name):
    from diff_gaussian_rasterization.densify import add_densification_stats_window
    g = torch.load(GOLD, weights_only=False)[name]
    c = synthetic_densify_case(g["P"], g["K"], g["seed"])
    st = _gpu_state(c)
    vs = [types.SimpleNamespace(grad=x.cuda()) for x in c["grads"]]
    add_densification_stats_window(st, vs, [r.cuda() for r in c["radii"]])
    assert torch.equal(st.max_radii2D.cpu(), g["max_radii2D"])
    assert torch.equal(st.denom.cpu(), g["denom"])
    assert T.rel_err(st.xyz_gradient_accum, g["xyz_gradient_accum"]) <= 1e-6


@pytest.mark.gpu
def test_kernel_more_views_than_one_launch_and_edge_cases():
    from diff_gaussian_rasterization.densify import add_densification_stats_window, occ_aware_visibility as occ_gpu
    c = synthetic_densify_case(100_003, 37, 3)           # 37 views = 3 launches, same per-Gaussian sequence
    ref = {k: c[k].clone() for k in KEYS}
    densification_stats_window(ref["max_radii2D"], ref["xyz_gradient_accum"], ref["denom"], c["radii"], c["grads"])
    st = _gpu_state(c)
    add_densification_stats_window(st, [x.cuda() for x in c["grads"]], [r.cuda() for r in c["radii"]])
    assert torch.equal(st.max_radii2D.cpu(), ref["max_radii2D"]) and torch.equal(st.denom.cpu(), ref["denom"])
    assert T.rel_err(st.xyz_gradient_accum, ref["xyz_gradient_accum"]) <= 1e-6
    # nothing visible: state untouched, bit for bit
    st2 = _gpu_state(c)
    add_densification_stats_window(st2, [x.cuda() for x in c["grads"][:2]],
                                   [torch.zeros(100_003, dtype=torch.int32, device="cuda")] * 2)
    for k in KEYS:
        assert torch.equal(getattr(st2, k).cpu(), c[k])
    add_densification_stats_window(st2, [], [])             # empty window is a no-op
    # visibility
    vis_o, n_obs_o = occ_aware_visibility(c["n_touched"])
    vis, n_obs = occ_gpu([t.cuda() for t in c["n_touched"]])
    assert all(v.dtype == torch.int64 and torch.equal(v.cpu(), o) for v, o in zip(vis, vis_o))
    assert torch.equal(n_obs.cpu(), n_obs_o)
    with pytest.raises(RuntimeError):
        add_densification_stats_window(_gpu_state(c), [c["grads"][0]], [c["radii"][0].cuda()])    # CPU gradient


@pytest.mark.gpu
def test_kernel_on_real_rasterizer_outputs():
    """radii and the screen-space gradient come straight from the rasterizer's forward/backward."""
    from diff_gaussian_rasterization.densify import add_densification_stats_window
    m = T.load_ours()
    P = 4000
    radii_l, grad_l = [], []
    for k in range(3):
        case = T.make_case(camera="small", P=P, view_k=k)
        S = T.settings_for(m, case, "cuda")
        leaves = {n: case[n].cuda().requires_grad_(True) for n in ("means3D", "opacities", "shs", "scales", "rotations")}
        means2D = torch.zeros(P, 3, device="cuda", requires_grad=True)
        color, radii, depth, opacity, n_touched = m.GaussianRasterizer(S)(
            means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opacities"], shs=leaves["shs"],
            scales=leaves["scales"], rotations=leaves["rotations"])
        (color.sum() + depth.sum()).backward()
        radii_l.append(radii)
        grad_l.append(means2D)
    st = types.SimpleNamespace(max_radii2D=torch.zeros(P, device="cuda"), xyz_gradient_accum=torch.zeros(P, 1, device="cuda"),
                               denom=torch.zeros(P, 1, device="cuda"))
    add_densification_stats_window(st, grad_l, radii_l)
    ref = dict(max_radii2D=torch.zeros(P), xyz_gradient_accum=torch.zeros(P, 1), denom=torch.zeros(P, 1))
    densification_stats_window(ref["max_radii2D"], ref["xyz_gradient_accum"], ref["denom"], [r.cpu() for r in radii_l],
                               [g.grad.cpu() for g in grad_l])
    assert torch.equal(st.max_radii2D.cpu(), ref["max_radii2D"]) and torch.equal(st.denom.cpu(), ref["denom"])
    assert T.rel_err(st.xyz_gradient_accum, ref["xyz_gradient_accum"]) <= 1e-6
    assert float(st.denom.sum()) > 0
