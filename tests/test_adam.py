"""Fused Adam. The algorithm lives in PyTorch (the reference calls torch.optim.Adam,
gaussian_model.py:302, mapper.py:1111), so the oracle is pinned against torch.optim.Adam itself, live, and the CUDA
kernel is checked against both through the C ABI."""
import numpy as np
import pytest
import torch

import parity as T  # noqa: F401  (sys.path)
from oracle.adam import AdamOracle

# the reference's parameter groups (gaussian_model.py:268-300) with configs/splat_slam.yaml-like learning rates
GROUPS = (("xyz", 3, 0.00016 * 6.0), ("f_dc", 3, 0.0025), ("f_rest", 0, 0.0025 / 20.0), ("opacity", 1, 0.05),
          ("scaling", 3, 0.001 * 6.0), ("rotation", 4, 0.001))


def _params(P, seed, device="cpu"):
    g = torch.Generator().manual_seed(seed)
    return {name: torch.randn(P, max(w, 0), generator=g).to(device) for name, w, _ in GROUPS}


def _grads(P, seed, it, device="cpu"):
    g = torch.Generator().manual_seed(1000 * seed + it)
    out = {}
    for name, w, _ in GROUPS:
        scale = 10.0 ** float(torch.randint(-6, 1, (1,), generator=g))        # gradients span many decades
        x = torch.randn(P, w, generator=g) * scale
        x[:: 7] = 0.0                                                        # untouched Gaussians: exact-zero gradients
        out[name] = x.to(device)
    return out


def _make(cls, params, **kw):
    leaves = {k: torch.nn.Parameter(v.clone()) for k, v in params.items()}
    groups = [{"params": [leaves[name]], "lr": lr, "name": name} for name, _, lr in GROUPS]
    return leaves, cls(groups, lr=0.0, eps=1e-15, **kw)


def _close(a, b, tol=3e-6):
    """max |a-b| <= tol * max |b|: lerp / addcmul results can cancel, so rounding is relative to the operands."""
    a = np.asarray(a.detach().cpu() if torch.is_tensor(a) else a, dtype=np.float64)
    b = np.asarray(b.detach().cpu() if torch.is_tensor(b) else b, dtype=np.float64)
    if a.size == 0:
        assert b.size == 0
        return
    assert np.abs(a - b).max() <= tol * max(np.abs(b).max(), 1e-30), (np.abs(a - b).max(), np.abs(b).max())


def test_oracle_matches_torch_adam_cpu():
    P = 257
    params = _params(P, 0)
    leaves, opt = _make(torch.optim.Adam, params)
    orc = AdamOracle({k: v.numpy() for k, v in params.items()}, {n: lr for n, _, lr in GROUPS}, eps=1e-15)
    for it in range(6):
        grads = _grads(P, 0, it)
        for k, p in leaves.items():
            p.grad = grads[k].clone()
        if it == 3:
            leaves["opacity"].grad = None          # a parameter without a gradient skips the step (and its count)
            grads["opacity"] = None
        opt.step()
        orc.step({k: (None if g is None else g.numpy()) for k, g in grads.items()})
        opt.zero_grad(set_to_none=True)
    for k in leaves:
        if leaves[k].numel():
            _close(orc.params[k], leaves[k])
            _close(orc.m[k], opt.state[leaves[k]]["exp_avg"])
            _close(orc.v[k], opt.state[leaves[k]]["exp_avg_sq"])
    assert orc.t["opacity"] == 5 and orc.t["xyz"] == 6


def test_constructor_rejects_what_is_not_built():
    from diff_gaussian_rasterization.adam import GaussianAdam
    p = [torch.nn.Parameter(torch.zeros(3))]
    for kw in (dict(weight_decay=0.1), dict(amsgrad=True), dict(maximize=True), dict(betas=(0.3, 0.999))):
        with pytest.raises(NotImplementedError):
            GaussianAdam(p, **kw)
    with pytest.raises(ValueError):
        GaussianAdam(p, lr=-1.0)
    opt = GaussianAdam(p, lr=0.1)
    p[0].grad = torch.ones(3)
    with pytest.raises(RuntimeError):       # CPU tensors: there is no CPU path
        opt.step()


def _ulp_diff(a, b):
    ai = a.detach().contiguous().view(torch.int32).long()
    bi = b.detach().contiguous().view(torch.int32).long()
    return int((ai - bi).abs().max()) if a.numel() else 0


@pytest.mark.gpu
@pytest.mark.parametrize("P", [1, 2049, 100_003])
def test_fused_adam_vs_torch_and_oracle(P):
    from diff_gaussian_rasterization.adam import GaussianAdam
    params = _params(P, 1)
    ours_leaves, ours = _make(GaussianAdam, {k: v.cuda() for k, v in params.items()})
    ref_leaves, ref = _make(torch.optim.Adam, {k: v.cuda() for k, v in params.items()})
    orc = AdamOracle({k: v.numpy() for k, v in params.items()}, {n: lr for n, _, lr in GROUPS}, eps=1e-15)
    for it in range(5):
        grads = _grads(P, 1, it)
        for k in ours_leaves:
            ours_leaves[k].grad = grads[k].cuda()
            ref_leaves[k].grad = grads[k].cuda()
        ours.step()
        ref.step()
        orc.step({k: g.numpy() for k, g in grads.items()})
        ours.zero_grad(set_to_none=True)
        ref.zero_grad(set_to_none=True)
    worst = 0
    for k in ours_leaves:
        if not ours_leaves[k].numel():
            continue
        so, sr = ours.state[ours_leaves[k]], ref.state[ref_leaves[k]]
        assert float(so["step"]) == float(sr["step"]) == 5.0
        # against torch's own CUDA kernels: tolerance of the acceptance bar, and the distance in ulps for the record
        _close(ours_leaves[k], ref_leaves[k], 1e-6)
        _close(so["exp_avg"], sr["exp_avg"], 1e-6)
        _close(so["exp_avg_sq"], sr["exp_avg_sq"], 1e-6)
        worst = max(worst, _ulp_diff(ours_leaves[k], ref_leaves[k]))
        # against the CPU oracle
        _close(ours_leaves[k], orc.params[k])
    print("max ulp distance to torch.optim.Adam on CUDA:", worst)
    assert worst <= 4


@pytest.mark.gpu
def test_fused_adam_many_small_unaligned_tensors_and_zeroing():
    """The keyframe optimizer (mapper.py:1089-1111): many tiny tensors with their own lr; more than one launch worth of
    segments, odd sizes, 4-byte-aligned views, and the in-kernel gradient clear."""
    from diff_gaussian_rasterization.adam import GaussianAdam
    g = torch.Generator().manual_seed(3)
    flat = torch.zeros(4096, device="cuda")
    groups_o, groups_r, views = [], [], []
    cursor = 1                                  # start off 16-byte alignment on purpose
    for i in range(45):
        n = [3, 3, 1, 1, 7, 130][i % 6]
        init = torch.randn(n, generator=g)
        po = torch.nn.Parameter(init.cuda())
        pr = torch.nn.Parameter(init.cuda())
        lr = [0.003, 0.001, 0.01, 0.01, 0.02, 0.0005][i % 6]
        groups_o.append({"params": [po], "lr": lr, "name": f"p{i}"})
        groups_r.append({"params": [pr], "lr": lr, "name": f"p{i}"})
        views.append(flat[cursor:cursor + n])
        cursor += n
    ours = GaussianAdam(groups_o, zero_grads=True)
    ref = torch.optim.Adam(groups_r)
    from diff_gaussian_rasterization import _cabi
    _cabi.launch_count(reset=True)
    for it in range(3):
        for go, gr, view in zip(groups_o, groups_r, views):
            val = torch.randn(view.numel(), generator=g).cuda()
            view.copy_(val)
            go["params"][0].grad = view               # gradient is a view into one flat buffer (window.FlatGrads)
            gr["params"][0].grad = val.clone()
        ours.step()
        ref.step()
        assert float(flat.abs().max()) == 0.0         # every gradient cleared by the same kernel
    assert _cabi.launch_count() == 3 * 2          # 45 tensors = 2 launches per step
    for go, gr in zip(groups_o, groups_r):
        _close(go["params"][0], gr["params"][0], 1e-6)


@pytest.mark.gpu
def test_fused_adam_survives_the_reference_optimizer_surgery():
    """Densification edits the optimizer by hand (gaussian_model.py:559-590 cat, :519-540 prune): swap the parameter,
    carry the state over. The same code must work on GaussianAdam, and a state_dict must load into torch's Adam."""
    from diff_gaussian_rasterization.adam import GaussianAdam
    P = 1000
    params = {k: v.cuda() for k, v in _params(P, 4).items()}
    res = {}
    for cls in (GaussianAdam, torch.optim.Adam):
        leaves, opt = _make(cls, params)
        for it in range(2):
            for k, gr in _grads(P, 4, it, "cuda").items():
                leaves[k].grad = gr
            opt.step()
            opt.zero_grad(set_to_none=True)
        mask = torch.arange(P, device="cuda") % 3 != 0
        for group in opt.param_groups:           # prune, then extend by 10 rows, as the reference does it
            old = group["params"][0]
            st = opt.state.get(old)
            ext = torch.full((10, old.shape[1]), 0.5, device="cuda")
            st["exp_avg"] = torch.cat((st["exp_avg"][mask], torch.zeros_like(ext)), dim=0)
            st["exp_avg_sq"] = torch.cat((st["exp_avg_sq"][mask], torch.zeros_like(ext)), dim=0)
            del opt.state[old]
            group["params"][0] = torch.nn.Parameter(torch.cat((old[mask], ext), dim=0).requires_grad_(True))
            opt.state[group["params"][0]] = st
        P2 = int(mask.sum()) + 10
        for it in range(2):
            grads = _grads(P2, 5, it, "cuda")
            for group in opt.param_groups:
                group["params"][0].grad = grads[group["name"]]
            opt.step()
            opt.zero_grad(set_to_none=True)
        res[cls.__name__] = ({g["name"]: g["params"][0] for g in opt.param_groups}, opt)
    for k in res["Adam"][0]:
        _close(res["GaussianAdam"][0][k], res["Adam"][0][k], 1e-6)
    res["Adam"][1].load_state_dict(res["GaussianAdam"][1].state_dict())     # state layouts are interchangeable
