"""One whole mapping iteration (src/mapper.py:418-562) on a B200, two arms on the same synthetic window:

  reference : the compiled reference rasterizer (oracle/_ref/ref_dgr) driven the way the reference drives it —
              activations inside every render call, torch mapping loss per view, isotropic loss, ONE backward over the
              summed loss, the boolean-index densification loop, torch.optim.Adam for the Gaussians and the keyframes;
  ours      : fused activations once per iteration, this library's rasterizer with one stream per view, the
              mapping loss evaluated inside each render (LOSS=kernel: stand-alone fused kernel), per-view backward, fused densification stats / visibility, GaussianAdam for both optimizers.

    gpurun -- 'python tests/perf/bench_iteration.py'          # prints ms per iteration for each arm
"""
import os
import sys
import types

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "splat-slam_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))
import diff_gaussian_rasterization as ours_mod  # noqa: E402
from diff_gaussian_rasterization import synthetic as syn  # noqa: E402
from diff_gaussian_rasterization.activations import fused_activations  # noqa: E402
from diff_gaussian_rasterization.adam import GaussianAdam  # noqa: E402
from diff_gaussian_rasterization.densify import add_densification_stats_window, occ_aware_visibility  # noqa: E402
from diff_gaussian_rasterization.mapping_loss import attach_loss, fused_mapping_loss  # noqa: E402
from diff_gaussian_rasterization.window import ViewStreams  # noqa: E402
from oracle.densify import densification_stats_window  # noqa: E402
from oracle.mapping_loss import mapping_loss  # noqa: E402

try:
    import ref_dgr as ref_mod
except Exception:
    ref_mod = None

CAM, P, NV = "replica", int(os.environ.get("P", 1_000_000)), 10
dev = torch.device("cuda:0")
W, H = syn.CAMERAS[CAM][:2]


def make_state(adam_cls):
    g = syn.make_gaussians(CAM, P, seed=0, sh_coeffs=1)
    st = types.SimpleNamespace()
    st.xyz = torch.nn.Parameter(g["means3D"].to(dev))
    st.f_dc = torch.nn.Parameter(g["shs"].to(dev))
    st.opacity = torch.nn.Parameter(torch.logit(g["opacities"].clamp(1e-4, 1 - 1e-4)).to(dev))
    st.scaling = torch.nn.Parameter(torch.log(g["scales"]).to(dev))
    st.rotation = torch.nn.Parameter(g["rotations"].to(dev))
    st.optimizer = adam_cls([{"params": [st.xyz], "lr": 1.6e-4, "name": "xyz"}, {"params": [st.f_dc], "lr": 2.5e-3, "name": "f_dc"},
                             {"params": [st.opacity], "lr": 0.05, "name": "opacity"},
                             {"params": [st.scaling], "lr": 1e-3, "name": "scaling"},
                             {"params": [st.rotation], "lr": 1e-3, "name": "rotation"}], lr=0.0, eps=1e-15)
    st.views, kf_params = [], []
    gen = torch.Generator().manual_seed(1)
    for v in range(NV):
        view = syn.make_view(CAM, syn.window_pose(v))
        vp = types.SimpleNamespace(view=view, mats=[m.to(dev) for m in (view.viewmatrix, view.projmatrix, view.projmatrix_raw, view.campos)])
        vp.theta = torch.nn.Parameter(torch.zeros(3, device=dev))
        vp.rho = torch.nn.Parameter(torch.zeros(3, device=dev))
        vp.exposure_a = torch.nn.Parameter(torch.zeros(1, device=dev))
        vp.exposure_b = torch.nn.Parameter(torch.zeros(1, device=dev))
        vp.gt_image = torch.rand(3, H, W, generator=gen).to(dev)
        vp.gt_depth = (1.0 + 3.0 * torch.rand(1, H, W, generator=gen)).to(dev)
        vp.original_image, vp.depth = vp.gt_image, vp.gt_depth[0]      # the names the reference's viewpoints use
        for prm, lr in ((vp.theta, 3e-3), (vp.rho, 1e-3), (vp.exposure_a, 0.01), (vp.exposure_b, 0.01)):
            kf_params.append({"params": [prm], "lr": lr})
        st.views.append(vp)
    st.keyframe_optimizers = adam_cls(kf_params)
    st.max_radii2D = torch.zeros(P, device=dev)
    st.xyz_gradient_accum = torch.zeros(P, 1, device=dev)
    st.denom = torch.zeros(P, 1, device=dev)
    st.bg = torch.zeros(3, device=dev)
    return st


def settings(mod, st, vp):
    v = vp.view
    return mod.GaussianRasterizationSettings(
        image_height=v.image_height, image_width=v.image_width, tanfovx=v.tanfovx, tanfovy=v.tanfovy, bg=st.bg,
        scale_modifier=1.0, viewmatrix=vp.mats[0], projmatrix=vp.mats[1], projmatrix_raw=vp.mats[2], sh_degree=0,
        campos=vp.mats[3], prefiltered=False, debug=False)


def iteration_reference(st, mod, do_step=True):
    loss, vsp, radii_acm, nt_acm = 0.0, [], [], []
    for vp in st.views:
        means2D = torch.zeros_like(st.xyz, requires_grad=True) + 0
        means2D.retain_grad()
        color, radii, depth, opacity, n_touched = mod.GaussianRasterizer(settings(mod, st, vp))(
            means3D=st.xyz, means2D=means2D, shs=st.f_dc, opacities=torch.sigmoid(st.opacity),
            scales=torch.exp(st.scaling), rotations=torch.nn.functional.normalize(st.rotation), theta=vp.theta, rho=vp.rho)
        loss = loss + mapping_loss(color, depth, vp.exposure_a, vp.exposure_b, vp.gt_image, vp.gt_depth, 0.95, 0.01)
        vsp.append(means2D); radii_acm.append(radii); nt_acm.append(n_touched)
    scaling = torch.exp(st.scaling)
    loss = loss + 10 * torch.abs(scaling - scaling.mean(dim=1).view(-1, 1)).mean()
    loss.backward()
    with torch.no_grad():
        vis = {k: (n > 0).long() for k, n in enumerate(nt_acm)}
        densification_stats_window(st.max_radii2D, st.xyz_gradient_accum, st.denom, radii_acm, [v.grad for v in vsp])
        if not do_step:
            return vis
        st.optimizer.step()
        st.optimizer.zero_grad(set_to_none=True)
        st.keyframe_optimizers.step()
        st.keyframe_optimizers.zero_grad(set_to_none=True)
    return vis


streams = ViewStreams(dev, int(os.environ.get("STREAMS", 10)))
LOSS_IN_RENDER = os.environ.get("LOSS", "attached") == "attached"     # LOSS=kernel: the stand-alone fused loss kernel
CFG = {"Training": {"alpha": 0.95, "rgb_boundary_threshold": 0.01, "ssim_loss": False}}


def iteration_ours(st, mod, do_step=True):
    scales, rots, opac = fused_activations(st.scaling, st.rotation, st.opacity)
    # detach the activated tensors so every view's backward stops there; their summed gradients go through the
    # activation backward once, after the last view
    leaves = [t.detach().requires_grad_(True) for t in (scales, rots, opac)]
    vsp, radii_acm, nt_acm = [], [], []
    streams.fork()
    for k, vp in enumerate(st.views):
        with streams.view(k):
            means2D = torch.zeros_like(st.xyz, requires_grad=True)

            def render():
                return mod.GaussianRasterizer(settings(mod, st, vp))(
                    means3D=st.xyz, means2D=means2D, shs=st.f_dc, opacities=leaves[2], scales=leaves[0],
                    rotations=leaves[1], theta=vp.theta, rho=vp.rho)

            if LOSS_IN_RENDER:
                with attach_loss(CFG, vp) as fused:
                    color, radii, depth, opacity, n_touched = render()
                fused.loss.backward()
            else:
                color, radii, depth, opacity, n_touched = render()
                fused_mapping_loss(color, depth, vp.exposure_a, vp.exposure_b, vp.gt_image, vp.gt_depth, 0.95, 0.01).backward()
            vsp.append(means2D); radii_acm.append(radii); nt_acm.append(n_touched)
    streams.join()
    iso = 10 * torch.abs(scales - scales.mean(dim=1).view(-1, 1)).mean()
    torch.autograd.backward([iso, scales, rots, opac], [None, leaves[0].grad, leaves[1].grad, leaves[2].grad])
    vis, _ = occ_aware_visibility(nt_acm, want_n_obs=False)
    add_densification_stats_window(st, vsp, radii_acm)
    if not do_step:
        return vis
    st.optimizer.step()
    st.optimizer.zero_grad(set_to_none=True)
    st.keyframe_optimizers.step()
    st.keyframe_optimizers.zero_grad(set_to_none=True)
    return vis


def timed(fn, st, mod, n=5, warm=2):
    for _ in range(warm):
        fn(st, mod)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn(st, mod)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


res = {}
if ref_mod is not None:
    st = make_state(torch.optim.Adam)
    res["reference"] = timed(iteration_reference, st, ref_mod)
    del st
    torch.cuda.empty_cache()
st = make_state(GaussianAdam)
res["ours"] = timed(iteration_ours, st, ours_mod)
# same iteration? from identical states, compare what reaches the optimizers (gradients) and the densification stats
# (relative to each tensor's largest entry; the first Adam step itself is +-lr * sign(grad), so parameters of
# Gaussians whose gradient is rounding noise around zero are not a meaningful comparison)
if ref_mod is not None:
    a, b = make_state(torch.optim.Adam), make_state(GaussianAdam)
    va = iteration_reference(a, ref_mod, do_step=False)
    vb = iteration_ours(b, ours_mod, do_step=False)
    rel = lambda x, y: float((x - y).abs().max() / y.abs().max().clamp_min(1e-30))
    for n in ("xyz", "f_dc", "opacity", "scaling", "rotation"):
        res["rel_diff_grad_" + n] = rel(getattr(b, n).grad, getattr(a, n).grad)
    res["rel_diff_grad_pose"] = max(rel(vb_.theta.grad, va_.theta.grad) for va_, vb_ in zip(a.views, b.views))
    res["rel_diff_grad_exposure"] = max(rel(vb_.exposure_a.grad, va_.exposure_a.grad) for va_, vb_ in zip(a.views, b.views))
    res["rel_diff_xyz_gradient_accum"] = rel(b.xyz_gradient_accum, a.xyz_gradient_accum)
    res["visibility_equal"] = all(torch.equal(x, y) for x, y in zip(vb, va.values()))
print({k: (round(v, 3) if "diff" not in k and not isinstance(v, bool) else v) for k, v in res.items()},
      f"P={P} views={NV} {W}x{H}; ms per mapping iteration")
