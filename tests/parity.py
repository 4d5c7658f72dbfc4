"""Shared harness for parity tests: build a synthetic case, run an implementation (this repo's CUDA path, the
compiled reference `ref_dgr`, or the CPU oracle) through the SAME public API, and compare.

Tolerances are the acceptance bars of BASELINE.md §5 / BASELINE.json north_star:
  sort keys, radii, tile ranges, n_touched (+ point_list, n_contrib, num_rendered): bit-exact
  color / depth / opacity: <= 1e-5 abs;   gradients: <= 1e-3 relative.
"""
from __future__ import annotations

import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "splat-slam_b200"), os.path.join(ROOT, "oracle", "_ref")):
    if p not in sys.path:
        sys.path.insert(0, p)

from diff_gaussian_rasterization import synthetic as syn  # noqa: E402

FWD_ATOL = 1e-5
GRAD_RTOL = 1e-3


def load_ours():
    import diff_gaussian_rasterization as m
    return m


def load_ref():
    """The compiled reference extension (oracle/_ref/ref_dgr), or None when it was not built."""
    try:
        import ref_dgr
        return ref_dgr
    except Exception:
        return None


def make_case(camera="small", P=2000, seed=0, sh_coeffs=1, sh_degree=0, view_k=0, bg=(0.0, 0.0, 0.0),
              use_colors_precomp=False, use_cov3d_precomp=False, scale_modifier=1.0, sigma_px=2.0):
    g = syn.make_gaussians(camera, P, seed=seed, sh_coeffs=sh_coeffs, sigma_px=sigma_px)
    v = syn.make_view(camera, syn.window_pose(view_k))
    Wc, Wd = syn.upstream_grads(camera)
    case = dict(camera=camera, P=P, view=v, bg=torch.tensor(bg, dtype=torch.float32), sh_degree=sh_degree,
                scale_modifier=scale_modifier, Wc=Wc, Wd=Wd, means3D=g["means3D"], opacities=g["opacities"],
                shs=None if use_colors_precomp else g["shs"],
                colors_precomp=g["colors_precomp"] if use_colors_precomp else None,
                scales=None if use_cov3d_precomp else g["scales"],
                rotations=None if use_cov3d_precomp else g["rotations"],
                cov3D_precomp=syn.cov3d_from_scale_rot(g["scales"], g["rotations"], scale_modifier)
                if use_cov3d_precomp else None)
    return case


def settings_for(mod, case, device):
    v = case["view"]
    return mod.GaussianRasterizationSettings(
        image_height=v.image_height, image_width=v.image_width, tanfovx=v.tanfovx, tanfovy=v.tanfovy,
        bg=case["bg"].to(device), scale_modifier=case["scale_modifier"], viewmatrix=v.viewmatrix.to(device),
        projmatrix=v.projmatrix.to(device), projmatrix_raw=v.projmatrix_raw.to(device), sh_degree=case["sh_degree"],
        campos=v.campos.to(device), prefiltered=False, debug=False)


GRAD_NAMES = ("means3D", "means2D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp",
              "theta", "rho")


def run_module(mod, case, device="cuda", with_grad=True):
    """Run forward (+ backward with the fixed upstream weights) through the package's public API."""
    S = settings_for(mod, case, device)
    leaf = {}
    for k in ("means3D", "opacities", "shs", "colors_precomp", "scales", "rotations", "cov3D_precomp"):
        t = case[k]
        leaf[k] = None if t is None else t.to(device).clone().requires_grad_(with_grad)
    leaf["means2D"] = torch.zeros_like(leaf["means3D"]).requires_grad_(with_grad)
    leaf["theta"] = torch.zeros(3, device=device).requires_grad_(with_grad)
    leaf["rho"] = torch.zeros(3, device=device).requires_grad_(with_grad)
    rast = mod.GaussianRasterizer(raster_settings=S)
    color, radii, depth, opacity, n_touched = rast(
        means3D=leaf["means3D"], means2D=leaf["means2D"], opacities=leaf["opacities"], shs=leaf["shs"],
        colors_precomp=leaf["colors_precomp"], scales=leaf["scales"], rotations=leaf["rotations"],
        cov3D_precomp=leaf["cov3D_precomp"], theta=leaf["theta"], rho=leaf["rho"])
    out = dict(color=color.detach(), radii=radii, depth=depth.detach(), opacity=opacity.detach(), n_touched=n_touched)
    if with_grad:
        loss = (color * case["Wc"].to(device)).sum() + (depth * case["Wd"].to(device)).sum()
        loss.backward()
        for k in GRAD_NAMES:
            t = leaf[k]
            out["grad_" + k] = None if (t is None or t.grad is None) else t.grad.detach()
    return out


def run_oracle(case, with_grad=True):
    from oracle import torch_rasterizer as orc
    v = case["view"]
    S = orc.Settings(v.image_height, v.image_width, v.tanfovx, v.tanfovy, case["bg"], case["scale_modifier"],
                     v.viewmatrix, v.projmatrix, v.projmatrix_raw, case["sh_degree"], v.campos)
    o = orc.rasterize(case["means3D"], case["opacities"], S, shs=case["shs"], colors_precomp=case["colors_precomp"],
                      scales=case["scales"], rotations=case["rotations"], cov3D_precomp=case["cov3D_precomp"],
                      dL_dcolor=case["Wc"] if with_grad else None, dL_ddepth=case["Wd"] if with_grad else None)
    if with_grad:
        ren = {"grad_shs": "grad_sh", "grad_cov3D_precomp": "grad_cov3Ds_precomp"}
        for k in GRAD_NAMES:
            src = ren.get("grad_" + k, "grad_" + k)
            o["grad_" + k] = o.get(src)
        # absent inputs have no gradient
        for k in ("shs", "colors_precomp", "scales", "rotations", "cov3D_precomp"):
            if case[k] is None:
                o["grad_" + k] = None
        o["grad_theta"] = o["grad_theta"].reshape(3)
        o["grad_rho"] = o["grad_rho"].reshape(3)
    return o


def rel_err(a, b):
    """max |a-b| / max(|b|_inf, tiny): the '1e-3 relative' of the acceptance bar, scale = the tensor's own max."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    if a.numel() == 0:
        return 0.0
    scale = max(float(b.abs().max()), 1e-12)
    return float((a - b).abs().max()) / scale


def compare_outputs(ours, ref, exact_ints=True, fwd_atol=FWD_ATOL, grad_rtol=GRAD_RTOL, int_mismatch_frac=0.0):
    """Returns a dict of error figures and raises AssertionError on violations."""
    rep = {}
    for k in ("radii", "n_touched"):
        a, b = ours[k].cpu().long(), ref[k].cpu().long()
        mism = int((a != b).sum())
        rep[k + "_mismatch"] = mism
        if exact_ints:
            assert mism == 0, f"{k}: {mism} of {a.numel()} differ"
        else:
            assert mism <= int_mismatch_frac * a.numel(), f"{k}: {mism} of {a.numel()} differ"
    for k in ("color", "depth", "opacity"):
        err = float((ours[k].cpu() - ref[k].cpu()).abs().max())
        rep[k + "_abs"] = err
        assert err <= fwd_atol, f"{k}: max abs err {err:.3e} > {fwd_atol}"
    for k in GRAD_NAMES:
        ga, gb = ours.get("grad_" + k), ref.get("grad_" + k)
        if ga is None or gb is None:
            continue
        e = rel_err(ga, gb)
        rep["grad_" + k + "_rel"] = e
        assert e <= grad_rtol, f"grad_{k}: rel err {e:.3e} > {grad_rtol}"
    return rep


def compare_vs_cpu_oracle(ours, orc, int_mismatch_frac=0.02, pix_tol=1e-4, bad_pix_frac=0.005, grad_rtol=5e-3):
    """GPU path vs the CPU oracle. CPU float32 has no FMA contraction and another expf, so a few borderline
    Gaussians/pixels take the other branch (cull, 1/255 cut, T < 1e-4 stop); everything else must agree to
    1e-5-class accuracy. Checked as: bounded integer mismatches, bounded fraction of outlier pixels, tiny median
    error, gradients within grad_rtol of the tensor's max."""
    rep = {}
    for k in ("radii", "n_touched"):
        a, b = ours[k].cpu().long(), orc[k].cpu().long()
        mism = int((a != b).sum())
        rep[k + "_mismatch"] = mism
        assert mism <= max(2, int_mismatch_frac * a.numel()), f"{k}: {mism} of {a.numel()} differ"
    for k in ("color", "depth", "opacity"):
        err = (ours[k].cpu() - orc[k].cpu()).abs()
        rep[k + "_median"] = float(err.median())
        rep[k + "_bad_frac"] = float((err > pix_tol).float().mean())
        assert rep[k + "_median"] <= FWD_ATOL, (k, rep)
        assert rep[k + "_bad_frac"] <= bad_pix_frac, (k, rep)
    for k in GRAD_NAMES:
        ga, gb = ours.get("grad_" + k), orc.get("grad_" + k)
        if ga is None or gb is None:
            continue
        e = rel_err(ga.reshape(gb.shape), gb)
        rep["grad_" + k + "_rel"] = e
        assert e <= grad_rtol, f"grad_{k}: rel err {e:.3e} > {grad_rtol}"
    return rep


# ---------------------------------------------------------------------------------------------------------
# intermediates of the reference, carved out of its opaque buffers (SURVEY.md Appendix B;
# DGR/cuda_rasterizer/rasterizer_impl.cu:155-194, rasterizer_impl.h:21-27)
# ---------------------------------------------------------------------------------------------------------
def _carve(buf, cursor, count, dtype):
    esize = torch.empty((), dtype=dtype).element_size()
    base = buf.data_ptr()
    addr = (base + cursor + 127) & ~127
    off = addr - base
    t = buf[off:off + count * esize].clone().view(dtype)
    return t, off + count * esize


def ref_forward_with_intermediates(ref_mod, case, device="cuda"):
    """Calls the reference's `_C.rasterize_gaussians` directly and decodes its three state buffers."""
    S = settings_for(ref_mod, case, device)
    e = torch.Tensor([])

    def opt(t):
        return e if t is None else t.to(device)

    args = (S.bg, case["means3D"].to(device), opt(case["colors_precomp"]), case["opacities"].to(device),
            opt(case["scales"]), opt(case["rotations"]), S.scale_modifier, opt(case["cov3D_precomp"]), S.viewmatrix,
            S.projmatrix, S.projmatrix_raw, S.tanfovx, S.tanfovy, S.image_height, S.image_width, opt(case["shs"]),
            S.sh_degree, S.campos, False, False)
    R, color, radii, geom, binning, img, depth, opacity, n_touched = ref_mod._C.rasterize_gaussians(*args)
    torch.cuda.synchronize()
    P = case["means3D"].shape[0]
    N = S.image_height * S.image_width
    c = 0
    depths, c = _carve(geom, c, P, torch.float32)
    clamped, c = _carve(geom, c, 3 * P, torch.uint8)
    _, c = _carve(geom, c, P, torch.int32)
    means2D, c = _carve(geom, c, 2 * P, torch.float32)
    cov3D, c = _carve(geom, c, 6 * P, torch.float32)
    conic_opacity, c = _carve(geom, c, 4 * P, torch.float32)
    rgb, c = _carve(geom, c, 3 * P, torch.float32)
    tiles_touched, c = _carve(geom, c, P, torch.int32)
    c = 0
    point_list, c = _carve(binning, c, R, torch.int32)
    _, c = _carve(binning, c, R, torch.int32)
    keys, c = _carve(binning, c, R, torch.int64)
    c = 0
    accum_alpha, c = _carve(img, c, N, torch.float32)
    n_contrib, c = _carve(img, c, N, torch.int32)
    ranges, c = _carve(img, c, 2 * N, torch.int32)
    tiles = ((S.image_width + 15) // 16) * ((S.image_height + 15) // 16)
    return dict(num_rendered=R, color=color, radii=radii, depth=depth, opacity=opacity, n_touched=n_touched,
                depths=depths, means2D=means2D.view(P, 2), cov3D=cov3D.view(P, 6),
                conic_opacity=conic_opacity.view(P, 4), rgb=rgb.view(P, 3), tiles_touched=tiles_touched,
                clamped=clamped.view(P, 3), keys=keys, point_list=point_list,
                final_T=accum_alpha.view(S.image_height, S.image_width),
                n_contrib=n_contrib.view(S.image_height, S.image_width), ranges=ranges.view(-1, 2)[:tiles])


def ours_forward_with_intermediates(case, device="cuda"):
    import diff_gaussian_rasterization as m
    from diff_gaussian_rasterization import _cabi
    S = settings_for(m, case, device)
    e = torch.Tensor([])

    def opt(t):
        return e if t is None else t.to(device)

    was = _cabi.keep_sort_output(True)     # the product path does not write the sorted keys / point list
    try:
        R, color, radii, geom, binning, img, depth, opacity, n_touched = m._forward_impl(
            S, case["means3D"].to(device), opt(case["shs"]), opt(case["colors_precomp"]), case["opacities"].to(device),
            opt(case["scales"]), opt(case["rotations"]), opt(case["cov3D_precomp"]))
        torch.cuda.synchronize()
    finally:
        _cabi.keep_sort_output(was)
    P = case["means3D"].shape[0]
    out = _cabi.debug_views(P, R, S.image_width, S.image_height, geom, binning, img)
    out.update(num_rendered=R, color=color, radii=radii, depth=depth, opacity=opacity, n_touched=n_touched)
    return out


def bits_equal(a, b):
    """Bitwise equality for float tensors (treats -0.0 == +0.0 as the reference's arithmetic can differ only there)."""
    return (a.cpu() == b.cpu()) | (torch.isnan(a.cpu()) & torch.isnan(b.cpu()))


def compare_intermediates(ours, ref):
    """Bit-exact comparison of the discrete chain against the reference. Per-Gaussian arrays are compared only
    where radii > 0 (the reference leaves the rest uninitialised)."""
    rep = {}
    assert ours["num_rendered"] == ref["num_rendered"], (ours["num_rendered"], ref["num_rendered"])
    vis = (ref["radii"].cpu() > 0)
    for k in ("radii", "tiles_touched", "n_touched"):
        a, b = ours[k].cpu().long(), ref[k].cpu().long()
        if k == "tiles_touched":
            pass
        mism = int((a != b).sum())
        rep[k] = mism
        assert mism == 0, f"{k}: {mism} mismatches"
    for k in ("depths", "means2D", "conic_opacity"):
        a, b = ours[k].cpu()[vis], ref[k].cpu()[vis]
        mism = int((~bits_equal(a, b)).sum())
        rep[k] = mism
        assert mism == 0, f"{k}: {mism} float bit mismatches among visible Gaussians"
    for k in ("keys", "point_list"):
        mism = int((ours[k].cpu().long() != ref[k].cpu().long()).sum())
        rep[k] = mism
        assert mism == 0, f"{k}: {mism} mismatches"
    mism = int((ours["ranges"].cpu().long() != ref["ranges"].cpu().long()).sum())
    rep["ranges"] = mism
    assert mism == 0, f"ranges: {mism} mismatches"
    mism = int((ours["n_contrib"].cpu().long() != ref["n_contrib"].cpu().long()).sum())
    rep["n_contrib"] = mism
    assert mism == 0, f"n_contrib: {mism} mismatches"
    return rep
