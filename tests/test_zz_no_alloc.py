"""GPU: `splat_raster_backward_noalloc` — the backward entry point for callers that cannot allocate device memory
where they call it (fixed-frame-time tracking loops, hosts that gave all device memory to fixed pools, thin
foreign-language bindings). Its scratch is a fixed 12 MiB array inside the library, shared by all calls on a GPU.

Called through the C ABI (ctypes) and compared with the regular `splat_raster_backward` on the same forward state,
twice in a row: the second call sees the first call's leftovers in the shared array and must still be right.
Tolerance: 1e-3 relative, the gradient bar used everywhere in this suite (float atomics reorder the sums).
The file name sorts last so a `-x` run has already exercised every other test when it gets here."""
import pytest
import torch

import parity as T

pytestmark = pytest.mark.gpu

OUT_NAMES = ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
             "dL_drotations", "dL_dtau_partial")


def _forward_state(case):
    import diff_gaussian_rasterization as m
    S = T.settings_for(m, case, "cuda")
    dev = {k: (None if case[k] is None else case[k].cuda()) for k in
           ("means3D", "opacities", "shs", "colors_precomp", "scales", "rotations", "cov3D_precomp")}
    e = torch.Tensor([])
    R, color, radii, geom, binning, img, depth, opacity, n_touched = m._forward_impl(
        S, dev["means3D"], e if dev["shs"] is None else dev["shs"],
        e if dev["colors_precomp"] is None else dev["colors_precomp"], dev["opacities"],
        e if dev["scales"] is None else dev["scales"], e if dev["rotations"] is None else dev["rotations"],
        e if dev["cov3D_precomp"] is None else dev["cov3D_precomp"])
    return S, dev, R, radii, geom, binning, img


def _outputs(P, M, fill):
    f = dict(dtype=torch.float32, device="cuda")
    shapes = dict(dL_dmeans2D=(P, 3), dL_dcolors=(P, 3), dL_dopacity=(P, 1), dL_dmeans3D=(P, 3), dL_dcov3D=(P, 6),
                  dL_dsh=(P, M, 3), dL_dscales=(P, 3), dL_drotations=(P, 4), dL_dtau_partial=((P + 255) // 256, 6))
    return {k: torch.full(s, fill, **f) for k, s in shapes.items()}


def _common_args(case, S, dev, R, radii, geom, binning, img):
    P = case["P"]
    M = dev["shs"].shape[1] if dev["shs"] is not None else 0
    W, H = S.image_width, S.image_height
    head = (P, S.sh_degree, M, R, W, H, S.bg, dev["means3D"], dev["shs"], dev["colors_precomp"], dev["scales"],
            S.scale_modifier, dev["rotations"], dev["cov3D_precomp"], S.viewmatrix, S.projmatrix, S.projmatrix_raw,
            S.campos, S.tanfovx, S.tanfovy, radii, geom, binning, img, case["Wc"].cuda(), case["Wd"].cuda())
    return P, M, head


def _run_regular(case, state):
    from diff_gaussian_rasterization import _cabi
    P, M, head = _common_args(case, *state)
    out = _outputs(P, M, float("nan"))
    scratch = torch.empty(P * 12, device="cuda")
    p = _cabi.ptr
    args = [a if isinstance(a, (int, float)) else p(a) for a in head]
    rc = _cabi.lib().splat_raster_backward(*args, p(scratch), *[p(out[k]) for k in OUT_NAMES],
                                           torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    torch.cuda.synchronize()
    return out


def _run_noalloc(case, state, fill):
    from diff_gaussian_rasterization import noalloc
    P, M, head = _common_args(case, *state)
    out = _outputs(P, M, fill)
    noalloc.backward_noalloc(*head, *[out[k] for k in OUT_NAMES], torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("kw", [dict(camera="small", P=20000, sh_coeffs=4, sh_degree=1, view_k=1),
                                dict(camera="tum", P=200000, view_k=2),
                                dict(camera="ragged", P=1200, use_colors_precomp=True, use_cov3d_precomp=True, seed=5)],
                         ids=["small_sh1", "tum_200k_tracking_size", "precomp"])
def test_noalloc_backward_matches_regular_twice_in_a_row(kw):
    case = T.make_case(**kw)
    state = _forward_state(case)
    ref = _run_regular(case, state)
    first = _run_noalloc(case, state, fill=float("nan"))
    second = _run_noalloc(case, state, fill=123.0)
    for name, got in (("first", first), ("second", second)):
        for k in OUT_NAMES:
            if ref[k].numel() == 0:
                continue
            if k in ("dL_dscales", "dL_drotations") and case["scales"] is None:
                continue          # not written without (scales, rotations), as for the regular entry point
            assert bool(torch.isfinite(got[k]).all()), (name, k)
            e = T.rel_err(got[k], ref[k])
            assert e <= T.GRAD_RTOL, (name, k, e)
        tau_ref, tau_got = ref["dL_dtau_partial"].sum(0), got["dL_dtau_partial"].sum(0)
        assert T.rel_err(tau_got, tau_ref) <= T.GRAD_RTOL, name


def test_noalloc_after_a_smaller_and_a_larger_call():
    """Leftovers from a LARGER earlier call must not leak into a smaller later one (and vice versa)."""
    big = T.make_case(camera="small", P=30000, view_k=3)
    small = T.make_case(camera="tiny", P=300)
    sb, ss = _forward_state(big), _forward_state(small)
    _run_noalloc(big, sb, fill=0.0)
    got = _run_noalloc(small, ss, fill=float("nan"))
    ref = _run_regular(small, ss)
    for k in OUT_NAMES:
        assert T.rel_err(got[k], ref[k]) <= T.GRAD_RTOL, k
    got = _run_noalloc(big, sb, fill=float("nan"))
    ref = _run_regular(big, sb)
    for k in OUT_NAMES:
        assert T.rel_err(got[k], ref[k]) <= T.GRAD_RTOL, k


def test_noalloc_rejects_clouds_beyond_its_capacity():
    from diff_gaussian_rasterization import noalloc
    case = T.make_case(camera="tiny", P=300)
    state = _forward_state(case)
    P, M, head = _common_args(case, *state)
    out = _outputs(P, M, 7.0)
    too_many = (noalloc.MAX_GAUSSIANS + 1,) + head[1:]
    with pytest.raises(noalloc.SplatRasterCapacityError, match="262144"):
        noalloc.backward_noalloc(*too_many, *[out[k] for k in OUT_NAMES], torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert all(bool((out[k] == 7.0).all()) for k in OUT_NAMES if out[k].numel()), "nothing may be launched"


def test_noalloc_pose_only_call_passes_null_for_the_per_gaussian_outputs():
    """A tracking loop only wants the pose gradient: every per-Gaussian output is NULL, as `splat_raster_backward`
    allows; the pose partials must equal those of the full call."""
    from diff_gaussian_rasterization import noalloc
    case = T.make_case(camera="small", P=5000, view_k=2)
    state = _forward_state(case)
    P, M, head = _common_args(case, *state)
    full = _run_noalloc(case, state, fill=float("nan"))
    tau = torch.full(((P + 255) // 256, 6), float("nan"), device="cuda")
    noalloc.backward_noalloc(*head, None, None, None, None, None, None, None, None, tau,
                             torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert bool(torch.isfinite(tau).all())
    assert T.rel_err(tau.sum(0), full["dL_dtau_partial"].sum(0)) <= T.GRAD_RTOL
