"""GPU (B200): parity of the CUDA path, called through the public API -> ctypes -> C ABI.

Three checkers, in order of strength:
  1. the compiled reference extension itself (oracle/_ref/ref_dgr) on identical inputs in the same process:
     sort keys, point list, radii, tile ranges, n_touched, n_contrib, final_T and the per-Gaussian geometry
     BIT-EXACT; color/depth/opacity <= 1e-5 abs; gradients <= 1e-3 relative;
  2. the committed golden vectors (outputs of that reference on a B200), same bars — works when ref_dgr is absent;
  3. the CPU oracle (1e-5-class agreement up to borderline flips), plus size-independent properties at full size.
"""
import glob
import os

import pytest
import torch

import parity as T
from test_oracle_golden import GOLDEN, case_from_fixture

pytestmark = pytest.mark.gpu

CASES = {
    "tiny_sh0": dict(camera="tiny", P=300),
    "single": dict(camera="tiny", P=1, seed=4, near_frac=0.0),
    "ragged_sh1_bg": dict(camera="ragged", P=1500, sh_coeffs=4, sh_degree=1, bg=(0.2, 0.5, 0.7), view_k=2),
    "ragged_sh2": dict(camera="ragged", P=900, sh_coeffs=9, sh_degree=2, view_k=4, seed=2),
    "small_sh3": dict(camera="small", P=20000, sh_coeffs=16, sh_degree=3, view_k=1),
    "small_M16_D0": dict(camera="small", P=3000, sh_coeffs=16, sh_degree=0, view_k=5),
    # padded SH layout with more rows than the per-Gaussian backward can stage in shared memory (M > 21)
    "small_M25_D3": dict(camera="small", P=3000, sh_coeffs=25, sh_degree=3, view_k=3),
    "small_precomp": dict(camera="small", P=20000, use_colors_precomp=True, use_cov3d_precomp=True, view_k=3),
    "small_colors_only": dict(camera="small", P=4000, use_colors_precomp=True, view_k=6),
    "big_splats": dict(camera="small", P=5000, scale_modifier=1.7, sigma_px=6.0),
    "huge_splats": dict(camera="ragged", P=64, sigma_px=200.0, seed=8),
    "tum_200k": dict(camera="tum", P=200000, view_k=1),
}


def _make(kw):
    kw = dict(kw)
    near = kw.pop("near_frac", None)
    case = T.make_case(**kw)
    if near is not None:
        from diff_gaussian_rasterization import synthetic as syn
        g = syn.make_gaussians(kw["camera"], kw["P"], seed=kw.get("seed", 0), near_frac=near)
        for k in ("means3D", "opacities", "scales", "rotations", "shs"):
            case[k] = g[k]
    return case


needs_ref = pytest.mark.skipif(T.load_ref() is None, reason="oracle/_ref/ref_dgr not built")


@needs_ref
@pytest.mark.parametrize("name", list(CASES))
def test_bit_exact_chain_vs_reference_extension(name):
    case = _make(CASES[name])
    ours = T.ours_forward_with_intermediates(case)
    ref = T.ref_forward_with_intermediates(T.load_ref(), case)
    T.compare_intermediates(ours, ref)
    assert int((~T.bits_equal(ours["final_T"], ref["final_T"])).sum()) == 0
    for k in ("color", "depth", "opacity"):
        assert float((ours[k] - ref[k]).abs().max()) <= T.FWD_ATOL


@needs_ref
@pytest.mark.parametrize("name", list(CASES))
def test_outputs_and_gradients_vs_reference_extension(name):
    case = _make(CASES[name])
    ours = T.run_module(T.load_ours(), case)
    ref = T.run_module(T.load_ref(), case)
    rep = T.compare_outputs(ours, ref, exact_ints=True)
    assert rep


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-3] for p in GOLDEN])
def test_vs_golden_vectors(path):
    fx = torch.load(path, weights_only=False)
    case = case_from_fixture(fx)
    inter = T.ours_forward_with_intermediates(case)
    assert inter["num_rendered"] == fx["num_rendered"]
    for k in ("radii", "n_touched", "keys", "point_list", "n_contrib", "tiles_touched"):
        assert torch.equal(inter[k].cpu().long().reshape(-1), fx[k].long().reshape(-1)), k
    assert torch.equal(inter["ranges"].cpu().long(), fx["ranges"].long())
    vis = fx["radii"] > 0
    for k in ("depths", "means2D", "conic_opacity"):
        assert bool(T.bits_equal(inter[k].cpu()[vis], fx[k][vis]).all()), k
    assert bool(T.bits_equal(inter["final_T"], fx["final_T"]).all())
    ours = T.run_module(T.load_ours(), case)
    T.compare_outputs(ours, fx, exact_ints=True)


@pytest.mark.parametrize("name", ["tiny_sh0", "ragged_sh1_bg", "small_precomp", "big_splats"])
def test_vs_cpu_oracle(name):
    case = _make(CASES[name])
    if case["P"] > 6000:       # keep the CPU side in seconds
        case = _make(dict(CASES[name], P=5000))
    ours = T.run_module(T.load_ours(), case)
    orc = T.run_oracle(case)
    T.compare_vs_cpu_oracle(ours, orc)


FULL_SIZE = {
    # BASELINE.json configs 2/4 (Replica shape), 3 (TUM tracking) and 5 (ScanNet refinement, SH degree 3)
    "replica_1M_sh0": dict(camera="replica", P=1_000_000, view_k=4),
    "tum_200k_view7": dict(camera="tum", P=200_000, view_k=7),
    "scannet_3M_sh3": dict(camera="scannet", P=3_000_000, sh_coeffs=16, sh_degree=3, view_k=2),
    # many instances per Gaussian: R = 15 M (tile lists up to 6.6 k) and R = 53 M (tile lists up to 22 k)
    "replica_1M_scale3": dict(camera="replica", P=1_000_000, view_k=3, scale_modifier=3.0),
    "replica_2M_big_splats": dict(camera="replica", P=2_000_000, view_k=5, sigma_px=8.0),
}


@needs_ref
@pytest.mark.parametrize("name", list(FULL_SIZE))
def test_full_size_bit_exact_and_gradients_vs_reference_extension(name):
    """The BASELINE.json shapes at full size against the live reference extension."""
    case = T.make_case(**FULL_SIZE[name])
    ours = T.ours_forward_with_intermediates(case)
    ref = T.ref_forward_with_intermediates(T.load_ref(), case)
    T.compare_intermediates(ours, ref)
    del ours, ref
    torch.cuda.empty_cache()
    a = T.run_module(T.load_ours(), case)
    b = T.run_module(T.load_ref(), case)
    T.compare_outputs(a, b, exact_ints=True)


def _stress_case(kind):
    case = T.make_case(camera="small", P=30000, sh_coeffs=4, sh_degree=1, view_k=2, seed=21)
    g = torch.Generator().manual_seed(77)
    if kind == "needles":          # anisotropy up to ~1e4, un-normalised quaternions
        case["scales"] = case["scales"] * torch.tensor([40.0, 1.0, 0.05])
        case["rotations"] = case["rotations"] * (0.5 + torch.rand(30000, 1, generator=g))
    elif kind == "faint":          # opacities straddling the 1/255 cut
        case["opacities"] = (1.0 / 255.0 - 0.001 + 0.004 * torch.rand(30000, 1, generator=g)).contiguous()
    elif kind == "opaque_big":     # early saturation: the T < 1e-4 stop and the 0.99 clamp everywhere
        case["opacities"] = torch.full((30000, 1), 0.999)
        case["scales"] = case["scales"] * 6.0
    elif kind == "far_outside":    # centres far outside the image but large enough to reach into it
        case["means3D"] = case["means3D"] * torch.tensor([6.0, 6.0, 1.0])
        case["scales"] = case["scales"] * 60.0
    return case


@needs_ref
@pytest.mark.parametrize("kind", ["needles", "faint", "opaque_big", "far_outside"])
def test_stress_footprint_binning_stays_bit_exact(kind):
    """The per-warp footprint boxes must never drop a pixel the reference blends: extreme anisotropy, opacities at
    the 1/255 cut, saturating splats and off-screen centres, all bit-exact against the reference extension."""
    case = _stress_case(kind)
    ours = T.ours_forward_with_intermediates(case)
    ref = T.ref_forward_with_intermediates(T.load_ref(), case)
    T.compare_intermediates(ours, ref)
    assert int((~T.bits_equal(ours["final_T"], ref["final_T"])).sum()) == 0
    for k in ("color", "depth", "opacity"):
        assert float((ours[k] - ref[k]).abs().max()) <= T.FWD_ATOL
    a = T.run_module(T.load_ours(), case)
    b = T.run_module(T.load_ref(), case)
    if kind != "needles":
        T.compare_outputs(a, b, exact_ints=True)
        return
    # Needle splats (anisotropy ~1e4, radius ~1000 px): the conic -> covariance gradient is a difference of terms
    # ~1e6 times larger than the result, so in float32 the reference's own numbers for those few Gaussians are
    # rounding noise (two runs of the reference differ by 3-20 % there, and a float32 CPU evaluation of the same
    # formulas differs from both). What must hold: everything the render backward produces agrees to 1e-3, and the
    # per-Gaussian chain agrees for all but a handful of ill-conditioned Gaussians.
    for k in ("means2D", "shs", "opacities"):
        assert T.rel_err(a["grad_" + k], b["grad_" + k]) <= T.GRAD_RTOL, k
    # The data behind that relaxation, measured here and printed into the log: the reference against ITSELF (a second
    # run on the same inputs; float atomics reorder its sums) on the same three tensors.
    b2 = T.run_module(T.load_ref(), case)
    for k in ("means3D", "scales", "rotations"):
        ga, gb, gb2 = a["grad_" + k].cpu(), b["grad_" + k].cpu(), b2["grad_" + k].cpu()
        scale = T.GRAD_RTOL * gb.abs().max()
        bad = ((ga - gb).abs().amax(1) > scale).float().mean()
        ref_bad = ((gb2 - gb).abs().amax(1) > scale).float().mean()
        worst_rows = (gb2 - gb).abs().amax(1).topk(min(8, gb.shape[0])).indices
        ref_rel = ((gb2 - gb).abs().amax(1)[worst_rows] / gb.abs().amax(1)[worst_rows].clamp_min(1e-30)).max()
        print(f"needles grad_{k}: ours-vs-ref rows beyond 1e-3 of max: {float(bad):.5f}; "
              f"ref-vs-ref rows beyond: {float(ref_bad):.5f}; ref-vs-ref worst per-row relative spread: {float(ref_rel):.3f}")
        # the reference-vs-reference figure is itself random from run to run (3e-5 ... 3e-4 of the rows observed), so it
        # is reported, not asserted against
        assert float(bad) <= 0.005, (k, float(bad))


# ---- edge cases -------------------------------------------------------------------------------------------
def test_zero_gaussians():
    m = T.load_ours()
    case = T.make_case(camera="tiny", P=4)
    S = T.settings_for(m, dict(case, bg=torch.tensor([0.3, 0.6, 0.9])), "cuda")
    z3 = torch.zeros(0, 3, device="cuda", requires_grad=True)
    color, radii, depth, opacity, n_touched = m.GaussianRasterizer(S)(
        means3D=z3, means2D=torch.zeros(0, 3, device="cuda"), opacities=torch.zeros(0, 1, device="cuda"),
        shs=torch.zeros(0, 1, 3, device="cuda"), scales=torch.zeros(0, 3, device="cuda"),
        rotations=torch.zeros(0, 4, device="cuda"))
    assert radii.numel() == 0 and n_touched.numel() == 0
    assert torch.allclose(color, torch.tensor([0.3, 0.6, 0.9], device="cuda")[:, None, None].expand_as(color))
    assert float(depth.abs().max()) == 0 and float(opacity.abs().max()) == 0
    color.sum().backward()
    assert z3.grad.shape == (0, 3)


def test_everything_culled():
    """All Gaussians behind the near plane: num_rendered = 0, background image, zero gradients."""
    m = T.load_ours()
    case = T.make_case(camera="tiny", P=200, bg=(0.5, 0.25, 0.125))
    case["means3D"] = case["means3D"].clone()
    case["means3D"][:, 2] = -case["means3D"][:, 2].abs() - 0.01
    out = T.run_module(m, case)
    assert int(out["radii"].abs().sum()) == 0 and int(out["n_touched"].sum()) == 0
    assert float((out["color"] - case["bg"].cuda()[:, None, None]).abs().max()) == 0
    for k in ("means3D", "shs", "opacities", "scales", "rotations", "theta", "rho"):
        assert float(out["grad_" + k].abs().max()) == 0, k


def test_near_plane_is_0p001():
    """As-installed near plane (Dockerfile:20): z = 0.0011 is kept, z = 0.001 is culled (p_view.z <= 0.001f)."""
    m = T.load_ours()
    case = T.make_case(camera="tiny", P=2)
    S = T.settings_for(m, case, "cuda")
    pts = torch.tensor([[0.0, 0.0, 0.0011], [0.0, 0.0, 0.001], [0.0, 0.0, 0.19], [0.0, 0.0, -1.0]], device="cuda")
    vis = m.GaussianRasterizer(S).markVisible(pts)
    assert vis.tolist() == [True, False, True, False]


@needs_ref
def test_mark_visible_vs_reference():
    m, ref = T.load_ours(), T.load_ref()
    case = T.make_case(camera="small", P=50000, view_k=2)
    pts = case["means3D"].cuda()
    a = m.GaussianRasterizer(T.settings_for(m, case, "cuda")).markVisible(pts)
    b = ref.GaussianRasterizer(T.settings_for(ref, case, "cuda")).markVisible(pts)
    assert torch.equal(a, b)


def test_pose_only_gradients():
    """BASELINE config 3 shape: only theta/rho require grad (MonoGS-style tracking)."""
    m = T.load_ours()
    case = T.make_case(camera="small", P=3000, view_k=1)
    S = T.settings_for(m, case, "cuda")
    theta = torch.zeros(3, device="cuda", requires_grad=True)
    rho = torch.zeros(3, device="cuda", requires_grad=True)
    color, radii, depth, opacity, n_touched = m.GaussianRasterizer(S)(
        means3D=case["means3D"].cuda(), means2D=torch.zeros(3000, 3, device="cuda"), opacities=case["opacities"].cuda(),
        shs=case["shs"].cuda(), scales=case["scales"].cuda(), rotations=case["rotations"].cuda(), theta=theta, rho=rho)
    ((color * case["Wc"].cuda()).sum() + (depth * case["Wd"].cuda()).sum()).backward()
    full = T.run_module(m, case)
    assert T.rel_err(theta.grad, full["grad_theta"]) < 1e-4 and T.rel_err(rho.grad, full["grad_rho"]) < 1e-4
    assert float(theta.grad.abs().max()) > 0


def test_strided_and_expanded_inputs_are_read_densely_in_both_passes():
    """Isotropic scales given as `s.expand(-1, 3)` and an SH tensor that is a strided slice: forward AND backward must see
    the dense values. Checked against dense copies of the same inputs, run through this library and through the compiled
    reference. (The reference itself only handles strided inputs in its forward: its backward takes `scales.data_ptr()`
    and `rotations.data_ptr()` without `.contiguous()`, DGR/rasterize_points.cu:180-182, and reads an expanded tensor
    out of bounds. That defect is not reproduced.)"""
    m = T.load_ours()
    case = T.make_case(camera="small", P=6000, sh_coeffs=4, sh_degree=1, view_k=2)
    dev = "cuda"
    iso = case["scales"][:, :1].contiguous()
    sh_wide = torch.zeros(6000, 8, 3)
    sh_wide[:, ::2] = case["shs"]

    def run(mod, dense):
        leaves = dict(means3D=case["means3D"].to(dev).requires_grad_(True), opacities=case["opacities"].to(dev).requires_grad_(True),
                      rotations=case["rotations"].to(dev).requires_grad_(True), iso=iso.to(dev).requires_grad_(True),
                      shw=sh_wide.to(dev).requires_grad_(True))
        scales = leaves["iso"].expand(-1, 3)
        shs = leaves["shw"][:, ::2]
        if dense:
            scales, shs = scales.contiguous(), shs.contiguous()
        else:
            assert not scales.is_contiguous() and not shs.is_contiguous()
        color, radii, depth, opacity, n_touched = mod.GaussianRasterizer(T.settings_for(mod, case, dev))(
            means3D=leaves["means3D"], means2D=torch.zeros(6000, 3, device=dev, requires_grad=True),
            opacities=leaves["opacities"], shs=shs, scales=scales, rotations=leaves["rotations"],
            theta=torch.zeros(3, device=dev, requires_grad=True), rho=torch.zeros(3, device=dev, requires_grad=True))
        ((color * case["Wc"].to(dev)).sum() + (depth * case["Wd"].to(dev)).sum()).backward()
        return color.detach(), {k: v.grad for k, v in leaves.items()}

    c_ours, g_ours = run(m, dense=False)
    checkers = [run(m, dense=True)]
    if T.load_ref() is not None:
        checkers.append(run(T.load_ref(), dense=True))
    for c_ref, g_ref in checkers:
        assert float((c_ours - c_ref).abs().max()) <= T.FWD_ATOL
        for k in g_ref:
            assert T.rel_err(g_ours[k], g_ref[k]) <= T.GRAD_RTOL, k
    assert float(g_ours["iso"].abs().max()) > 0 and float(g_ours["shw"][:, 1::2].abs().max()) == 0.0


def test_debug_mode_checks_every_stage_and_names_the_failing_one(tmp_path, monkeypatch):
    """`debug=True` (DGR/cuda_rasterizer/auxiliary.h:166-173: synchronise and check after EACH stage, report where it
    failed; DGR __init__.py:59-67,126-138: dump the inputs). The failure is injected through the library's test hook,
    which makes a stage report an error without faulting the device."""
    import diff_gaussian_rasterization as m
    from diff_gaussian_rasterization import _cabi
    monkeypatch.chdir(tmp_path)
    case = T.make_case(camera="small", P=2000, view_k=1)
    plain = T.run_module(m, case)
    case_dbg = dict(case)
    S = T.settings_for(m, case, "cuda")._replace(debug=True)
    monkeypatch.setattr(T, "settings_for", lambda mod, c, device="cuda": S)
    _cabi.debug_sync_count(reset=True)
    dbg = T.run_module(m, case_dbg)
    # forward: preprocess, depth sort, key emission, tile sort (+ records, ranges), render; backward: clear, render, per-Gaussian
    assert _cabi.debug_sync_count(reset=True) == 8
    assert not _cabi.set_debug(False), "debug mode must be switched off again after the call"
    for k in ("color", "depth", "radii", "n_touched"):
        assert torch.equal(plain[k], dbg[k]), k
    assert T.rel_err(dbg["grad_means3D"], plain["grad_means3D"]) <= T.GRAD_RTOL
    assert not os.path.exists("snapshot_fw.dump") and not os.path.exists("snapshot_bw.dump")
    # a failing forward stage: named in the exception, inputs dumped
    monkeypatch.setattr(m, "_debug_fail_stage", _cabi.stage_index("tile_sort_records_ranges"))
    with pytest.raises(_cabi.SplatRasterError, match="tile_sort"):
        T.run_module(m, case_dbg)
    dump = torch.load("snapshot_fw.dump", weights_only=False)
    assert torch.equal(dump[0], case["means3D"])
    assert not _cabi.set_debug(False)
    # a failing backward stage
    monkeypatch.setattr(m, "_debug_fail_stage", _cabi.stage_index("render_backward"))
    with pytest.raises(_cabi.SplatRasterError, match="render_backward"):
        T.run_module(m, case_dbg)
    assert os.path.exists("snapshot_bw.dump")
    monkeypatch.setattr(m, "_debug_fail_stage", -1)
    torch.cuda.synchronize()
    again = T.run_module(m, case)      # the library is healthy afterwards
    assert torch.equal(again["color"], plain["color"])


def test_without_the_warp_mask_in_a_fresh_process():
    """Clouds of 2^24 Gaussians or more cannot carry the forward-to-backward warp mask in their id words and use the
    footprint test in the backward instead. SPLAT_NO_WARP_MASK=1 forces that path; it is read once per process, so the
    cases run in a child process."""
    import subprocess
    import sys
    env = dict(os.environ, SPLAT_NO_WARP_MASK="1", SPLAT_BWD_CHUNK="16")      # also the other chunk size of the backward
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-x", "-q", "-m", "gpu", "-k",
                        "outputs_and_gradients_vs_reference_extension or vs_golden_vectors or stress_footprint"],
                       capture_output=True, text=True, env=env, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:]
    assert " passed" in r.stdout
    # chunk 16 WITH the mask
    env = dict(os.environ, SPLAT_BWD_CHUNK="16")
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.abspath(__file__), "-x", "-q", "-m", "gpu", "-k",
                        "outputs_and_gradients_vs_reference_extension or vs_golden_vectors"],
                       capture_output=True, text=True, env=env, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:]


def test_non_default_stream_and_repeatability():
    m = T.load_ours()
    case = T.make_case(camera="small", P=4000, view_k=2)
    a = T.run_module(m, case, with_grad=False)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        b = T.run_module(m, case, with_grad=False)
    s.synchronize()
    for k in ("color", "depth", "opacity", "radii", "n_touched"):
        assert torch.equal(a[k], b[k]), k


# ---- size-independent properties at full BASELINE size --------------------------------------------------
def test_full_size_properties():
    case = T.make_case(camera="replica", P=1_000_000)
    o = T.ours_forward_with_intermediates(case)
    R = o["num_rendered"]
    keys = o["keys"]
    assert R == int(o["tiles_touched"].long().sum())
    assert bool((keys[1:] >= keys[:-1]).all()), "keys sorted"
    tiles = (keys >> 32)
    ranges = o["ranges"].long()
    counts = torch.bincount(tiles, minlength=ranges.shape[0])
    assert torch.equal(ranges[:, 1] - ranges[:, 0], counts), "ranges = per-tile histogram"
    nz = counts > 0
    assert torch.equal(ranges[nz, 0], (torch.cumsum(counts, 0) - counts)[nz])
    # the sort is a permutation of the emitted instances: per-Gaussian multiplicity = tiles_touched
    mult = torch.bincount(o["point_list"].long(), minlength=case["P"])
    assert torch.equal(mult, o["tiles_touched"].long())
    # depth bits in the key equal the Gaussian's stored depth
    d = o["depths"][o["point_list"].long()].view(torch.int32).long() & 0xFFFFFFFF
    assert torch.equal(keys & 0xFFFFFFFF, d)
    # stability: equal keys keep ascending Gaussian id (emission order)
    same = keys[1:] == keys[:-1]
    assert bool((o["point_list"].long()[1:][same] > o["point_list"].long()[:-1][same]).all())
    # compositing invariants
    T_ = o["final_T"]
    assert float(T_.min()) >= 0.0 and float(T_.max()) <= 1.0
    assert torch.allclose(o["opacity"][0], 1.0 - T_, atol=1e-6)
    lens = (ranges[:, 1] - ranges[:, 0])
    H, W = T_.shape
    ty = torch.arange(H, device="cuda") // 16
    tx = torch.arange(W, device="cuda") // 16
    tile_of_pix = ty[:, None] * ((W + 15) // 16) + tx[None, :]
    assert bool((o["n_contrib"].long() <= lens[tile_of_pix]).all())
    assert int(o["n_touched"].min()) >= 0 and int((o["n_touched"][o["radii"] == 0]).sum()) == 0
    assert bool(torch.isfinite(o["color"]).all() and torch.isfinite(o["depth"]).all())


def test_linearity_in_upstream_gradient():
    """Backward is linear in (dL/dcolor, dL/ddepth): grads(2*Wc, 2*Wd) = 2*grads(Wc, Wd)."""
    m = T.load_ours()
    case = T.make_case(camera="small", P=5000, view_k=3)
    a = T.run_module(m, case)
    case2 = dict(case, Wc=2 * case["Wc"], Wd=2 * case["Wd"])
    b = T.run_module(m, case2)
    for k in ("means3D", "shs", "opacities", "scales", "rotations", "theta", "rho"):
        assert T.rel_err(b["grad_" + k], 2 * a["grad_" + k]) < 1e-4, k


@pytest.mark.gpu
def test_window_step_on_several_streams_matches_one_stream():
    """window.ViewStreams issues the views of a step round-robin on side streams; the accumulated Gaussian gradients and
    the per-view pose gradients must be those of the plain sequential loop (up to the backward's atomic-add order)."""
    from diff_gaussian_rasterization import synthetic as syn
    from diff_gaussian_rasterization.window import FlatGrads, ViewStreams, window_step
    m = T.load_ours()
    cam, P, n_views = "small", 20_000, 7
    g = syn.make_gaussians(cam, P, seed=3, sh_coeffs=1)
    Wc, Wd = (t.cuda() for t in syn.upstream_grads(cam))
    res = []
    for n_streams in (1, 3):
        params = {k: g[k].cuda().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
        flat = FlatGrads(params)
        theta = [torch.zeros(3, device="cuda", requires_grad=True) for _ in range(n_views)]
        rho = [torch.zeros(3, device="cuda", requires_grad=True) for _ in range(n_views)]
        means2D = torch.zeros(P, 3, device="cuda", requires_grad=True)

        def render_view(v):
            view = syn.make_view(cam, syn.window_pose(v))
            S = m.GaussianRasterizationSettings(
                image_height=view.image_height, image_width=view.image_width, tanfovx=view.tanfovx, tanfovy=view.tanfovy,
                bg=torch.zeros(3, device="cuda"), scale_modifier=1.0, viewmatrix=view.viewmatrix.cuda(),
                projmatrix=view.projmatrix.cuda(), projmatrix_raw=view.projmatrix_raw.cuda(), sh_degree=0,
                campos=view.campos.cuda(), prefiltered=False, debug=False)
            color, radii, depth, opacity, n_touched = m.GaussianRasterizer(S)(
                means3D=params["means3D"], means2D=means2D, opacities=params["opacities"], shs=params["shs"],
                scales=params["scales"], rotations=params["rotations"], theta=theta[v], rho=rho[v])
            return color, depth

        for _ in range(2):          # second pass re-uses the cached allocator blocks across streams
            window_step(render_view, lambda v: (Wc, Wd), flat, n_views,
                        streams=ViewStreams("cuda:0", n_streams, priority_lanes=True))
            torch.cuda.synchronize()
        res.append((flat.flat.clone(), torch.stack([t.grad for t in theta]), torch.stack([r.grad for r in rho]),
                    means2D.grad.clone()))
    for a, b in zip(res[1], res[0]):
        assert T.rel_err(a, b) <= 1e-4


@needs_ref
def test_priority_lanes_keep_the_chain_bit_exact():
    """With the short kernels on the high-priority side stream the forward is the same computation, event-ordered."""
    from diff_gaussian_rasterization import _cabi
    was = _cabi.priority_lanes(True)
    try:
        assert _cabi.priority_lanes(None) is True
        for name in ("ragged_sh1_bg", "small_sh3", "tum_200k"):
            case = _make(CASES[name])
            with torch.cuda.stream(torch.cuda.Stream()):
                ours = T.ours_forward_with_intermediates(case)
                a = T.run_module(T.load_ours(), case)
            torch.cuda.synchronize()
            ref = T.ref_forward_with_intermediates(T.load_ref(), case)
            T.compare_intermediates(ours, ref)
            T.compare_outputs(a, T.run_module(T.load_ref(), case), exact_ints=True)
    finally:
        _cabi.priority_lanes(was)
