"""CPU: the C-ABI library loads and exports every symbol include/*.h declares (no compute calls without a GPU),
and the host-side mirror of the reference interface keeps its names, argument meaning and error behaviour."""
import ctypes
import glob
import os
import re

import pytest
import torch

import parity as T  # noqa: F401  (sets sys.path)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = []
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        src = open(h).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names += re.findall(r"\b(splat_raster_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def test_header_declares_the_path():
    syms = declared_symbols()
    for must in ("splat_raster_forward_preprocess", "splat_raster_forward_render", "splat_raster_backward",
                 "splat_raster_mark_visible", "splat_raster_geom_bytes", "splat_raster_binning_bytes",
                 "splat_raster_image_bytes"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    from diff_gaussian_rasterization import _cabi
    L = _cabi.lib()
    for s in declared_symbols():
        assert hasattr(L, s), f"{s} declared in include/ but not exported"
    assert set(_cabi.EXPORTS) <= set(declared_symbols())
    assert L.splat_raster_abi_version() == _cabi.ABI_VERSION


def test_size_queries_are_monotone_and_aligned():
    from diff_gaussian_rasterization import _cabi
    L = _cabi.lib()
    prev = 0
    for P in (0, 1, 255, 256, 257, 100_000, 3_000_000):
        b = L.splat_raster_geom_bytes(P)
        assert b % 256 == 0 and b >= prev
        prev = b
    assert L.splat_raster_geom_bytes(-1) == 0
    assert L.splat_raster_image_bytes(1200, 680) >= 1200 * 680 * 8 + 75 * 43 * 8
    assert L.splat_raster_binning_bytes(0) > 0
    assert L.splat_raster_binning_bytes(4_000_000) >= 4_000_000 * (48 + 4 * 4)     # records + the two sort pairs


def test_argument_errors_without_touching_the_gpu():
    from diff_gaussian_rasterization import _cabi
    L = _cabi.lib()
    # NULL state buffer / bad sizes are rejected before any launch
    assert L.splat_raster_forward_preprocess(10, 0, 1, 64, 64, None, None, 1.0, None, None, None, None, None, None, None,
                                             None, 1.0, 1.0, 0, None, None, None, None, None) == -1
    assert L.splat_raster_forward_preprocess(-1, 0, 1, 64, 64, None, None, 1.0, None, None, None, None, None, None, None,
                                             None, 1.0, 1.0, 0, None, None, None, None, None) == -2
    assert L.splat_raster_forward_render(1, 1, 0, 64, *([None] * 11)) == -2
    assert L.splat_raster_mark_visible(-5, None, None, None, None, None) == -2
    assert L.splat_raster_mark_visible(0, None, None, None, None, None) == 0
    assert L.splat_raster_backward(0, 0, 0, 0, 64, 64, *([None] * 5), 1.0, *([None] * 6), 1.0, 1.0, *([None] * 17)) == 0


def _settings(mod, dev="cpu"):
    return mod.GaussianRasterizationSettings(
        image_height=48, image_width=64, tanfovx=0.5, tanfovy=0.4, bg=torch.zeros(3, device=dev), scale_modifier=1.0,
        viewmatrix=torch.eye(4, device=dev), projmatrix=torch.eye(4, device=dev),
        projmatrix_raw=torch.eye(4, device=dev), sh_degree=0, campos=torch.zeros(3, device=dev), prefiltered=False,
        debug=False)


def test_public_names_match_the_reference_package():
    import diff_gaussian_rasterization as m
    assert m.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "projmatrix_raw", "sh_degree", "campos", "prefiltered", "debug")
    import inspect
    sig = inspect.signature(m.GaussianRasterizer.forward)
    assert list(sig.parameters)[1:] == ["means3D", "means2D", "opacities", "shs", "colors_precomp", "scales",
                                        "rotations", "cov3D_precomp", "theta", "rho"]
    assert callable(m.rasterize_gaussians) and hasattr(m.GaussianRasterizer, "markVisible")
    assert issubclass(m._RasterizeGaussians, torch.autograd.Function)


def test_argument_combination_errors_like_the_reference():
    import diff_gaussian_rasterization as m
    r = m.GaussianRasterizer(_settings(m))
    x = torch.zeros(4, 3)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(means3D=x, means2D=x, opacities=torch.ones(4, 1), scales=x, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(means3D=x, means2D=x, opacities=torch.ones(4, 1), shs=torch.zeros(4, 1, 3), colors_precomp=x, scales=x,
          rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=x, means2D=x, opacities=torch.ones(4, 1), colors_precomp=x)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=x, means2D=x, opacities=torch.ones(4, 1), colors_precomp=x, scales=x, rotations=torch.zeros(4, 4),
          cov3D_precomp=torch.zeros(4, 6))


def test_no_cpu_fallback():
    """A CPU tensor must fail loudly instead of silently taking another path."""
    import diff_gaussian_rasterization as m
    from diff_gaussian_rasterization._cabi import SplatRasterError
    r = m.GaussianRasterizer(_settings(m))
    x = torch.zeros(4, 3)
    with pytest.raises(SplatRasterError, match="CUDA tensor"):
        r(means3D=x, means2D=x, opacities=torch.ones(4, 1), colors_precomp=x, scales=x, rotations=torch.zeros(4, 4))
    with pytest.raises(RuntimeError, match="num_points, 3"):
        r(means3D=torch.zeros(4, 2), means2D=x, opacities=torch.ones(4, 1), colors_precomp=x, scales=x,
          rotations=torch.zeros(4, 4))


def test_product_path_never_imports_the_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may touch oracle/."""
    pkg = os.path.join(ROOT, "splat-slam_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                path = os.path.join(dirpath, f)
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), path
                assert not re.search(r"[\"']oracle[\"'/]", src), path
                assert "torch_rasterizer" not in src, path


def test_view_partition():
    from diff_gaussian_rasterization.window import partition_views
    for world in (1, 2, 4, 8):
        parts = [partition_views(10, world, r) for r in range(world)]
        assert sorted(sum(parts, [])) == list(range(10))
        assert max(len(p) for p in parts) == -(-10 // world)
