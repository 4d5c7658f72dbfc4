"""ctypes binding of the C ABI in include/splat_raster.h (libsplat_raster.so, hand-written sm_100a CUDA).

This replaces the pybind module `_C` of the reference extension (DGR/ext.cpp:15-19): the host side only
allocates torch tensors, picks the device/stream and forwards raw pointers.  There is no CPU fallback: if the
shared library is missing or a CUDA tensor is not supplied, calls fail loudly.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_float, c_int, c_size_t, c_void_p

import torch

_LIB_NAME = "libsplat_raster.so"
_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), _LIB_NAME)
ABI_VERSION = 1

_lib = None


class SplatRasterError(RuntimeError):
    pass


class _Views(ctypes.Structure):
    _fields_ = [(n, c_void_p) for n in ("depths", "means2D", "conic_opacity", "rgb", "cov3D", "tiles_touched", "keys",
                                        "point_list", "ranges", "final_T", "n_contrib")]


class _LossDesc(ctypes.Structure):
    """splat_raster_loss_desc (include/splat_raster.h)."""
    _fields_ = [("gt_image", c_void_p), ("gt_depth", c_void_p), ("exposure_a", c_void_p), ("exposure_b", c_void_p),
                ("alpha", c_float), ("rgb_boundary_threshold", c_float), ("initialization", c_int),
                ("workspace", c_void_p), ("out", c_void_p), ("dL_dimage", c_void_p), ("dL_ddepth", c_void_p)]


EXPORTS = (
    "splat_raster_abi_version", "splat_raster_geom_bytes", "splat_raster_image_bytes", "splat_raster_binning_bytes",
    "splat_raster_forward_preprocess", "splat_raster_forward_wait", "splat_raster_forward_render", "splat_raster_backward",
    "splat_raster_mark_visible", "splat_raster_debug_views", "splat_raster_launch_count", "splat_raster_profile",
    "splat_raster_stage_count", "splat_raster_stage_name", "splat_raster_profile_read",
    "splat_raster_mapping_loss_blocks", "splat_raster_mapping_loss", "splat_raster_adam_step",
    "splat_raster_densify_stats", "splat_raster_visibility_count",
    "splat_raster_knn_scratch_bytes", "splat_raster_knn_mean_dist2",
    "splat_raster_deform_points", "splat_raster_activations_forward", "splat_raster_activations_backward", "splat_raster_priority_lanes",
    "splat_raster_backward_noalloc", "splat_raster_set_debug", "splat_raster_last_failed_stage",
    "splat_raster_debug_sync_count", "splat_raster_allreduce_peer", "splat_raster_allreduce_multimem",
    "splat_raster_keep_sort_output", "splat_raster_forward_render_loss", "splat_raster_loss_partials",
)


def lib() -> ctypes.CDLL:
    """Load libsplat_raster.so (built by `__graft_entry__.build()` / `make -C splat-slam_b200/csrc`)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise SplatRasterError(
            f"{_LIB_PATH} not found: build the CUDA library first (python -c 'import __graft_entry__ as g; g.build()'). "
            "There is no CPU fallback for the rasterizer.")
    L = ctypes.CDLL(_LIB_PATH)
    L.splat_raster_abi_version.restype = c_int
    for fn in ("splat_raster_geom_bytes", "splat_raster_image_bytes", "splat_raster_binning_bytes"):
        getattr(L, fn).restype = c_size_t
    L.splat_raster_geom_bytes.argtypes = [c_int]
    L.splat_raster_image_bytes.argtypes = [c_int, c_int]
    L.splat_raster_binning_bytes.argtypes = [c_int]
    P = c_void_p
    L.splat_raster_forward_preprocess.restype = c_int
    L.splat_raster_forward_preprocess.argtypes = [c_int] * 5 + [P, P, c_float, P, P, P, P, P, P, P, P, c_float, c_float,
                                                                c_int, P, P, P, P, P]
    L.splat_raster_forward_wait.restype = c_int
    L.splat_raster_forward_wait.argtypes = []
    L.splat_raster_forward_render.restype = c_int
    L.splat_raster_forward_render.argtypes = [c_int] * 4 + [P] * 11
    L.splat_raster_forward_render_loss.restype = c_int
    L.splat_raster_forward_render_loss.argtypes = [c_int] * 4 + [P] * 10 + [POINTER(_LossDesc), P]
    L.splat_raster_loss_partials.restype = c_int
    L.splat_raster_loss_partials.argtypes = [c_int, c_int]
    L.splat_raster_backward.restype = c_int
    L.splat_raster_backward.argtypes = [c_int] * 6 + [P, P, P, P, P, c_float, P, P, P, P, P, P, c_float, c_float] + [P] * 17
    L.splat_raster_mark_visible.restype = c_int
    L.splat_raster_mark_visible.argtypes = [c_int, P, P, P, P, P]
    L.splat_raster_debug_views.restype = c_int
    L.splat_raster_debug_views.argtypes = [c_int] * 4 + [P, P, P, POINTER(_Views)]
    L.splat_raster_launch_count.restype = ctypes.c_ulonglong
    L.splat_raster_launch_count.argtypes = [c_int]
    L.splat_raster_profile.restype = c_int
    L.splat_raster_profile.argtypes = [c_int]
    L.splat_raster_stage_count.restype = c_int
    L.splat_raster_stage_name.restype = ctypes.c_char_p
    L.splat_raster_stage_name.argtypes = [c_int]
    L.splat_raster_profile_read.restype = c_int
    L.splat_raster_profile_read.argtypes = [POINTER(c_float), POINTER(c_int), c_int]
    L.splat_raster_set_debug.restype = c_int
    L.splat_raster_set_debug.argtypes = [c_int, c_int]
    L.splat_raster_last_failed_stage.restype = ctypes.c_char_p
    L.splat_raster_last_failed_stage.argtypes = [c_int]
    L.splat_raster_debug_sync_count.restype = ctypes.c_ulonglong
    L.splat_raster_debug_sync_count.argtypes = [c_int]
    L.splat_raster_allreduce_peer.restype = c_int
    L.splat_raster_allreduce_peer.argtypes = [POINTER(c_void_p), c_int, c_int, c_size_t, P]
    L.splat_raster_allreduce_multimem.restype = c_int
    L.splat_raster_allreduce_multimem.argtypes = [P, c_int, c_int, c_size_t, P]
    if L.splat_raster_abi_version() != ABI_VERSION:
        raise SplatRasterError("libsplat_raster.so ABI version mismatch; rebuild")
    _lib = L
    return L


def check(status: int, what: str) -> None:
    if status == 0:
        return
    if status > 0:
        stage = lib().splat_raster_last_failed_stage(1).decode()
        where = f" in stage '{stage}'" if stage else ""
        raise SplatRasterError(f"{what}: CUDA error {status}{where} "
                               f"({torch.cuda.get_device_name() if torch.cuda.is_available() else 'no device'})")
    raise SplatRasterError(f"{what}: invalid argument (code {status})")


def ptr(t) -> int | None:
    """Device pointer of a tensor, or None (NULL) for absent/empty inputs — the reference's convention of passing
    empty tensors for unused optionals (DGR/diff_gaussian_rasterization/__init__.py:214-228)."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def require_cuda(name: str, t: torch.Tensor) -> torch.Tensor:
    if not t.is_cuda:
        raise SplatRasterError(f"{name} must be a CUDA tensor (got {t.device}); the rasterizer has no CPU path")
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def stream_ptr(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def keep_sort_output(enable: bool | None = None) -> bool:
    """Make forwards materialise the sorted 64-bit keys and the point list for `debug_views` (tests only)."""
    L = lib()
    L.splat_raster_keep_sort_output.restype = c_int
    L.splat_raster_keep_sort_output.argtypes = [c_int]
    return bool(L.splat_raster_keep_sort_output(-1 if enable is None else int(bool(enable))))


def debug_views(P, R, W, H, geom, binning, image):
    """Tensor copies of the private intermediate arrays (tests only; keys / point_list need `keep_sort_output(True)`
    during the forward)."""
    v = _Views()
    check(lib().splat_raster_debug_views(P, R, W, H, geom.data_ptr(), binning.data_ptr(), image.data_ptr(),
                                         ctypes.byref(v)), "debug_views")
    dev = geom.device
    tiles = ((W + 15) // 16) * ((H + 15) // 16)

    def grab(base_t, addr, count, dtype):
        nbytes = count * torch.empty((), dtype=dtype).element_size()
        off = addr - base_t.data_ptr()
        return base_t[off:off + nbytes].clone().view(dtype)

    out = dict(
        depths=grab(geom, v.depths, P, torch.float32),
        means2D=grab(geom, v.means2D, 2 * P, torch.float32).view(P, 2),
        conic_opacity=grab(geom, v.conic_opacity, 4 * P, torch.float32).view(P, 4),
        rgb=grab(geom, v.rgb, 3 * P, torch.float32).view(P, 3),
        cov3D=grab(geom, v.cov3D, 6 * P, torch.float32).view(P, 6),
        tiles_touched=grab(geom, v.tiles_touched, P, torch.int32),
        keys=grab(binning, v.keys, R, torch.int64),
        point_list=grab(binning, v.point_list, R, torch.int32),
        ranges=grab(image, v.ranges, 2 * tiles, torch.int32).view(tiles, 2),
        final_T=grab(image, v.final_T, H * W, torch.float32).view(H, W),
        n_contrib=grab(image, v.n_contrib, H * W, torch.int32).view(H, W),
    )
    assert dev == out["depths"].device
    return out


def priority_lanes(enable: bool | None = None) -> bool:
    """Run the short kernels of a forward on a high-priority side stream (see include/splat_raster.h). Returns the
    previous setting; None only queries."""
    L = lib()
    L.splat_raster_priority_lanes.restype = c_int
    L.splat_raster_priority_lanes.argtypes = [c_int]
    return bool(L.splat_raster_priority_lanes(-1 if enable is None else int(bool(enable))))


def set_debug(enable: bool, fail_stage: int = -1) -> bool:
    """Per-stage synchronisation and error check, the reference's `debug` (see include/splat_raster.h)."""
    return bool(lib().splat_raster_set_debug(1 if enable else 0, int(fail_stage)))


def stage_index(name: str) -> int:
    L = lib()
    names = [L.splat_raster_stage_name(i).decode() for i in range(L.splat_raster_stage_count())]
    return names.index(name)


def debug_sync_count(reset: bool = False) -> int:
    return int(lib().splat_raster_debug_sync_count(1 if reset else 0))


def launch_count(reset: bool = False) -> int:
    return int(lib().splat_raster_launch_count(1 if reset else 0))


def profile(enable: bool) -> bool:
    return bool(lib().splat_raster_profile(1 if enable else 0))


def profile_read(reset: bool = True):
    """{stage name: (total ms, calls)} measured with CUDA events on the launching stream."""
    L = lib()
    n = L.splat_raster_stage_count()
    ms = (c_float * n)()
    calls = (c_int * n)()
    check(L.splat_raster_profile_read(ms, calls, 1 if reset else 0), "profile_read")
    return {L.splat_raster_stage_name(i).decode(): (float(ms[i]), int(calls[i])) for i in range(n)}
