"""CPU, world_size 2 over gloo: the view-sharded window step (partition + flat-buffer all-reduce) gives every rank
the same Gaussian gradients as a single process rendering all views. The renderer plugged in here is a small
differentiable stand-in built from the CPU oracle's per-tile compositing, so the test covers the host logic of the
N>1 path without a GPU."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import parity as T  # noqa: F401

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _toy_render(params, k):
    """Differentiable toy 'view k': depends on every leaf parameter, different per view."""
    w = 1.0 + 0.37 * k
    color = (params["means3D"] * w).sum(1, keepdim=True) * params["opacities"] + params["shs"][:, 0, :].sum(1, keepdim=True)
    depth = (params["scales"] * params["rotations"][:, :3]).sum(1, keepdim=True) * w
    return color.sum(0, keepdim=True).expand(3, 1).reshape(3, 1, 1).contiguous() * torch.ones(3, 2, 2), \
        depth.sum().reshape(1, 1, 1) * torch.ones(1, 2, 2)


def _make_params(seed=0, P=64):
    g = torch.Generator().manual_seed(seed)
    return {"means3D": torch.randn(P, 3, generator=g).requires_grad_(True),
            "shs": torch.randn(P, 1, 3, generator=g).requires_grad_(True),
            "opacities": torch.rand(P, 1, generator=g).requires_grad_(True),
            "scales": torch.rand(P, 3, generator=g).requires_grad_(True),
            "rotations": torch.randn(P, 4, generator=g).requires_grad_(True)}


def _upstream(k):
    g = torch.Generator().manual_seed(100 + k)
    return torch.rand(3, 2, 2, generator=g), torch.rand(1, 2, 2, generator=g)


def _worker(rank, world, port, n_views, out_dir):
    sys.path.insert(0, os.path.join(ROOT, "splat-slam_b200"))
    from diff_gaussian_rasterization.window import FlatGrads, window_step
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    params = _make_params()
    flat = FlatGrads(params)
    mine = window_step(lambda v: _toy_render(params, v), _upstream, flat, n_views)
    torch.save({"flat": flat.flat.clone(), "mine": mine}, os.path.join(out_dir, f"r{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.parametrize("n_views", [10, 3])
def test_window_step_world2_matches_single_process(tmp_path, n_views):
    from diff_gaussian_rasterization.window import FlatGrads, window_step
    world = 2
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n_views, str(tmp_path)), nprocs=world, join=True)
    r0 = torch.load(tmp_path / "r0.pt")
    r1 = torch.load(tmp_path / "r1.pt")
    assert sorted(r0["mine"] + r1["mine"]) == list(range(n_views))
    assert torch.equal(r0["flat"], r1["flat"])               # replicas stay identical
    params = _make_params()
    flat = FlatGrads(params)
    window_step(lambda v: _toy_render(params, v), _upstream, flat, n_views)   # single process: all views
    assert torch.allclose(flat.flat, r0["flat"], rtol=1e-5, atol=1e-6)
    # grads are views into the flat buffer, in PARAM_ORDER
    assert params["means3D"].grad.data_ptr() == flat.flat.data_ptr()
    assert flat.flat.numel() == sum(p.numel() for p in params.values())


def test_view_streams_without_side_streams_is_a_no_op():
    """n <= 1: no CUDA stream is created, fork/join do nothing, view() is a null context (CPU-safe)."""
    from diff_gaussian_rasterization.window import ViewStreams
    vs = ViewStreams("cpu", 1)
    assert len(vs) == 0
    vs.fork()
    with vs.view(3):
        pass
    vs.join()
