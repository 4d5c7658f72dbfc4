"""Run by tests/test_multi_gpu.py under torch.distributed.run with one process per GPU (NCCL).

A 6-view mapping window of the REAL rasterizer is sharded over the ranks with `window.window_step`; after the exchange
every rank's flat gradient buffer must equal the gradient of all six views rendered by one rank alone (1e-3 relative,
the gradient bar of the suite), be bit-identical on all ranks, and the pose gradients (theta, rho) of a view must exist
on the rank that owns the view only.  argv[1] = exchange kind ("nccl" | "auto" | "peer")."""
import os
import sys

import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (HERE, os.path.join(ROOT, "splat-slam_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import diff_gaussian_rasterization as m  # noqa: E402
from diff_gaussian_rasterization import synthetic as syn  # noqa: E402
from diff_gaussian_rasterization.window import FlatGrads, ViewStreams, partition_views, window_step  # noqa: E402

N_VIEWS, CAM, P = 6, "small", 20000


def main():
    exchange = sys.argv[1] if len(sys.argv) > 1 else "nccl"
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    g = syn.make_gaussians(CAM, P, seed=3, sh_coeffs=4)
    Wc, Wd = (t.to(dev) for t in syn.upstream_grads(CAM))
    views = [syn.make_view(CAM, syn.window_pose(v)) for v in range(N_VIEWS)]
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)

    def build(exchange_kind):
        params = {k: g[k].to(dev).requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
        return params, FlatGrads(params, exchange=exchange_kind)

    def renderer(params, theta, rho):
        def render_view(v):
            vw = views[v]
            S = m.GaussianRasterizationSettings(
                image_height=vw.image_height, image_width=vw.image_width, tanfovx=vw.tanfovx, tanfovy=vw.tanfovy, bg=bg,
                scale_modifier=1.0, viewmatrix=vw.viewmatrix.to(dev), projmatrix=vw.projmatrix.to(dev),
                projmatrix_raw=vw.projmatrix_raw.to(dev), sh_degree=1, campos=vw.campos.to(dev), prefiltered=False,
                debug=False)
            color, radii, depth, opacity, n_touched = m.GaussianRasterizer(S)(
                means3D=params["means3D"], means2D=torch.zeros(P, 3, device=dev, requires_grad=True),
                opacities=params["opacities"], shs=params["shs"], scales=params["scales"], rotations=params["rotations"],
                theta=theta[v], rho=rho[v])
            return color, depth
        return render_view

    def upstream(v):
        return (1.0 + 0.1 * v) * Wc, Wd

    # sharded over the ranks, exchanged
    params, flat = build(exchange)
    theta = [torch.zeros(3, device=dev, requires_grad=True) for _ in range(N_VIEWS)]
    rho = [torch.zeros(3, device=dev, requires_grad=True) for _ in range(N_VIEWS)]
    for streams in (None, ViewStreams(dev, 3)):
        for t in theta + rho:
            t.grad = None
        mine = window_step(renderer(params, theta, rho), upstream, flat, N_VIEWS, streams=streams)
        torch.cuda.synchronize(dev)
        assert mine == partition_views(N_VIEWS, world, rank)
        sharded = flat.flat.clone()
        for v in range(N_VIEWS):
            owned = v in mine
            assert (theta[v].grad is not None and float(theta[v].grad.abs().max()) > 0) == owned, (rank, v)
            assert (rho[v].grad is not None) == owned, (rank, v)
        # bit-identical on every rank
        gathered = [torch.empty_like(sharded) for _ in range(world)]
        dist.all_gather(gathered, sharded)
        for r in range(world):
            assert torch.equal(gathered[r], gathered[0]), f"rank {r} differs from rank 0 ({flat.exchange_kind})"
        # equal to one rank rendering all views
        p1, f1 = build("nccl")
        th1 = [torch.zeros(3, device=dev, requires_grad=True) for _ in range(N_VIEWS)]
        rh1 = [torch.zeros(3, device=dev, requires_grad=True) for _ in range(N_VIEWS)]
        rv = renderer(p1, th1, rh1)
        f1.zero_()
        for v in range(N_VIEWS):
            torch.autograd.backward(list(rv(v)), list(upstream(v)))
        torch.cuda.synchronize(dev)
        rel = float((sharded - f1.flat).abs().max() / f1.flat.abs().max())
        assert rel <= 1e-3, (rank, rel)
        for v in mine:
            assert float((theta[v].grad - th1[v].grad).abs().max() / th1[v].grad.abs().max()) <= 1e-3
    kinds = [None] * world
    dist.all_gather_object(kinds, flat.exchange_kind)
    if rank == 0:
        print(f"MULTIRANK_OK world={world} exchange={kinds[0]} rel={rel:.2e}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
