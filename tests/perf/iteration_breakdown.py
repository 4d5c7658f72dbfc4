"""Where one mapping iteration of the 'ours' arm of bench_iteration.py spends its time (CUDA events, one stream for
the sections outside the view loop). gpurun -- 'python tests/perf/iteration_breakdown.py'"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
os.environ.setdefault("P", "1000000")
import bench_iteration as B  # noqa: E402  (runs its own measurement first)

st = B.make_state(B.GaussianAdam)
mod = B.ours_mod
ev = [torch.cuda.Event(enable_timing=True) for _ in range(7)]
acc = [0.0] * 6
N = 5
for it in range(N + 2):
    ev[0].record()
    scales, rots, opac = B.fused_activations(st.scaling, st.rotation, st.opacity)
    leaves = [t.detach().requires_grad_(True) for t in (scales, rots, opac)]
    ev[1].record()
    vsp, radii_acm, nt_acm = [], [], []
    B.streams.fork()
    for k, vp in enumerate(st.views):
        with B.streams.view(k):
            means2D = torch.zeros_like(st.xyz, requires_grad=True)
            color, radii, depth, opacity, n_touched = mod.GaussianRasterizer(B.settings(mod, st, vp))(
                means3D=st.xyz, means2D=means2D, shs=st.f_dc, opacities=leaves[2], scales=leaves[0], rotations=leaves[1],
                theta=vp.theta, rho=vp.rho)
            B.fused_mapping_loss(color, depth, vp.exposure_a, vp.exposure_b, vp.gt_image, vp.gt_depth, 0.95, 0.01).backward()
            vsp.append(means2D); radii_acm.append(radii); nt_acm.append(n_touched)
    B.streams.join()
    ev[2].record()
    iso = 10 * torch.abs(scales - scales.mean(dim=1).view(-1, 1)).mean()
    torch.autograd.backward([iso, scales, rots, opac], [None, leaves[0].grad, leaves[1].grad, leaves[2].grad])
    ev[3].record()
    vis, _ = B.occ_aware_visibility(nt_acm, want_n_obs=False)
    B.add_densification_stats_window(st, vsp, radii_acm)
    ev[4].record()
    st.optimizer.step()
    st.optimizer.zero_grad(set_to_none=True)
    ev[5].record()
    st.keyframe_optimizers.step()
    st.keyframe_optimizers.zero_grad(set_to_none=True)
    ev[6].record()
    torch.cuda.synchronize()
    if it >= 2:
        for i in range(6):
            acc[i] += ev[i].elapsed_time(ev[i + 1]) / N
names = ["activations fwd", "10 views: render + loss + backward", "isotropic loss + activations bwd", "visibility + densify stats",
         "Gaussian Adam", "keyframe Adam"]
print({n: round(a, 3) for n, a in zip(names, acc)}, "total", round(sum(acc), 3), "ms")
