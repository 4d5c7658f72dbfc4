"""TEST INFRASTRUCTURE — CPU oracle of the densification bookkeeping that follows backward in a mapping iteration.

Plain-PyTorch restatement of src/mapper.py:494-499 (occlusion-aware visibility), :507-509 (n_obs), :519-529
(max_radii2D) and GaussianModel.add_densification_stats (thirdparty/gaussian_splatting/scene/gaussian_model.py:
738-742). Pinned by tests/golden/make_golden_densify.py, which calls the reference's own add_densification_stats
(imported in the build container) on the same seeded inputs. Used only by tests.
"""
import torch


def densification_stats_window(max_radii2D, xyz_gradient_accum, denom, radii_list, viewspace_grad_list):
    """In place, view by view, exactly the reference's indexing ops. Shapes: [P], [P,1], [P,1]; radii [P] int,
    viewspace grads [P,3]."""
    for radii, grad in zip(radii_list, viewspace_grad_list):
        vis = radii > 0                                                      # render_pkg["visibility_filter"]
        max_radii2D[vis] = torch.max(max_radii2D[vis], radii[vis])           # mapper.py:520-523
        xyz_gradient_accum[vis] += torch.norm(grad[vis, :2], dim=-1, keepdim=True)   # gaussian_model.py:739-741
        denom[vis] += 1                                                      # gaussian_model.py:742
    return max_radii2D, xyz_gradient_accum, denom


def occ_aware_visibility(n_touched_list):
    """mapper.py:494-499: one int64 0/1 vector per window view; :507-509: n_obs = their sum."""
    vis = [(n > 0).long() for n in n_touched_list]
    n_obs = torch.zeros_like(n_touched_list[0], dtype=torch.int32)
    for v in vis:
        n_obs += v.to(torch.int32)
    return vis, n_obs


def synthetic_densify_case(P=500, K=5, seed=0):
    g = torch.Generator().manual_seed(seed)
    radii = [(torch.randint(-3, 40, (P,), generator=g).clamp_min(0)).to(torch.int32) for _ in range(K)]
    grads = [torch.randn(P, 3, generator=g) * 10.0 ** float(torch.randint(-7, -2, (1,), generator=g)) for _ in range(K)]
    n_touched = [torch.randint(-1, 3, (P,), generator=g).clamp_min(0).to(torch.int32) * 50 for _ in range(K)]
    state = dict(max_radii2D=torch.randint(0, 30, (P,), generator=g).float(),
                 xyz_gradient_accum=torch.rand(P, 1, generator=g) * 1e-4, denom=torch.randint(0, 9, (P, 1), generator=g).float())
    return dict(radii=radii, grads=grads, n_touched=n_touched, **state)
