"""Generates the golden fixtures in this directory by running the UNMODIFIED reference extension
(oracle/_ref/ref_dgr, built by oracle/build_ref.sh) on a B200.

    gpurun -- 'python tests/golden/make_golden.py gpurun_out/golden'   # then copy gpurun_out/golden/*.pt here

Each fixture holds the inputs (so the test does not depend on RNG stability), the reference's public outputs,
its intermediates (sort keys, point list, tile ranges, n_contrib, final_T carved from the opaque buffers) and
all gradients for loss = (color*Wc).sum() + (depth*Wd).sum().
The reference has no tests or golden vectors of its own for this path; these files are what pins
the CPU oracle and the CUDA path when the reference extension itself is not importable.
"""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import parity as T  # noqa: E402

CASES = {
    "tiny_sh0": dict(camera="tiny", P=300),
    "ragged_sh1_bg": dict(camera="ragged", P=1500, sh_coeffs=4, sh_degree=1, bg=(0.2, 0.5, 0.7), view_k=2),
    "tiny_sh3": dict(camera="tiny", P=400, sh_coeffs=16, sh_degree=3, view_k=1, seed=3),
    "ragged_precomp": dict(camera="ragged", P=1200, use_colors_precomp=True, use_cov3d_precomp=True, view_k=3, seed=5),
    "ragged_big_splats": dict(camera="ragged", P=400, scale_modifier=1.7, sigma_px=9.0, seed=9),
}


def main(out_dir):
    ref = T.load_ref()
    assert ref is not None, "reference extension not built (oracle/build_ref.sh)"
    os.makedirs(out_dir, exist_ok=True)
    for name, kw in CASES.items():
        case = T.make_case(**kw)
        inter = T.ref_forward_with_intermediates(ref, case)
        full = T.run_module(ref, case)
        fx = {"kwargs": kw, "gpu": torch.cuda.get_device_name()}
        for k in ("means3D", "opacities", "shs", "colors_precomp", "scales", "rotations", "cov3D_precomp", "Wc", "Wd", "bg"):
            fx["in_" + k] = case[k]
        for k in ("color", "depth", "opacity", "radii", "n_touched"):
            fx[k] = full[k].cpu()
        for k in T.GRAD_NAMES:
            g = full.get("grad_" + k)
            fx["grad_" + k] = None if g is None else g.cpu()
        fx["num_rendered"] = int(inter["num_rendered"])
        for k in ("keys", "point_list", "ranges", "n_contrib", "final_T", "tiles_touched", "depths", "means2D",
                  "conic_opacity"):
            fx[k] = inter[k].cpu()
        # sanity: the two reference runs agree with each other
        assert torch.equal(fx["radii"], inter["radii"].cpu()) and torch.equal(fx["n_touched"], inter["n_touched"].cpu())
        torch.save(fx, os.path.join(out_dir, name + ".pt"))
        print(name, "R =", fx["num_rendered"], "visible =", int((fx["radii"] > 0).sum()))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(HERE, "..", "..", "gpurun_out", "golden"))
