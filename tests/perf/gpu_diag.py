"""Development diagnostic (GPU box): compare this repo's CUDA path with the compiled reference on several cases,
print every error figure without stopping at the first failure, and time both. Writes gpurun_out/diag.json."""
import json
import os
import sys
import time
import traceback

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import parity as T  # noqa: E402

os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
report = {}


def inter_report(case, name):
    ref = T.load_ref()
    o = T.ours_forward_with_intermediates(case)
    r = T.ref_forward_with_intermediates(ref, case)
    rep = {"R_ours": o["num_rendered"], "R_ref": r["num_rendered"]}
    vis = (r["radii"].cpu() > 0)
    rep["visible"] = int(vis.sum())
    for k in ("radii", "tiles_touched", "n_touched"):
        rep[k] = int((o[k].cpu().long() != r[k].cpu().long()).sum())
    for k in ("depths", "means2D", "conic_opacity", "cov3D", "rgb"):
        a, b = o[k].cpu()[vis], r[k].cpu()[vis]
        rep[k + "_bits"] = int((~T.bits_equal(a, b)).sum())
        rep[k + "_maxabs"] = float((a - b).abs().max()) if a.numel() else 0.0
    if o["num_rendered"] == r["num_rendered"]:
        for k in ("keys", "point_list"):
            rep[k] = int((o[k].cpu().long() != r[k].cpu().long()).sum())
        rep["ranges"] = int((o["ranges"].cpu().long() != r["ranges"].cpu().long()).sum())
    rep["n_contrib"] = int((o["n_contrib"].cpu().long() != r["n_contrib"].cpu().long()).sum())
    rep["final_T_bits"] = int((~T.bits_equal(o["final_T"], r["final_T"])).sum())
    for k in ("color", "depth", "opacity"):
        rep[k + "_abs"] = float((o[k].cpu() - r[k].cpu()).abs().max())
    print(name, "INTERMEDIATES", json.dumps(rep))
    return rep


def grad_report(case, name):
    ref = T.load_ref()
    o = T.run_module(T.load_ours(), case)
    r = T.run_module(ref, case)
    rep = {}
    for k in ("radii", "n_touched"):
        rep[k] = int((o[k].cpu().long() != r[k].cpu().long()).sum())
    for k in ("color", "depth", "opacity"):
        rep[k + "_abs"] = float((o[k].cpu() - r[k].cpu()).abs().max())
    for k in T.GRAD_NAMES:
        ga, gb = o.get("grad_" + k), r.get("grad_" + k)
        if ga is None or gb is None:
            rep["grad_" + k] = None if (ga is None and gb is None) else "one-sided"
            continue
        rep["grad_" + k] = T.rel_err(ga, gb)
    print(name, "GRADS", json.dumps(rep))
    return rep


def timeit(mod, case, iters=20, warm=5):
    times = []
    for i in range(warm + iters):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        T.run_module(mod, case)
        torch.cuda.synchronize()
        if i >= warm:
            times.append(time.perf_counter() - t0)
    times.sort()
    return times[len(times) // 2] * 1e3


cases = {
    "tiny_sh0": dict(camera="tiny", P=300),
    "ragged_sh1": dict(camera="ragged", P=1500, sh_coeffs=4, sh_degree=1, bg=(0.2, 0.5, 0.7), view_k=2),
    "small_sh3": dict(camera="small", P=20000, sh_coeffs=16, sh_degree=3, view_k=1),
    "small_precomp": dict(camera="small", P=20000, use_colors_precomp=True, use_cov3d_precomp=True, view_k=3),
    "small_mod": dict(camera="small", P=5000, scale_modifier=1.7, sigma_px=6.0),
    "tum_200k": dict(camera="tum", P=200000, view_k=1),
    "replica_500k": dict(camera="replica", P=500000, view_k=0),
}
only = sys.argv[1:] or list(cases)
for name in only:
    try:
        case = T.make_case(**cases[name])
        report[name] = {"inter": inter_report(case, name), "grads": grad_report(case, name)}
    except Exception:
        traceback.print_exc()
        report[name] = {"error": traceback.format_exc()[-2000:]}
    torch.cuda.empty_cache()

try:
    case = T.make_case(camera="replica", P=1000000)
    t_ours = timeit(T.load_ours(), case)
    t_ref = timeit(T.load_ref(), case)
    report["timing_replica_1M_fwd_bwd_ms"] = {"ours": t_ours, "ref": t_ref}
    print("TIMING replica 1M fwd+bwd ms: ours", t_ours, "ref", t_ref)
except Exception:
    traceback.print_exc()

with open(os.path.join(ROOT, "gpurun_out", "diag.json"), "w") as f:
    json.dump(report, f, indent=1)
