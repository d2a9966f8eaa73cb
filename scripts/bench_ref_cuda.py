"""Reported baseline (north_star): the reference's OWN CUDA extensions, compiled unmodified for sm_100a into oracle/_ref
(oracle/build_ref.py), timed with CUDA events on the same box for config 2 (field query fwd + dL/dx bwd) and config 3
(march_rays_train + field + composite_rays_train, fwd + bwd).  The reference's cuda_ray path has no pose gradient; its
backward reaches sigma/rgb -> MLP -> table (nerf/utils.py:446-450) and that is what is timed.
  python scripts/bench_ref_cuda.py [--points 4194304] [--rays 32768]   -> one JSON line"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_pipeline as RP
from nerf_navigation_b200 import closures, synthetic
from nerf_navigation_b200.nerf.network import NeRFNetwork


def timeit(fn, reps, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=1 << 22)
    ap.add_argument("--rays", type=int, default=32768)
    a = ap.parse_args()
    if not RP.available():
        print(json.dumps({"unavailable": "oracle/_ref extensions not built"}))
        return
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    ours = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10).to(dev)
    ours.encoder.embeddings.data.uniform_(-0.5, 0.5)
    synthetic.install_stonehenge(ours)
    emb = ours.encoder.embeddings.detach().clone().requires_grad_(True)
    sw = [l.weight.detach().clone().requires_grad_(True) for l in ours.sigma_net]
    cw = [l.weight.detach().clone().requires_grad_(True) for l in ours.color_net]
    field = RP.RefField(emb, ours.encoder.offsets, ours.encoder.per_level_scale, sw, cw, 2.0)
    out = {"what": "reference CUDA extensions (unmodified sources, sm_100a build) on this box"}

    # ---- config 2: density fwd + dL/dx bwd at `points` uniform points (planner collision cost)
    B = a.points
    x = (torch.rand(B, 3, device=dev) * 2 - 1)

    def cfg2():
        xin = x.detach().requires_grad_(True)
        sigma, _ = field.density(xin)
        (sigma ** 2).mean().backward()
        for p in field.params():
            p.grad = None

    ms = timeit(cfg2, 5)
    out["cfg2_field_fwd_bwd"] = {"points": B, "ms": ms, "Msamples_per_s": B / ms / 1e3}

    # ---- config 3: photometric loss through the cuda_ray train branch, `rays` rays per call
    gt = closures.look_at_pose((1.2, 0.0, 0.4))
    hyp = max(1, a.rays // 4096)
    poses = closures.perturb_poses(gt, hyp, generator=torch.Generator().manual_seed(0)).to(dev)
    pix = torch.randint(0, 640000, (4096,), generator=torch.Generator().manual_seed(0)).to(dev)
    ro, rd = closures.rays_from_poses(poses, pix, (1111.1, 1111.1, 400.0, 400.0), 800)
    ro, rd = ro.reshape(-1, 3).contiguous(), rd.reshape(-1, 3).contiguous()
    target = torch.rand(ro.shape[0], 3, device=dev)
    info = {}

    def cfg3(backward=True):
        image, m = RP.ref_render_train(field, ro, rd, ours.density_bitfield, ours.cascade, ours.grid_size, ours.aabb_train, 0.2)
        info["m"] = m
        if backward:
            ((image - target) ** 2).mean().backward()
            for p in field.params():
                p.grad = None

    with torch.no_grad():
        ms_f = timeit(lambda: cfg3(False), 3)
    ms_fb = timeit(cfg3, 3)
    N = ro.shape[0]
    out["cfg3_estimator"] = {"rays_per_call": N, "samples_per_ray": info["m"] / N, "fwd_ms": ms_f, "fwd_bwd_ms": ms_fb,
                             "fwd_rays_per_s": N / ms_f * 1e3, "fwd_bwd_rays_per_s": N / ms_fb * 1e3,
                             "note": "backward = sigma/rgb -> MLPs -> table atomics (no pose gradient exists in the reference's cuda_ray path)"}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
