"""BASELINE config 4 (dev harness; bench.py imports `setup` / `step`): full-frame cuda_ray render, 800x800, 128^3 x 2
occupancy grid, F camera poses on an orbit (r = 1.2), inference path — fused kernel (csrc/render.cu) vs the reference-style
host loop of march_rays / forward / composite_rays / compaction (nerf/renderer.py:325-374).
  python scripts/bench_render.py [--frames 8] [--steps 3] [--loop]"""
import argparse
import json
import math
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_navigation_b200 import closures, synthetic
from nerf_navigation_b200.nerf.network import NeRFNetwork

H = W = 800
INTR = (1111.1, 1111.1, 400.0, 400.0)


def setup(dev, frames, first=0, total=64, seed=0):
    """Model + the rays of frames [first, first+frames) of a `total`-pose orbit: rays_o, rays_d [frames*H*W, 3]."""
    torch.manual_seed(seed)
    model = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10).to(dev)
    model.encoder.embeddings.data.uniform_(-0.5, 0.5)
    synthetic.install_stonehenge(model)
    model.eval()
    poses = torch.stack([closures.look_at_pose((1.2 * math.cos(2 * math.pi * k / total), 1.2 * math.sin(2 * math.pi * k / total), 0.4))
                         for k in range(first, first + frames)]).to(dev)
    pix = torch.arange(H * W, device=dev)
    ro, rd = closures.rays_from_poses(poses, pix, INTR, W)
    return model, ro.reshape(-1, 3).contiguous(), rd.reshape(-1, 3).contiguous()


def step(model, ro, rd, fused=True):
    model.fused_render = fused
    with torch.no_grad():
        out = model.render(ro.unsqueeze(0), rd.unsqueeze(0), bg_color=1.0, perturb=False, dt_gamma=0, max_steps=1024)
    model.fused_render = True
    return out


def run(dev, frames=8, steps=3, fused=True):
    from nerf_navigation_b200.field import render_rays_infer
    model, ro, rd = setup(dev, frames)
    step(model, ro, rd, fused)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        out = step(model, ro, rd, fused)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    _, _, _, stats = render_rays_infer(model, ro, rd, model.aabb_infer, 1.0, return_stats=True)
    ns = int(stats[1].item())
    n = ro.shape[0]
    return {"path": "fused" if fused else "loop", "frames": frames, "rays": n, "ms_per_step": ms, "ms_per_frame": ms / frames,
            "rays_per_s": n / (ms / 1e3), "samples": ns, "samples_per_ray": ns / n, "samples_per_s": ns / (ms / 1e3),
            "image_mean": float(out["image"].mean())}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=8)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--loop", action="store_true")
    a = ap.parse_args()
    t0 = time.time()
    print(json.dumps(run(torch.device("cuda", 0), a.frames, a.steps, not a.loop)), f"wall {time.time() - t0:.1f}s")
