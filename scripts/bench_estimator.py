"""BASELINE config 3 (dev harness; bench.py imports `run`): pose-estimator photometric loss, 800x800 synthetic frame,
4096 rays x P pose hypotheses, march_rays_train / composite_rays_train forward + backward to the poses.
  python scripts/bench_estimator.py [--hyp 256] [--rays 4096] [--chunk 16] [--steps 3]"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_navigation_b200 import closures, synthetic
from nerf_navigation_b200.nerf.network import NeRFNetwork

H = W = 800
INTR = (1111.1, 1111.1, 400.0, 400.0)


def setup(dev, hyp, rays, seed=0):
    torch.manual_seed(seed)
    model = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10).to(dev)
    model.encoder.embeddings.data.uniform_(-0.5, 0.5)
    synthetic.install_stonehenge(model)
    for p in model.parameters():
        p.requires_grad_(False)
    model.train()  # selects the march_rays_train / composite_rays_train branch (renderer.py:282); the field is frozen
    g = torch.Generator().manual_seed(seed)
    gt = closures.look_at_pose((1.2, 0.0, 0.4))
    poses = closures.perturb_poses(gt, hyp, generator=g).to(dev)
    pix = torch.randint(0, H * W, (rays,), generator=g).to(dev)
    # the same hypotheses in the estimator's own coordinates: 12-states around the state that sees `gt`
    # (position +- 0.02 m, rotation vector +- 0.02 rad: SURVEY §8d config 3)
    st0 = closures.pose_to_state(gt)
    noise = torch.zeros(hyp, 12)
    noise[:, 0:3] = torch.randn(hyp, 3, generator=g) * 0.02
    noise[:, 6:9] = torch.randn(hyp, 3, generator=g) * 0.02
    noise[0] = 0
    model.estimator_states = (st0[None, :] + noise).to(dev)
    kw = dict(bg_color=1.0, perturb=False, dt_gamma=0, max_steps=1024, force_all_rays=True)
    with torch.no_grad():
        ro, rd = closures.rays_from_poses(poses[:1], pix, INTR, W)
        target = model.render(ro.reshape(1, -1, 3), rd.reshape(1, -1, 3), **kw)["image"].reshape(rays, 3)
    return model, poses, pix, target, kw


def step(model, poses, pix, target, kw, chunk):
    """losses [P] and d loss_p / d pose_p [P,4,4], evaluated `chunk` hypotheses at a time."""
    P = poses.shape[0]
    losses = torch.empty(P, device=poses.device)
    grads = torch.empty_like(poses)
    for a in range(0, P, chunk):
        p = poses[a:a + chunk].detach().requires_grad_(True)
        l = closures.photometric_loss(model, p, pix, target, INTR, W, **kw)
        l.sum().backward()
        losses[a:a + chunk] = l.detach()
        grads[a:a + chunk] = p.grad
    return losses, grads


def step_states(model, states, pix, target, kw):
    """The estimator step as nav/ runs it, for ALL hypotheses in one launch chain: 12-states -> poses (closures.state_to_pose)
    -> rays of the sampled pixels (one kernel) -> march_rays_train -> fused field + colour -> composite -> per-hypothesis MSE
    -> backward to the states.  Returns losses [P], dL_p / d state_p [P,12]."""
    s = states.detach().requires_grad_(True)
    l = closures.measurement_loss(model, s, pix, target, INTR, W, **kw)
    l.sum().backward()
    return l.detach(), s.grad


def run(dev, hyp=256, rays=4096, chunk=16, steps=3, warmup=1):
    model, poses, pix, target, kw = setup(dev, hyp, rays)
    for _ in range(warmup):
        step(model, poses, pix, target, kw, chunk)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        losses, grads = step(model, poses, pix, target, kw, chunk)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    # samples per ray of this scene (one chunk)
    with torch.no_grad():
        ro, rd = closures.rays_from_poses(poses[:chunk], pix, INTR, W)
        ro, rd = ro.reshape(-1, 3).contiguous(), rd.reshape(-1, 3).contiguous()
        from nerf_navigation_b200 import raymarching as rm
        nears, fars = rm.near_far_from_aabb(ro, rd, model.aabb_train, model.min_near)
        xyzs, _, _, _ = rm.march_rays_train(ro, rd, model.bound, model.density_bitfield, model.cascade, model.grid_size,
                                            nears, fars, None, -1, False, -1, True, 0, 1024)
        spr = xyzs.shape[0] / ro.shape[0]
    return {"rays_per_s": hyp * rays / (ms / 1e3), "ms_per_step": ms, "hypotheses": hyp, "rays_per_hypothesis": rays,
            "chunk_hypotheses": chunk, "samples_per_ray": spr, "samples_per_s": hyp * rays * spr / (ms / 1e3),
            "loss_gt": float(losses[0]), "loss_mean": float(losses.mean()), "grad_abs_mean": float(grads.abs().mean())}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--hyp", type=int, default=256)
    ap.add_argument("--rays", type=int, default=4096)
    ap.add_argument("--chunk", type=int, default=16)
    ap.add_argument("--steps", type=int, default=3)
    a = ap.parse_args()
    t0 = time.time()
    print(json.dumps(run(torch.device("cuda", 0), a.hyp, a.rays, a.chunk, a.steps)), f"wall {time.time() - t0:.1f}s")
