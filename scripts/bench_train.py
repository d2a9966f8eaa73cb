"""BASELINE config 5 (dev harness; bench.py imports `setup` / `step`): stonehenge-shaped synthetic NeRF training step,
bound = 2, 2^19 table, 4096 rays per GPU from random poses on the orbit, run_cuda train branch, MSE against analytic
target colours, backward to table + MLPs, NCCL all-reduce of the gradients, fused Adam (betas 0.9/0.99, eps 1e-15,
lr 1e-2; main_nerf.py:126).
  python scripts/bench_train.py [--rays 4096] [--steps 20]           (1 GPU)
  torchrun --nproc-per-node N scripts/bench_train.py                  (N GPUs, NCCL)"""
import argparse
import json
import math
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_navigation_b200 import closures, synthetic, training
from nerf_navigation_b200.nerf.network import NeRFNetwork

W = 800
INTR = (1111.1, 1111.1, 400.0, 400.0)


def setup(dev, rays, rank=0, seed=0):
    torch.manual_seed(seed)                       # identical initial weights on every rank (DDP broadcasts rank 0's)
    model = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10).to(dev)
    synthetic.install_stonehenge(model)
    model.mean_count = rays * 80                  # as after the first update_extra_state (renderer.py:533-536)
    opt = training.make_optimizer(model, lr=1e-2, ema_decay=0.95)
    g = torch.Generator().manual_seed(1000 + rank)  # each rank draws its own rays
    ang = torch.rand(rays, generator=g) * 2 * math.pi
    poses = torch.stack([closures.look_at_pose((1.2 * math.cos(a), 1.2 * math.sin(a), 0.4)) for a in ang[:16].tolist()]).to(dev)
    pix = torch.randint(0, W * W, (rays // 16,), generator=g).to(dev)
    ro, rd = closures.rays_from_poses(poses, pix, INTR, W)
    ro, rd = ro.reshape(-1, 3).contiguous(), rd.reshape(-1, 3).contiguous()
    target = (0.5 + 0.5 * torch.sin(3.0 * rd)).contiguous()   # analytic colours
    return model, opt, ro, rd, target


def step(model, opt, ro, rd, target, world):
    return training.train_step(model, opt, ro, rd, target, world=world, perturb=True, force_all_rays=False, dt_gamma=0,
                               max_steps=1024)


def parity(dev, rays, rank, world, steps=3):
    """SURVEY §8(d) config 5 on hardware: `steps` optimiser steps with the gradients averaged over the ranks (each rank its own
    rays) against a single-process run of the same steps on the CONCATENATED batch (every rank computes it: the inputs are
    seeded).  perturb = False and force_all_rays so that a ray takes the same samples in both runs.  Compared: the reduced
    gradient of the first step (well conditioned: 1e-5 of the tensor's largest entry) and the parameters after the last
    step.  Adam with eps = 1e-15 turns a gradient into +-lr whatever its size, so a parameter whose gradient is at the
    rounding level of the sum can land on the other side; reported as the fraction of parameters within 1e-5 (of the
    tensor's largest entry) plus the worst relative difference."""
    kw = dict(perturb=False, force_all_rays=True, dt_gamma=0, max_steps=1024)
    shards = [setup(dev, rays, r)[2:] for r in range(world)]
    model_d, opt_d = setup(dev, rays, rank)[:2]
    model_s, opt_s = setup(dev, rays, rank)[:2]
    ro, rd, target = (torch.cat([s[i] for s in shards]) for i in range(3))
    my = shards[rank]
    grad_rel, grad_rel_by = 0.0, {}
    names = [n for n, p_ in model_d.named_parameters() if p_.requires_grad]
    for it in range(steps):
        training.train_step(model_d, opt_d, my[0], my[1], my[2], world=world, **kw)
        training.train_step(model_s, opt_s, ro, rd, target, world=1, **kw)
        if it == 0:
            for i, (a, b) in enumerate(zip(training.trainable(model_d), training.trainable(model_s))):
                ga, gb = a.grad / world, b.grad
                e = float((ga - gb).abs().max() / gb.abs().max().clamp_min(1e-30))
                grad_rel_by[names[i] if i < len(names) else str(i)] = e
                grad_rel = max(grad_rel, e)
    worst, inside, total = 0.0, 0, 0
    for a, b in zip(training.trainable(model_d), training.trainable(model_s)):
        d = (a.detach() - b.detach()).abs() / b.detach().abs().max().clamp_min(1e-30)
        worst = max(worst, float(d.max()))
        inside += int((d <= 1e-5).sum()); total += d.numel()
    ok = grad_rel < 1e-5 and inside / total > 0.999
    if world > 1:
        t = torch.tensor([1.0 if ok else 0.0], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        ok = bool(t.item() > 0)
    return {"train_parity": ok, "reduced_gradient_rel_err": grad_rel, "reduced_gradient_rel_err_by_tensor": grad_rel_by,
            "params_within_1e-5": inside / total,
            "params_worst_rel_diff": worst, "steps": steps, "rays": world * rays,
            "what": "rank-mean gradients / post-step parameters vs a single-process run on the concatenated batch"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rays", type=int, default=4096)
    ap.add_argument("--steps", type=int, default=20)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    model, opt, ro, rd, target = setup(dev, a.rays, rank)
    losses = [float(step(model, opt, ro, rd, target, world)) for _ in range(5)]
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        loss = step(model, opt, ro, rd, target, world)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    if world > 1:
        t = torch.tensor([ms], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t)
    if rank == 0:
        print(json.dumps({"world": world, "rays_per_gpu": a.rays, "ms_per_step": ms, "steps_per_s": 1e3 / ms,
                          "rays_per_s": world * a.rays / (ms / 1e3), "loss_first": losses[0], "loss_last": float(loss),
                          "grad_allreduce_bytes": sum(p.numel() for p in training.trainable(model)) * 4}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
