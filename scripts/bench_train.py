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
