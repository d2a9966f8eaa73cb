"""One optimiser step of the training configuration (BASELINE config 5): the run_cuda train branch
(march_rays_train -> field -> composite_rays_train) on this rank's rays, MSE against the target colours, backward to the
hash table and the MLPs, ONE all-reduce (sum) of the gradients over the ranks, fused Adam with the 1/world mean folded in.

Reference: Trainer.train_step + train_one_epoch (nerf/utils.py:425-450, 775-791) under DistributedDataParallel
(nerf/utils.py:368-369), whose gradient averaging is what the explicit all-reduce restates; optimiser main_nerf.py:126."""
import torch
import torch.distributed as dist

from .optim import FusedAdam


def make_optimizer(model, lr=1e-2, ema_decay=None):
    """main_nerf.py:126 — Adam(betas=(0.9, 0.99), eps=1e-15) over model.get_params(lr)."""
    groups = [{"params": list(g["params"]), "lr": g["lr"]} for g in model.get_params(lr)]
    groups = [g for g in groups if g["params"]]
    return FusedAdam(groups, lr=lr, betas=(0.9, 0.99), eps=1e-15, ema_decay=ema_decay)


def trainable(model):
    return [p for p in model.parameters() if p.requires_grad]


class GradBucket:
    """All gradients of the model as views of ONE flat fp32 tensor: the table-gradient kernels write their result straight
    into the table's slice (gridencoder.table_grad.set_output), the small MLP gradients are copied in after backward, the
    DDP mean is a single in-place all-reduce of the bucket (no torch.cat, no copy back), and the optimiser reads the views."""

    def __init__(self, params):
        from .gridencoder import table_grad
        self.params = list(params)
        total = sum(p.numel() for p in self.params)
        self.flat = torch.zeros(total, dtype=torch.float32, device=self.params[0].device)
        self.views, o = [], 0
        for p in self.params:
            self.views.append(self.flat[o:o + p.numel()].view_as(p))
            o += p.numel()
        self.big = [i for i, p in enumerate(self.params) if p.dim() == 2 and p.numel() >= (1 << 20)]   # hash tables
        for i in self.big:
            table_grad.set_output(self.params[i], self.views[i])

    def collect(self):
        """After backward: make every p.grad the bucket view (tables already are; MLP gradients are copied in)."""
        for p, v in zip(self.params, self.views):
            if p.grad is None:
                v.zero_()
            elif p.grad.data_ptr() != v.data_ptr():
                v.copy_(p.grad)
            p.grad = v

    def allreduce_(self, world):
        if world > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)


def allreduce_grads_(params, world):
    """Sum the gradients over the ranks in one flat fp32 bucket (table 50.6 MB + MLPs 37 KB); the mean's 1/world is
    applied inside the Adam kernel.  No-op for world == 1.  (Fallback for callers without a GradBucket.)"""
    if world == 1:
        return
    grads = [p.grad for p in params if p.grad is not None]
    flat = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    o = 0
    for g in grads:
        g.copy_(flat[o:o + g.numel()].view_as(g))
        o += g.numel()


def grad_bucket(model):
    b = getattr(model, "_grad_bucket", None)
    params = trainable(model)
    if b is None or len(b.params) != len(params) or any(x is not y for x, y in zip(b.params, params)) or \
            b.flat.device != params[0].device:
        b = GradBucket(params)
        model._grad_bucket = b
    return b


def train_step(model, optimizer, rays_o, rays_d, target, world=1, bg_color=1.0, perturb=True, force_all_rays=False,
               **render_kwargs):
    """rays_o, rays_d [N,3] (this rank's rays), target [N,3] -> loss (detached scalar tensor, this rank's mean)."""
    model.train()
    optimizer.zero_grad(set_to_none=True)
    out = model.render(rays_o.unsqueeze(0), rays_d.unsqueeze(0), staged=False, bg_color=bg_color, perturb=perturb,
                       force_all_rays=force_all_rays, **render_kwargs)
    loss = ((out["image"].reshape(-1, 3) - target) ** 2).mean()
    bucket = grad_bucket(model)
    loss.backward()
    bucket.collect()
    bucket.allreduce_(world)
    optimizer.grad_mult = 1.0 / world
    optimizer.step()
    return loss.detach()
