"""TEST / BASELINE INFRASTRUCTURE ONLY — torch-CPU restatement of the reference's CPU-runnable path (BASELINE config 1)
and a torch transcription of the hash-grid lookup.  Used by tests/ (checked against tests/golden/cfg1.npz, which was
produced by importing the real reference: scripts/make_golden.py) and by bench.py's cpu_baseline / --impl reference
legs.  The product path never imports this.

Restated (reference file:line):
  freq_encode      encoding.py:5-43           [x, sin(2^k x), cos(2^k x)], k = 0..5
  Cfg1Field        nerf/network.py:33-68, 95-143, 163-191   sigma net 39-64-16, colour net 54-64-64-3, no bias, ReLU,
                   trunc_exp (activation.py:5-17), sigmoid; colour evaluated only where weights > 1e-4
  render_run       nerf/renderer.py:125-254 (upsample_steps = 0 branch) + :542-575 staging
  hashgrid_encode  gridencoder/src/gridencoder.cu:75-224 (forward) with autograd supplying :227-343
"""
import numpy as np
import torch

from . import oracle as O


def freq_encode(x, n_freqs=6):
    out = [x]
    for k in range(n_freqs):
        f = float(2.0 ** k)
        out += [torch.sin(x * f), torch.cos(x * f)]
    return torch.cat(out, dim=-1)


class _TruncExp(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        ctx.save_for_backward(x)
        return torch.exp(x)

    @staticmethod
    def backward(ctx, g):
        return g * torch.exp(ctx.saved_tensors[0].clamp(-15, 15))


class Cfg1Field:
    """Weights: dict with sigma_net.{0,1}.weight, color_net.{0,1,2}.weight (the reference state-dict names)."""

    def __init__(self, sd, bound=2.0, density_scale=1.0, min_near=0.2):
        self.Ws = [torch.as_tensor(sd[f"sigma_net.{i}.weight"]) for i in range(2)]
        self.Wc = [torch.as_tensor(sd[f"color_net.{i}.weight"]) for i in range(3)]
        self.bound, self.density_scale, self.min_near = float(bound), float(density_scale), float(min_near)
        self.aabb = np.array([-bound] * 3 + [bound] * 3, np.float32)

    @staticmethod
    def random(seed=0, **kw):
        g = torch.Generator().manual_seed(seed)
        shapes = {"sigma_net.0.weight": (64, 39), "sigma_net.1.weight": (16, 64), "color_net.0.weight": (64, 54),
                  "color_net.1.weight": (64, 64), "color_net.2.weight": (3, 64)}
        sd = {k: (torch.rand(s, generator=g) * 2 - 1) / np.sqrt(s[1]) for k, s in shapes.items()}
        return Cfg1Field(sd, **kw)

    def density(self, x):
        h = torch.relu(freq_encode(x) @ self.Ws[0].t()) @ self.Ws[1].t()
        return _TruncExp.apply(h[..., 0]), h[..., 1:]

    def color(self, d, geo, mask):
        rgb = torch.zeros(mask.shape[0], 3, dtype=d.dtype)
        if mask.any():
            h = torch.cat([freq_encode(d[mask]), geo[mask]], dim=-1)
            h = torch.relu(h @ self.Wc[0].t())
            h = torch.relu(h @ self.Wc[1].t())
            rgb[mask] = torch.sigmoid(h @ self.Wc[2].t())
        return rgb


def render_run(field, rays_o, rays_d, num_steps=512, bg_color=1.0, max_ray_batch=4096):
    """rays_o, rays_d [N,3] -> image [N,3], depth [N] (uniform sampler, no perturbation, no upsampling)."""
    images, depths = [], []
    lo, hi = torch.from_numpy(field.aabb[:3]), torch.from_numpy(field.aabb[3:])
    for a in range(0, rays_o.shape[0], max_ray_batch):
        o, d = rays_o[a:a + max_ray_batch], rays_d[a:a + max_ray_batch]
        n = o.shape[0]
        nr, fr = O.near_far_from_aabb(o.detach().numpy(), d.detach().numpy(), field.aabb, field.min_near)
        near, far = torch.from_numpy(nr)[:, None], torch.from_numpy(fr)[:, None]
        z = near + (far - near) * torch.linspace(0.0, 1.0, num_steps)[None, :]
        dist = (far - near) / num_steps
        xyz = torch.min(torch.max(o[:, None, :] + d[:, None, :] * z[..., None], lo), hi)
        sigma, geo = field.density(xyz.reshape(-1, 3))
        sigma = sigma.view(n, num_steps)
        deltas = torch.cat([z[:, 1:] - z[:, :-1], dist * torch.ones_like(z[:, :1])], dim=-1)
        alpha = 1 - torch.exp(-deltas * field.density_scale * sigma)
        trans = torch.cumprod(torch.cat([torch.ones_like(alpha[:, :1]), 1 - alpha + 1e-15], dim=-1), dim=-1)[:, :-1]
        w = alpha * trans
        rgb = field.color(d[:, None, :].expand_as(xyz).reshape(-1, 3), geo, (w > 1e-4).reshape(-1)).view(n, num_steps, 3)
        wsum = w.sum(-1)
        depths.append((w * ((z - near) / (far - near)).clamp(0, 1)).sum(-1))
        images.append((w[..., None] * rgb).sum(-2) + (1 - wsum)[:, None] * bg_color)
    return torch.cat(images), torch.cat(depths)


def hashgrid_encode(x01, table, offsets, scales, align_corners=False):
    """Torch transcription of the hash-grid forward for D = 3 (hash type): x01 [B,3] in [0,1], table [rows,C] ->
    [B, L*C].  Differentiable w.r.t. x01 and table through autograd."""
    B = x01.shape[0]
    L = len(offsets) - 1
    outs = []
    primes = (1, 2654435761, 805459861)
    for l in range(L):
        rows = int(offsets[l + 1] - offsets[l])
        scale = float(scales[l])
        res = int(np.ceil(scale)) + 1
        # one rounding like the kernel's fma (gridencoder.cu:134 contracts under nvcc's default -fmad=true)
        pos = (x01.double() * scale + (0.0 if align_corners else 0.5)).float()
        pg = torch.floor(pos)
        f = pos - pg
        pg = pg.long()
        step = res if align_corners else res + 1
        dense = step ** 3 <= rows
        acc = 0
        for k in range(8):
            w = 1.0
            idx = 0
            stride = 1
            for dd in range(3):
                bit = (k >> dd) & 1
                c = pg[:, dd] + bit
                w = w * (f[:, dd] if bit else 1 - f[:, dd])
                if dense:
                    idx = idx + c * stride
                    stride *= step
                else:
                    idx = (idx ^ ((c * primes[dd]) & 0xFFFFFFFF)) if dd else (c & 0xFFFFFFFF)
            idx = (idx % rows) + int(offsets[l])
            acc = acc + w[:, None] * table[idx]
        outs.append(acc)
    return torch.cat(outs, dim=-1).view(B, -1)
