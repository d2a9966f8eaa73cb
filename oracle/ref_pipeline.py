"""TEST / BASELINE INFRASTRUCTURE ONLY — drives the reference's OWN CUDA extensions (oracle/_ref/*.so, compiled unmodified
from /root/reference by oracle/build_ref.py) in the call sequence of the reference's Python wrappers, so that its CUDA
build can be timed and compared on the GPU box, where /root/reference itself does not exist.

Call sequences restated (not imported) from: gridencoder/grid.py:19-87 (`_grid_encode`: [L,B,C] output + permute,
dy_dx buffer, zero-filled grad_embeddings, atomics scatter), shencoder/sphere_harmonics.py:14-56, raymarching/
raymarching.py:161-286 (march_rays_train with its N*max_steps zero-filled buffers + .item() slice, composite_rays_train),
nerf/network.py:95-143 (Linear stacks, trunc_exp, sigmoid), nerf/renderer.py:282-323 (run_cuda train branch).
The product never imports this module.
"""
import numpy as np
import torch
import torch.nn.functional as F
from torch.autograd import Function

from . import build_ref

_mods = {}


def ext(name):
    if name not in _mods:
        _mods[name] = build_ref.load_ref(name)
    return _mods[name]


def available():
    return all(ext(n) is not None for n in ("gridencoder", "raymarching", "shencoder"))


class _RefGridEncode(Function):  # grid.py:19-87
    @staticmethod
    def forward(ctx, inputs, embeddings, offsets, per_level_scale, base_resolution, calc_grad_inputs):
        inputs = inputs.contiguous()
        B, D = inputs.shape
        L, C = offsets.shape[0] - 1, embeddings.shape[1]
        S, H = float(np.log2(per_level_scale)), base_resolution
        outputs = torch.empty(L, B, C, device=inputs.device, dtype=embeddings.dtype)
        dy_dx = torch.empty(B, L * D * C, device=inputs.device, dtype=embeddings.dtype) if calc_grad_inputs else \
            torch.empty(1, device=inputs.device, dtype=embeddings.dtype)
        ext("gridencoder").grid_encode_forward(inputs, embeddings, offsets, outputs, B, D, C, L, S, H, calc_grad_inputs, dy_dx, 0, False)
        outputs = outputs.permute(1, 0, 2).reshape(B, L * C)
        ctx.save_for_backward(inputs, embeddings, offsets, dy_dx)
        ctx.dims = (B, D, C, L, S, H, calc_grad_inputs)
        return outputs

    @staticmethod
    def backward(ctx, grad):
        inputs, embeddings, offsets, dy_dx = ctx.saved_tensors
        B, D, C, L, S, H, calc = ctx.dims
        grad = grad.view(B, L, C).permute(1, 0, 2).contiguous()
        grad_embeddings = torch.zeros_like(embeddings)
        grad_inputs = torch.zeros_like(inputs) if calc else torch.zeros(1, device=inputs.device, dtype=embeddings.dtype)
        ext("gridencoder").grid_encode_backward(grad, inputs, embeddings, offsets, grad_embeddings, B, D, C, L, S, H, calc, dy_dx, grad_inputs, 0, False)
        return (grad_inputs if calc else None), grad_embeddings, None, None, None, None


class _RefSH(Function):  # sphere_harmonics.py:14-56
    @staticmethod
    def forward(ctx, inputs, degree, calc_grad_inputs):
        inputs = inputs.contiguous()
        B = inputs.shape[0]
        out = torch.empty(B, degree ** 2, dtype=inputs.dtype, device=inputs.device)
        dy_dx = torch.empty(B, 3 * degree ** 2, dtype=inputs.dtype, device=inputs.device) if calc_grad_inputs else \
            torch.empty(1, dtype=inputs.dtype, device=inputs.device)
        ext("shencoder").sh_encode_forward(inputs, out, B, 3, degree, calc_grad_inputs, dy_dx)
        ctx.save_for_backward(inputs, dy_dx)
        ctx.cfg = (B, degree, calc_grad_inputs)
        return out

    @staticmethod
    def backward(ctx, grad):
        B, degree, calc = ctx.cfg
        if not calc:
            return None, None, None
        inputs, dy_dx = ctx.saved_tensors
        gi = torch.zeros_like(inputs)
        ext("shencoder").sh_encode_backward(grad.contiguous(), inputs, B, 3, degree, dy_dx, gi)
        return gi, None, None


class _RefTruncExp(Function):  # activation.py:5-17
    @staticmethod
    def forward(ctx, x):
        ctx.save_for_backward(x)
        return torch.exp(x)

    @staticmethod
    def backward(ctx, g):
        return g * torch.exp(ctx.saved_tensors[0].clamp(-15, 15))


class _RefComposite(Function):  # raymarching.py:233-286
    @staticmethod
    def forward(ctx, sigmas, rgbs, deltas, rays):
        sigmas, rgbs = sigmas.contiguous(), rgbs.contiguous()
        M, N = sigmas.shape[0], rays.shape[0]
        ws = torch.empty(N, dtype=sigmas.dtype, device=sigmas.device)
        depth = torch.empty(N, dtype=sigmas.dtype, device=sigmas.device)
        image = torch.empty(N, 3, dtype=sigmas.dtype, device=sigmas.device)
        ext("raymarching").composite_rays_train_forward(sigmas, rgbs, deltas, rays, M, N, ws, depth, image)
        ctx.save_for_backward(sigmas, rgbs, deltas, rays, ws, depth, image)
        ctx.dims = (M, N)
        return ws, depth, image

    @staticmethod
    def backward(ctx, g_ws, g_depth, g_image):
        sigmas, rgbs, deltas, rays, ws, depth, image = ctx.saved_tensors
        M, N = ctx.dims
        gs, gr = torch.zeros_like(sigmas), torch.zeros_like(rgbs)
        ext("raymarching").composite_rays_train_backward(g_ws.contiguous(), g_image.contiguous(), sigmas, rgbs, deltas, rays, ws, image, M, N, gs, gr)
        return gs, gr, None, None


class RefField:
    """The reference's torch-variant NeRFNetwork (network.py:10-143) over its CUDA encoders, with given parameters."""

    def __init__(self, embeddings, offsets, per_level_scale, sigma_w, color_w, bound):
        self.embeddings, self.offsets, self.pls, self.bound = embeddings, offsets, per_level_scale, bound
        self.sigma_w, self.color_w = sigma_w, color_w

    def params(self):
        return [self.embeddings] + list(self.sigma_w) + list(self.color_w)

    def density(self, x):
        x01 = (x + self.bound) / (2 * self.bound)
        h = _RefGridEncode.apply(x01, self.embeddings, self.offsets, self.pls, 16, x.requires_grad)
        for i, w in enumerate(self.sigma_w):
            h = F.linear(h, w)
            if i != len(self.sigma_w) - 1:
                h = F.relu(h, inplace=True)
        return _RefTruncExp.apply(h[..., 0]), h[..., 1:]

    def __call__(self, x, d):
        sigma, geo = self.density(x)
        h = torch.cat([_RefSH.apply(d, 4, d.requires_grad), geo], dim=-1)
        for i, w in enumerate(self.color_w):
            h = F.linear(h, w)
            if i != len(self.color_w) - 1:
                h = F.relu(h, inplace=True)
        return sigma, torch.sigmoid(h)


def ref_march_train(rays_o, rays_d, bound, bitfield, C, H, nears, fars, max_steps=1024):
    """raymarching.py:161-230 with force_all_rays: zero-filled N*max_steps buffers, then slice by counter[0].item()."""
    N = rays_o.shape[0]
    M = N * max_steps
    dev = rays_o.device
    xyzs = torch.zeros(M, 3, device=dev); dirs = torch.zeros(M, 3, device=dev); deltas = torch.zeros(M, 2, device=dev)
    rays = torch.empty(N, 3, dtype=torch.int32, device=dev)
    counter = torch.zeros(2, dtype=torch.int32, device=dev)
    ext("raymarching").march_rays_train(rays_o, rays_d, bitfield, bound, 0.0, max_steps, N, C, H, M, nears, fars, xyzs, dirs, deltas, rays, counter, 0)
    m = counter[0].item()
    return xyzs[:m], dirs[:m], deltas[:m], rays


def ref_render_train(field, rays_o, rays_d, bitfield, C, H, aabb, min_near, bg_color=1.0):
    """renderer.py:257-323 (train branch)."""
    N = rays_o.shape[0]
    nears = torch.empty(N, device=rays_o.device); fars = torch.empty(N, device=rays_o.device)
    ext("raymarching").near_far_from_aabb(rays_o, rays_d, aabb, N, min_near, nears, fars)
    xyzs, dirs, deltas, rays = ref_march_train(rays_o, rays_d, field.bound, bitfield, C, H, nears, fars)
    sigmas, rgbs = field(xyzs, dirs)
    ws, depth, image = _RefComposite.apply(sigmas, rgbs, deltas, rays)
    return image + (1 - ws).unsqueeze(-1) * bg_color, xyzs.shape[0]


def ref_run(field, rays_o, rays_d, aabb, min_near, num_steps=512, bg_color=1.0, density_scale=1.0):
    """renderer.py:125-254 with upsample_steps = 0, perturb = False: the cuda_ray=False sampler the pose estimator uses
    under `simulate.py -O` — uniform depths between the AABB hits, density() on every sample, colour on the samples with
    weight > 1e-4 (network.py:163-191), alpha compositing in torch.  Its autograd reaches rays_o / rays_d through the
    reference's CUDA GridEncoder (dy_dx) and SHEncoder: the like-for-like comparator for a photometric loss fwd + bwd to
    the pose."""
    N = rays_o.shape[0]
    dev = rays_o.device
    nears = torch.empty(N, device=dev); fars = torch.empty(N, device=dev)
    with torch.no_grad():
        ext("raymarching").near_far_from_aabb(rays_o.detach().contiguous(), rays_d.detach().contiguous(), aabb, N, min_near, nears, fars)
    nears = nears.unsqueeze(-1); fars = fars.unsqueeze(-1)
    z = torch.linspace(0.0, 1.0, num_steps, device=dev).unsqueeze(0).expand(N, num_steps)
    z = nears + (fars - nears) * z
    sample_dist = (fars - nears) / num_steps
    xyzs = rays_o.unsqueeze(-2) + rays_d.unsqueeze(-2) * z.unsqueeze(-1)
    xyzs = torch.min(torch.max(xyzs, aabb[:3]), aabb[3:])
    sigma, geo = field.density(xyzs.reshape(-1, 3))
    sigma = sigma.view(N, num_steps)
    deltas = torch.cat([z[..., 1:] - z[..., :-1], sample_dist * torch.ones_like(z[..., :1])], dim=-1)
    alphas = 1 - torch.exp(-deltas * density_scale * sigma)
    shifted = torch.cat([torch.ones_like(alphas[..., :1]), 1 - alphas + 1e-15], dim=-1)
    weights = alphas * torch.cumprod(shifted, dim=-1)[..., :-1]
    mask = (weights > 1e-4).reshape(-1)
    dirs = rays_d.view(-1, 1, 3).expand_as(xyzs).reshape(-1, 3)
    rgbs = torch.zeros(N * num_steps, 3, device=dev)
    if mask.any():
        h = torch.cat([_RefSH.apply(dirs[mask], 4, dirs.requires_grad), geo[mask]], dim=-1)
        for i, w in enumerate(field.color_w):
            h = F.linear(h, w)
            if i != len(field.color_w) - 1:
                h = F.relu(h, inplace=True)
        rgbs[mask] = torch.sigmoid(h)
    rgbs = rgbs.view(N, num_steps, 3)
    ws = weights.sum(-1)
    depth = (weights * ((z - nears) / (fars - nears)).clamp(0, 1)).sum(-1)
    image = (weights.unsqueeze(-1) * rgbs).sum(-2) + (1 - ws).unsqueeze(-1) * bg_color
    return image, depth


@torch.no_grad()
def ref_render_infer(field, rays_o, rays_d, bitfield, C, H, aabb, min_near, bg_color=1.0, max_steps=1024, density_scale=1.0):
    """renderer.py:325-374 (eval branch): the alive-ray loop of march_rays -> forward(x, d) -> composite_rays with the
    host-side compaction (one device->host sync per iteration for the alive count).  Returns image, depth, samples evaluated."""
    N = rays_o.shape[0]
    dev = rays_o.device
    R = ext("raymarching")
    nears = torch.empty(N, device=dev); fars = torch.empty(N, device=dev)
    R.near_far_from_aabb(rays_o, rays_d, aabb, N, min_near, nears, fars)
    ws = torch.zeros(N, device=dev); depth = torch.zeros(N, device=dev); image = torch.zeros(N, 3, device=dev)
    alive = torch.arange(N, dtype=torch.int32, device=dev)
    rays_t = nears.clone()
    step, evaluated = 0, 0
    while step < max_steps:
        n_alive = alive.shape[0]
        if n_alive <= 0:
            break
        n_step = max(min(N // n_alive, 8), 1)
        M = n_alive * n_step
        M += 128 - (M % 128)
        xyzs = torch.zeros(M, 3, device=dev); dirs = torch.zeros(M, 3, device=dev); deltas = torch.zeros(M, 2, device=dev)
        R.march_rays(n_alive, n_step, alive, rays_t, rays_o, rays_d, field.bound, 0.0, max_steps, C, H, bitfield, nears, fars,
                     xyzs, dirs, deltas, 0)
        sigmas, rgbs = field(xyzs, dirs)
        sigmas = density_scale * sigmas
        R.composite_rays(n_alive, n_step, alive, rays_t, sigmas.contiguous(), rgbs.contiguous(), deltas, ws, depth, image)
        alive = alive[alive >= 0]
        evaluated += M
        step += n_step
    image = image + (1 - ws).unsqueeze(-1) * bg_color
    depth = torch.clamp(depth - nears, min=0) / (fars - nears)
    return image, depth, evaluated
