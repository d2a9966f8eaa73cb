"""NeRFRenderer — host-side mirror of the reference's nerf/renderer.py:61-575 (same ctor kwargs, buffers, kwargs contract of
run / run_cuda / render / update_extra_state), on top of the B200 ray kernels in ..raymarching.

Kept semantics: `run` = uniform sampler through density()/color() (renderer.py:125-254); `run_cuda` train branch =
near_far -> march_rays_train -> field -> composite_rays_train (:282-323), eval branch = the alive-ray loop of
march_rays / field / composite_rays (:325-374); `render` = dispatcher + max_ray_batch staging (:542-575);
`update_extra_state` = occupancy-grid refresh + packbits (:446-539).  `sample_pdf` (:13-48) supports upsample_steps > 0.
Out of scope: bg_radius > 0 background sphere, plot_pointcloud, mark_untrained_grid's camera culling is kept (pure torch).
"""
import math

import numpy as np
import torch
import torch.nn as nn

from .. import raymarching


def sample_pdf(bins, weights, n_samples, det=False):
    """Inverse-CDF sampling of `n_samples` depths per ray from piecewise-constant `weights` (renderer.py:13-48)."""
    weights = weights + 1e-5
    pdf = weights / torch.sum(weights, -1, keepdim=True)
    cdf = torch.cumsum(pdf, -1)
    cdf = torch.cat([torch.zeros_like(cdf[..., :1]), cdf], -1)
    if det:
        u = torch.linspace(0.5 / n_samples, 1.0 - 0.5 / n_samples, steps=n_samples, device=weights.device)
        u = u.expand(list(cdf.shape[:-1]) + [n_samples])
    else:
        u = torch.rand(list(cdf.shape[:-1]) + [n_samples], device=weights.device)
    u = u.contiguous()
    inds = torch.searchsorted(cdf, u, right=True)
    below = torch.max(torch.zeros_like(inds - 1), inds - 1)
    above = torch.min((cdf.shape[-1] - 1) * torch.ones_like(inds), inds)
    inds_g = torch.stack([below, above], -1)
    matched_shape = [inds_g.shape[0], inds_g.shape[1], cdf.shape[-1]]
    cdf_g = torch.gather(cdf.unsqueeze(1).expand(matched_shape), 2, inds_g)
    bins_g = torch.gather(bins.unsqueeze(1).expand(matched_shape), 2, inds_g)
    denom = cdf_g[..., 1] - cdf_g[..., 0]
    denom = torch.where(denom < 1e-5, torch.ones_like(denom), denom)
    t = (u - cdf_g[..., 0]) / denom
    return bins_g[..., 0] + t * (bins_g[..., 1] - bins_g[..., 0])


class NeRFRenderer(nn.Module):
    def __init__(self, bound=1, cuda_ray=False, density_scale=1, min_near=0.2, density_thresh=0.01, bg_radius=-1):
        super().__init__()
        self.bound = bound
        self.cascade = 1 + math.ceil(math.log2(bound))
        self.grid_size = 128
        self.density_scale = density_scale
        self.min_near = min_near
        self.density_thresh = density_thresh
        self.bg_radius = bg_radius
        aabb_train = torch.FloatTensor([-bound, -bound, -bound, bound, bound, bound])
        self.register_buffer("aabb_train", aabb_train)
        self.register_buffer("aabb_infer", aabb_train.clone())
        self.cuda_ray = cuda_ray
        if cuda_ray:
            self.register_buffer("density_grid", torch.zeros([self.cascade, self.grid_size ** 3]))
            self.register_buffer("density_bitfield",
                                 torch.zeros(self.cascade * self.grid_size ** 3 // 8, dtype=torch.uint8))
            self.mean_density = 0
            self.iter_density = 0
            self.register_buffer("step_counter", torch.zeros(16, 2, dtype=torch.int32))
            self.mean_count = 0
            self.local_step = 0

    def forward(self, x, d):
        raise NotImplementedError()

    def density(self, x):
        raise NotImplementedError()

    def color(self, x, d, mask=None, **kwargs):
        raise NotImplementedError()

    def _fused_render_supported(self, rays_o, rays_d, perturb):
        """The fused inference renderer covers nav's model (NeRFNetwork: 16x2 hash grid, 64-wide torch MLPs, SH 4) with a
        frozen field, fp32, no perturbation; anything else takes the op-by-op loop below."""
        if perturb or not getattr(self, "fused_eval", False) or not getattr(self, "fused_render", True) or torch.is_grad_enabled() and (
                rays_o.requires_grad or rays_d.requires_grad):
            return False
        try:
            from ..field import fused_color_supported
            return (rays_o.is_cuda and getattr(self, "num_layers", 0) == 2 and hasattr(self.encoder, "embeddings")
                    and hasattr(self, "color_net") and fused_color_supported(self, rays_o, rays_d))
        except AttributeError:
            return False

    def reset_extra_state(self):
        if not self.cuda_ray:
            return
        self.density_grid.zero_()
        self.mean_density = 0
        self.iter_density = 0
        self.step_counter.zero_()
        self.mean_count = 0
        self.local_step = 0

    # ------------------------------------------------------------------ torch sampler (cuda_ray=False)
    def _alpha_weights(self, z_vals, sample_dist, sigma):
        deltas = z_vals[..., 1:] - z_vals[..., :-1]
        deltas = torch.cat([deltas, sample_dist * torch.ones_like(deltas[..., :1])], dim=-1)
        alphas = 1 - torch.exp(-deltas * self.density_scale * sigma)
        alphas_shifted = torch.cat([torch.ones_like(alphas[..., :1]), 1 - alphas + 1e-15], dim=-1)
        return deltas, alphas * torch.cumprod(alphas_shifted, dim=-1)[..., :-1]

    def run(self, rays_o, rays_d, num_steps=128, upsample_steps=128, bg_color=None, perturb=False, **kwargs):
        prefix = rays_o.shape[:-1]
        rays_o = rays_o.contiguous().view(-1, 3)
        rays_d = rays_d.contiguous().view(-1, 3)
        N = rays_o.shape[0]
        device = rays_o.device
        aabb = self.aabb_train if self.training else self.aabb_infer
        with torch.no_grad():
            nears, fars = raymarching.near_far_from_aabb(rays_o, rays_d, aabb, self.min_near)
        nears = nears.unsqueeze(-1)
        fars = fars.unsqueeze(-1)
        z_vals = torch.linspace(0.0, 1.0, num_steps, device=device).unsqueeze(0).expand((N, num_steps))
        z_vals = nears + (fars - nears) * z_vals
        sample_dist = (fars - nears) / num_steps
        if perturb:
            z_vals = z_vals + (torch.rand(z_vals.shape, device=device) - 0.5) * sample_dist
        xyzs = rays_o.unsqueeze(-2) + rays_d.unsqueeze(-2) * z_vals.unsqueeze(-1)
        xyzs = torch.min(torch.max(xyzs, aabb[:3]), aabb[3:])
        density_outputs = self.density(xyzs.reshape(-1, 3))
        for k, v in density_outputs.items():
            density_outputs[k] = v.view(N, num_steps, -1)
        if upsample_steps > 0:
            with torch.no_grad():
                deltas, weights = self._alpha_weights(z_vals, sample_dist, density_outputs["sigma"].squeeze(-1))
                z_vals_mid = z_vals[..., :-1] + 0.5 * deltas[..., :-1]
                new_z_vals = sample_pdf(z_vals_mid, weights[:, 1:-1], upsample_steps, det=not self.training).detach()
                new_xyzs = rays_o.unsqueeze(-2) + rays_d.unsqueeze(-2) * new_z_vals.unsqueeze(-1)
                new_xyzs = torch.min(torch.max(new_xyzs, aabb[:3]), aabb[3:])
            new_density_outputs = self.density(new_xyzs.reshape(-1, 3))
            for k, v in new_density_outputs.items():
                new_density_outputs[k] = v.view(N, upsample_steps, -1)
            z_vals = torch.cat([z_vals, new_z_vals], dim=1)
            z_vals, z_index = torch.sort(z_vals, dim=1)
            xyzs = torch.cat([xyzs, new_xyzs], dim=1)
            xyzs = torch.gather(xyzs, dim=1, index=z_index.unsqueeze(-1).expand_as(xyzs))
            for k in density_outputs:
                tmp = torch.cat([density_outputs[k], new_density_outputs[k]], dim=1)
                density_outputs[k] = torch.gather(tmp, dim=1, index=z_index.unsqueeze(-1).expand_as(tmp))
        _, weights = self._alpha_weights(z_vals, sample_dist, density_outputs["sigma"].squeeze(-1))
        dirs = rays_d.view(-1, 1, 3).expand_as(xyzs)
        for k, v in density_outputs.items():
            density_outputs[k] = v.view(-1, v.shape[-1])
        mask = weights > 1e-4
        rgbs = self.color(xyzs.reshape(-1, 3), dirs.reshape(-1, 3), mask=mask.reshape(-1), **density_outputs)
        rgbs = rgbs.view(N, -1, 3)
        weights_sum = weights.sum(dim=-1)
        ori_z_vals = ((z_vals - nears) / (fars - nears)).clamp(0, 1)
        depth = torch.sum(weights * ori_z_vals, dim=-1)
        image = torch.sum(weights.unsqueeze(-1) * rgbs, dim=-2)
        if bg_color is None:
            bg_color = 1
        image = image + (1 - weights_sum).unsqueeze(-1) * bg_color
        return {"depth": depth.view(*prefix), "image": image.view(*prefix, 3), "weights_sum": weights_sum}

    # ------------------------------------------------------------------ occupancy-grid ray marcher (cuda_ray=True)
    def run_cuda(self, rays_o, rays_d, dt_gamma=0, bg_color=None, perturb=False, force_all_rays=False, max_steps=1024,
                 **kwargs):
        prefix = rays_o.shape[:-1]
        rays_o = rays_o.contiguous().view(-1, 3)
        rays_d = rays_d.contiguous().view(-1, 3)
        N = rays_o.shape[0]
        device = rays_o.device
        aabb = self.aabb_train if self.training else self.aabb_infer
        with torch.no_grad():
            nears, fars = raymarching.near_far_from_aabb(rays_o, rays_d, aabb, self.min_near)
        if bg_color is None:
            bg_color = 1
        results = {}
        if self.training:
            counter = self.step_counter[self.local_step % 16]
            counter.zero_()
            self.local_step += 1
            xyzs, dirs, deltas, rays = raymarching.march_rays_train(
                rays_o, rays_d, self.bound, self.density_bitfield, self.cascade, self.grid_size, nears, fars, counter,
                self.mean_count, perturb, 128, force_all_rays, dt_gamma, max_steps)
            sigmas, rgbs = self(xyzs, dirs)
            sigmas = self.density_scale * sigmas
            weights_sum, depth, image = raymarching.composite_rays_train(sigmas, rgbs, deltas, rays)
            image = image + (1 - weights_sum).unsqueeze(-1) * bg_color
            depth = torch.clamp(depth - nears, min=0) / (fars - nears)
            results["weights_sum"] = weights_sum
        elif self._fused_render_supported(rays_o, rays_d, perturb):
            # one persistent kernel for the whole alive-ray loop + post-processing (csrc/render.cu)
            from ..field import render_rays_infer
            image, depth, weights_sum = render_rays_infer(self, rays_o, rays_d, aabb, bg_color, dt_gamma, max_steps)
        else:
            weights_sum = torch.zeros(N, dtype=torch.float32, device=device)
            depth = torch.zeros(N, dtype=torch.float32, device=device)
            image = torch.zeros(N, 3, dtype=torch.float32, device=device)
            rays_alive = torch.arange(N, dtype=torch.int32, device=device)
            rays_t = nears.clone()
            step = 0
            while step < max_steps:
                n_alive = rays_alive.shape[0]
                if n_alive <= 0:
                    break
                n_step = max(min(N // n_alive, 8), 1)
                xyzs, dirs, deltas = raymarching.march_rays(
                    n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, self.bound, self.density_bitfield, self.cascade,
                    self.grid_size, nears, fars, 128, perturb, dt_gamma, max_steps)
                sigmas, rgbs = self(xyzs, dirs)
                sigmas = self.density_scale * sigmas
                raymarching.composite_rays(n_alive, n_step, rays_alive, rays_t, sigmas, rgbs, deltas, weights_sum, depth,
                                           image)
                rays_alive = raymarching.compact_alive(rays_alive)
                step += n_step
            image = image + (1 - weights_sum).unsqueeze(-1) * bg_color
            depth = torch.clamp(depth - nears, min=0) / (fars - nears)
        results["depth"] = depth.view(*prefix)
        results["image"] = image.view(*prefix, 3)
        return results

    @torch.no_grad()
    def update_extra_state(self, decay=0.95, S=128):
        """Occupancy refresh, renderer.py:446-539: full sweep for the first 16 calls, then 1/4 random + 1/4 occupied
        cells per cascade; grid <- max(grid*decay, sigma); thresh = min(mean, density_thresh); packbits."""
        if not self.cuda_ray:
            return
        dev = self.density_bitfield.device
        G = self.grid_size
        tmp_grid = -torch.ones_like(self.density_grid)

        def query(cas, coords):
            xyzs = 2 * coords.float() / (G - 1) - 1
            bound = min(2 ** cas, self.bound)
            half = bound / G
            cas_xyzs = xyzs * (bound - half)
            cas_xyzs += (torch.rand_like(cas_xyzs) * 2 - 1) * half
            sig = self.density(cas_xyzs)["sigma"].reshape(-1).detach()
            return sig * self.density_scale

        if self.iter_density < 16 and getattr(self, "fused_occ_sweep", True) and self.density_grid.is_cuda:
            return self._update_extra_state_sweep(decay)
        if self.iter_density < 16:
            ar = torch.arange(G, dtype=torch.int32, device=dev)
            for xs in ar.split(S):
                for ys in ar.split(S):
                    for zs in ar.split(S):
                        xx, yy, zz = torch.meshgrid(xs, ys, zs, indexing="ij")
                        coords = torch.stack([xx.reshape(-1), yy.reshape(-1), zz.reshape(-1)], dim=-1)
                        indices = raymarching.morton3D(coords).long()
                        for cas in range(self.cascade):
                            tmp_grid[cas, indices] = query(cas, coords)
        else:
            N = G ** 3 // 4
            for cas in range(self.cascade):
                coords = torch.randint(0, G, (N, 3), device=dev)
                indices = raymarching.morton3D(coords).long()
                occ_indices = torch.nonzero(self.density_grid[cas] > 0).squeeze(-1)
                rand_mask = torch.randint(0, occ_indices.shape[0], [N], dtype=torch.long, device=dev)
                occ_indices = occ_indices[rand_mask]
                occ_coords = raymarching.morton3D_invert(occ_indices)
                indices = torch.cat([indices, occ_indices], dim=0)
                coords = torch.cat([coords, occ_coords], dim=0)
                tmp_grid[cas, indices] = query(cas, coords)
        valid = (self.density_grid >= 0) & (tmp_grid >= 0)
        self.density_grid[valid] = torch.maximum(self.density_grid[valid] * decay, tmp_grid[valid])
        self.mean_density = torch.mean(self.density_grid.clamp(min=0)).item()
        self.iter_density += 1
        density_thresh = min(self.mean_density, self.density_thresh)
        self.density_bitfield = raymarching.packbits(self.density_grid, density_thresh, self.density_bitfield)
        total_step = min(16, self.local_step)
        if total_step > 0:
            self.mean_count = int(self.step_counter[:total_step, 0].sum().item() / total_step)
        self.local_step = 0

    @torch.no_grad()
    def _update_extra_state_sweep(self, decay, jitter=1.0):
        """Full-sweep branch of update_extra_state (renderer.py:457-486, 519-539) on the occupancy kernels (csrc/occgrid.cu):
        positions in Morton order -> density() (the fused field query) -> EMA-max + mean partials -> packbits."""
        from .. import _lib
        L = _lib.lib()
        G, n = self.grid_size, self.grid_size ** 3
        dev = self.density_grid.device
        xyz = torch.empty(n, 3, device=dev)
        nblk = L.ngp_occ_grid_update_blocks(n)
        sums = torch.empty(self.cascade, nblk, device=dev)
        seed = (int(self.iter_density) * 2654435761 + 12345) & 0xFFFFFFFF
        with torch.cuda.device(dev):
            for cas in range(self.cascade):
                _lib.check(L.ngp_occ_grid_positions(cas, G, float(self.bound), float(jitter), seed, _lib.ptr(xyz), _lib.stream()),
                           "occ_grid_positions")
                sigma = self.density(xyz)["sigma"].reshape(-1).float().contiguous()
                _lib.check(L.ngp_occ_grid_update(self.density_grid[cas].data_ptr(), _lib.ptr(sigma), n, float(decay),
                                                 float(self.density_scale), sums[cas].data_ptr(), _lib.stream()), "occ_grid_update")
        self.mean_density = float(sums.double().sum().item() / (self.cascade * n))
        self.iter_density += 1
        density_thresh = min(self.mean_density, self.density_thresh)
        self.density_bitfield = raymarching.packbits(self.density_grid, density_thresh, self.density_bitfield)
        total_step = min(16, self.local_step)
        if total_step > 0:
            self.mean_count = int(self.step_counter[:total_step, 0].sum().item() / total_step)
        self.local_step = 0

    def render(self, rays_o, rays_d, staged=False, max_ray_batch=4096, **kwargs):
        """renderer.py:542-575.  rays_o, rays_d: [B, N, 3] -> {'image': [B, N, 3], 'depth': [B, N]}."""
        _run = self.run_cuda if self.cuda_ray else self.run
        B, N = rays_o.shape[:2]
        device = rays_o.device
        if staged and not self.cuda_ray:
            depth = torch.empty((B, N), device=device)
            image = torch.empty((B, N, 3), device=device)
            for b in range(B):
                head = 0
                while head < N:
                    tail = min(head + max_ray_batch, N)
                    r = _run(rays_o[b:b + 1, head:tail], rays_d[b:b + 1, head:tail], **kwargs)
                    depth[b:b + 1, head:tail] = r["depth"]
                    image[b:b + 1, head:tail] = r["image"]
                    head += max_ray_batch
            return {"depth": depth, "image": image}
        return _run(rays_o, rays_d, **kwargs)
