"""NeRFRenderer for the B200 kernels — the kept API of the reference's nerf/renderer.py:61-575 (constructor kwargs, buffers
`aabb_train / aabb_infer / density_grid / density_bitfield / step_counter`, and the kwargs contracts of `run`, `run_cuda`,
`render`, `update_extra_state`, `mark_untrained_grid`), organised around what runs on the GPU:

* cuda_ray = True, eval: ONE persistent kernel renders the rays (`field.render_rays_infer`, csrc/render5.cu) — the
  reference's alive-ray loop of march_rays / forward / composite_rays with a host sync per iteration (:325-374) is kept as
  `_render_eval_loop` for the cases the kernel does not cover (perturb, other network shapes);
* cuda_ray = True, train: near/far -> march_rays_train -> fused field (+ colour) -> composite_rays_train (:282-323), all
  differentiable to the rays (the estimator's pose gradient) and to the parameters;
* cuda_ray = False: the uniform (+ importance) sampler in torch over density() / color() (:125-254), whose density() is
  the fused field query;
* occupancy upkeep (`update_extra_state` :446-539, `mark_untrained_grid` :381-444): cell positions, the draw of the cells to
  refresh, the EMA-max / mean and the frustum test are kernels (csrc/occgrid.cu); the densities come from density().

Out of scope: the background model (bg_radius > 0) and plot_pointcloud (trimesh viewer).
"""
import math

import numpy as np
import torch
import torch.nn as nn

from .. import _lib, raymarching


def sample_pdf(bins, weights, n_samples, det=False):
    """Draw `n_samples` depths per ray from the piecewise-constant density `weights` [B, T-1] over `bins` [B, T] by
    inverting its CDF (renderer.py:13-48; `det` = evenly spaced quantiles instead of uniform noise)."""
    pdf = weights + 1e-5
    pdf = pdf / pdf.sum(-1, keepdim=True)
    cdf = torch.nn.functional.pad(pdf.cumsum(-1), (1, 0))                       # [B, T], starts at 0
    shape = (*cdf.shape[:-1], n_samples)
    if det:
        u = torch.linspace(0.5 / n_samples, 1.0 - 0.5 / n_samples, n_samples, device=cdf.device).expand(shape)
    else:
        u = torch.rand(shape, device=cdf.device)
    u = u.contiguous()
    hi = torch.searchsorted(cdf, u, right=True)
    lo = (hi - 1).clamp_min(0)
    hi = hi.clamp_max(cdf.shape[-1] - 1)
    c0, c1 = cdf.gather(-1, lo), cdf.gather(-1, hi)
    b0, b1 = bins.gather(-1, lo), bins.gather(-1, hi)
    width = c1 - c0
    width = torch.where(width < 1e-5, torch.ones_like(width), width)            # empty bin: take its left edge
    return b0 + (u - c0) / width * (b1 - b0)


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
        box = torch.tensor([-bound] * 3 + [bound] * 3, dtype=torch.float32)
        self.register_buffer("aabb_train", box)
        self.register_buffer("aabb_infer", box.clone())
        self.cuda_ray = cuda_ray
        if cuda_ray:
            cells = self.grid_size ** 3
            self.register_buffer("density_grid", torch.zeros(self.cascade, cells))
            self.register_buffer("density_bitfield", torch.zeros(self.cascade * cells // 8, dtype=torch.uint8))
            self.register_buffer("step_counter", torch.zeros(16, 2, dtype=torch.int32))   # last 16 (samples, rays) counts
            self.mean_density = 0
            self.iter_density = 0
            self.mean_count = 0
            self.local_step = 0

    def forward(self, x, d):
        raise NotImplementedError()

    def density(self, x):
        raise NotImplementedError()

    def color(self, x, d, mask=None, **kwargs):
        raise NotImplementedError()

    def reset_extra_state(self):
        if self.cuda_ray:
            self.density_grid.zero_()
            self.step_counter.zero_()
            self.mean_density = self.iter_density = self.mean_count = self.local_step = 0

    def _box(self):
        return self.aabb_train if self.training else self.aabb_infer

    # ------------------------------------------------------------------ cuda_ray = False: sampler in torch
    def _weights(self, z, last_gap, sigma):
        """Alpha-compositing weights of the samples at depths z [N, T] (renderer.py:206-210); gaps, weights."""
        gaps = torch.cat([z[:, 1:] - z[:, :-1], last_gap.expand(z.shape[0], 1)], -1)
        alpha = 1 - torch.exp(-gaps * self.density_scale * sigma)
        through = torch.cat([torch.ones_like(alpha[:, :1]), 1 - alpha + 1e-15], -1).cumprod(-1)[:, :-1]
        return gaps, alpha * through

    def _points(self, rays_o, rays_d, z, box):
        p = rays_o[:, None, :] + rays_d[:, None, :] * z[..., None]
        return torch.min(torch.max(p, box[:3]), box[3:])

    def run(self, rays_o, rays_d, num_steps=128, upsample_steps=128, bg_color=None, perturb=False, **kwargs):
        lead = rays_o.shape[:-1]
        rays_o = rays_o.contiguous().view(-1, 3)
        rays_d = rays_d.contiguous().view(-1, 3)
        N = rays_o.shape[0]
        box = self._box()
        with torch.no_grad():
            near, far = raymarching.near_far_from_aabb(rays_o, rays_d, box, self.min_near)
        near, far = near[:, None], far[:, None]
        span = far - near
        gap = span / num_steps
        z = near + span * torch.linspace(0.0, 1.0, num_steps, device=rays_o.device)[None, :].expand(N, num_steps)
        if perturb:
            z = z + (torch.rand_like(z) - 0.5) * gap
        pts = self._points(rays_o, rays_d, z, box)
        field = {k: v.view(N, num_steps, -1) for k, v in self.density(pts.reshape(-1, 3)).items()}
        if upsample_steps > 0:   # importance samples where the coarse pass puts its weight, merged in depth order
            with torch.no_grad():
                gaps, w = self._weights(z, gap, field["sigma"].squeeze(-1))
                mids = z[:, :-1] + 0.5 * gaps[:, :-1]
                z_new = sample_pdf(mids, w[:, 1:-1], upsample_steps, det=not self.training).detach()
                pts_new = self._points(rays_o, rays_d, z_new, box)
            extra = {k: v.view(N, upsample_steps, -1) for k, v in self.density(pts_new.reshape(-1, 3)).items()}
            z, order = torch.cat([z, z_new], 1).sort(1)
            pts = torch.cat([pts, pts_new], 1).gather(1, order[..., None].expand(-1, -1, 3))
            for k in field:
                both = torch.cat([field[k], extra[k]], 1)
                field[k] = both.gather(1, order[..., None].expand_as(both))
        _, w = self._weights(z, gap, field["sigma"].squeeze(-1))
        flat = {k: v.reshape(-1, v.shape[-1]) for k, v in field.items()}
        dirs = rays_d[:, None, :].expand_as(pts).reshape(-1, 3)
        rgb = self.color(pts.reshape(-1, 3), dirs, mask=(w > 1e-4).reshape(-1), **flat).view(N, -1, 3)
        opacity = w.sum(-1)
        depth = (w * ((z - near) / span).clamp(0, 1)).sum(-1)
        image = (w[..., None] * rgb).sum(-2) + (1 - opacity)[:, None] * (1 if bg_color is None else bg_color)
        return {"depth": depth.view(*lead), "image": image.view(*lead, 3), "weights_sum": opacity}

    # ------------------------------------------------------------------ cuda_ray = True: occupancy-grid marcher
    def _fused_render_supported(self, rays_o, rays_d, perturb):
        """The fused inference renderer covers nav's model (NeRFNetwork: 16x2 hash grid, 64-wide torch MLPs, SH 4) with a
        frozen field, fp32, no perturbation; anything else takes the op-by-op loop."""
        if perturb or not getattr(self, "fused_eval", False) or not getattr(self, "fused_render", True):
            return False
        if torch.is_grad_enabled() and (rays_o.requires_grad or rays_d.requires_grad):
            return False
        try:
            from ..field import fused_color_supported
            return (rays_o.is_cuda and getattr(self, "num_layers", 0) == 2 and hasattr(self.encoder, "embeddings")
                    and hasattr(self, "color_net") and fused_color_supported(self, rays_o, rays_d))
        except AttributeError:
            return False

    def _finish(self, image, depth, opacity, near, far, bg_color):
        return image + (1 - opacity)[:, None] * bg_color, torch.clamp(depth - near, min=0) / (far - near)

    def _render_train(self, rays_o, rays_d, near, far, bg_color, perturb, force_all_rays, dt_gamma, max_steps):
        counter = self.step_counter[self.local_step % 16]
        counter.zero_()
        self.local_step += 1
        xyzs, dirs, deltas, rays = raymarching.march_rays_train(
            rays_o, rays_d, self.bound, self.density_bitfield, self.cascade, self.grid_size, near, far, counter,
            self.mean_count, perturb, 128, force_all_rays, dt_gamma, max_steps)
        sigma, rgb = self(xyzs, dirs)
        opacity, depth, image = raymarching.composite_rays_train(self.density_scale * sigma, rgb, deltas, rays)
        return (*self._finish(image, depth, opacity, near, far, bg_color), opacity)

    def _render_eval_loop(self, rays_o, rays_d, near, far, bg_color, perturb, dt_gamma, max_steps):
        N, dev = rays_o.shape[0], rays_o.device
        opacity = torch.zeros(N, device=dev)
        depth = torch.zeros(N, device=dev)
        image = torch.zeros(N, 3, device=dev)
        alive = torch.arange(N, dtype=torch.int32, device=dev)
        t = near.clone()
        done = 0
        while done < max_steps and alive.shape[0] > 0:
            n_alive = alive.shape[0]
            chunk = max(min(N // n_alive, 8), 1)          # more samples per round as rays die
            xyzs, dirs, deltas = raymarching.march_rays(n_alive, chunk, alive, t, rays_o, rays_d, self.bound,
                                                        self.density_bitfield, self.cascade, self.grid_size, near, far, 128,
                                                        perturb, dt_gamma, max_steps)
            sigma, rgb = self(xyzs, dirs)
            raymarching.composite_rays(n_alive, chunk, alive, t, self.density_scale * sigma, rgb, deltas, opacity, depth, image)
            alive = raymarching.compact_alive(alive)
            done += chunk
        return (*self._finish(image, depth, opacity, near, far, bg_color), opacity)

    def run_cuda(self, rays_o, rays_d, dt_gamma=0, bg_color=None, perturb=False, force_all_rays=False, max_steps=1024,
                 **kwargs):
        lead = rays_o.shape[:-1]
        rays_o = rays_o.contiguous().view(-1, 3)
        rays_d = rays_d.contiguous().view(-1, 3)
        box = self._box()
        bg = 1 if bg_color is None else bg_color
        out = {}
        if not self.training and self._fused_render_supported(rays_o, rays_d, perturb):
            from ..field import render_rays_infer
            image, depth, _ = render_rays_infer(self, rays_o, rays_d, box, bg, dt_gamma, max_steps)
        else:
            with torch.no_grad():
                near, far = raymarching.near_far_from_aabb(rays_o, rays_d, box, self.min_near)
            if self.training:
                image, depth, out["weights_sum"] = self._render_train(rays_o, rays_d, near, far, bg, perturb, force_all_rays,
                                                                      dt_gamma, max_steps)
            else:
                image, depth, _ = self._render_eval_loop(rays_o, rays_d, near, far, bg, perturb, dt_gamma, max_steps)
        out["depth"] = depth.view(*lead)
        out["image"] = image.view(*lead, 3)
        return out

    # ------------------------------------------------------------------ occupancy upkeep
    @torch.no_grad()
    def mark_untrained_grid(self, poses, intrinsic, S=64):
        """Cells no training camera sees get density -1: never refreshed, never occupied (renderer.py:381-444).  poses
        [B,4,4] cam2world (tensor or ndarray), intrinsic (fx, fy, cx, cy).  One kernel over all cells and cameras."""
        if not self.cuda_ray:
            return
        if isinstance(poses, np.ndarray):
            poses = torch.from_numpy(poses)
        grid = self.density_grid
        if not grid.is_cuda:
            raise RuntimeError("mark_untrained_grid: the occupancy grid must be on a CUDA device (no CPU path)")
        poses = poses.to(device=grid.device, dtype=torch.float32).contiguous()
        fx, fy, cx, cy = (float(v) for v in intrinsic)
        with torch.cuda.device(grid.device):
            _lib.check(_lib.lib().ngp_occ_grid_mark_untrained(_lib.ptr(poses), poses.shape[0], fx, fy, cx, cy, self.cascade,
                                                              self.grid_size, float(self.bound), _lib.ptr(grid),
                                                              _lib.stream()), "mark_untrained_grid")

    def _refresh_seed(self):
        return (int(self.iter_density) * 2654435761 + 12345) & 0xFFFFFFFF

    @torch.no_grad()
    def _refresh_full(self, decay, sums, jitter=1.0):
        """Every cell of every cascade (the first 16 refreshes, renderer.py:457-486)."""
        L = _lib.lib()
        cells = self.grid_size ** 3
        xyz = torch.empty(cells, 3, device=self.density_grid.device)
        for cas in range(self.cascade):
            _lib.check(L.ngp_occ_grid_positions(cas, self.grid_size, float(self.bound), float(jitter), self._refresh_seed(),
                                                _lib.ptr(xyz), _lib.stream()), "occ_grid_positions")
            sigma = self.density(xyz)["sigma"].reshape(-1).float().contiguous()
            _lib.check(L.ngp_occ_grid_update(self.density_grid[cas].data_ptr(), _lib.ptr(sigma), cells, float(decay),
                                             float(self.density_scale), sums[cas].data_ptr(), _lib.stream()), "occ_grid_update")

    @torch.no_grad()
    def _refresh_partial(self, decay, sums, jitter=1.0):
        """A quarter of the cells at random plus as many draws from the occupied ones, per cascade (renderer.py:488-517);
        returns the drawn cell indices [cascade, 2N] (for inspection / tests)."""
        L = _lib.lib()
        dev = self.density_grid.device
        cells = self.grid_size ** 3
        N = cells // 4
        work = torch.empty(L.ngp_occ_grid_partial_workspace_bytes(self.grid_size), dtype=torch.uint8, device=dev)
        picked = torch.empty(self.cascade, 2 * N, dtype=torch.int32, device=dev)
        xyz = torch.empty(2 * N, 3, device=dev)
        tmp = torch.empty(cells, device=dev)
        for cas in range(self.cascade):
            _lib.check(L.ngp_occ_grid_partial_select(self.density_grid[cas].data_ptr(), cas, self.grid_size, float(self.bound),
                                                     float(jitter), self._refresh_seed(), N, _lib.ptr(work),
                                                     picked[cas].data_ptr(), _lib.ptr(xyz), _lib.stream()), "occ_grid_partial_select")
            sigma = self.density(xyz)["sigma"].reshape(-1).float().contiguous()
            _lib.check(L.ngp_occ_grid_scatter_update(self.density_grid[cas].data_ptr(), picked[cas].data_ptr(), _lib.ptr(sigma),
                                                     2 * N, cells, float(decay), float(self.density_scale), _lib.ptr(tmp),
                                                     sums[cas].data_ptr(), _lib.stream()), "occ_grid_scatter_update")
        return picked

    @torch.no_grad()
    def update_extra_state(self, decay=0.95, S=128):
        """Refresh the occupancy grid and its bitfield (renderer.py:446-539): grid <- max(grid * decay, density) on the
        queried cells (all of them for the first 16 calls, then the partial draw), threshold min(mean density,
        density_thresh), pack the bits; then fold the marcher's sample counters into mean_count."""
        if not self.cuda_ray:
            return
        if not self.density_grid.is_cuda:
            raise RuntimeError("update_extra_state: the occupancy grid must be on a CUDA device (no CPU path)")
        L = _lib.lib()
        cells = self.grid_size ** 3
        sums = torch.empty(self.cascade, L.ngp_occ_grid_update_blocks(cells), device=self.density_grid.device)
        with torch.cuda.device(self.density_grid.device):
            if self.iter_density < 16:
                self._refresh_full(decay, sums)
            else:
                self._refresh_partial(decay, sums)
        self.mean_density = float(sums.double().sum().item() / (self.cascade * cells))
        self.iter_density += 1
        self.density_bitfield = raymarching.packbits(self.density_grid, min(self.mean_density, self.density_thresh),
                                                     self.density_bitfield)
        recent = min(16, self.local_step)
        if recent > 0:
            self.mean_count = int(self.step_counter[:recent, 0].sum().item() / recent)
        self.local_step = 0

    # kept name of the round-1 entry point (bench / tests call it to force a full sweep without jitter)
    @torch.no_grad()
    def _update_extra_state_sweep(self, decay, jitter=1.0):
        cells = self.grid_size ** 3
        sums = torch.empty(self.cascade, _lib.lib().ngp_occ_grid_update_blocks(cells), device=self.density_grid.device)
        with torch.cuda.device(self.density_grid.device):
            self._refresh_full(decay, sums, jitter)
        self.mean_density = float(sums.double().sum().item() / (self.cascade * cells))
        self.iter_density += 1
        self.density_bitfield = raymarching.packbits(self.density_grid, min(self.mean_density, self.density_thresh),
                                                     self.density_bitfield)
        recent = min(16, self.local_step)
        if recent > 0:
            self.mean_count = int(self.step_counter[:recent, 0].sum().item() / recent)
        self.local_step = 0

    # ------------------------------------------------------------------ dispatcher
    def render(self, rays_o, rays_d, staged=False, max_ray_batch=4096, **kwargs):
        """rays_o, rays_d [B, N, 3] -> {'image': [B, N, 3], 'depth': [B, N]} (renderer.py:542-575).  The torch sampler is
        run in batches of max_ray_batch rays when `staged` (its per-sample tensors are large); the marcher never is."""
        if self.cuda_ray:
            return self.run_cuda(rays_o, rays_d, **kwargs)
        if not staged:
            return self.run(rays_o, rays_d, **kwargs)
        B, N = rays_o.shape[:2]
        depth = torch.empty(B, N, device=rays_o.device)
        image = torch.empty(B, N, 3, device=rays_o.device)
        for b in range(B):
            for a in range(0, N, max_ray_batch):
                part = self.run(rays_o[b:b + 1, a:a + max_ray_batch], rays_d[b:b + 1, a:a + max_ray_batch], **kwargs)
                depth[b:b + 1, a:a + max_ray_batch] = part["depth"]
                image[b:b + 1, a:a + max_ray_batch] = part["image"]
        return {"depth": depth, "image": image}
