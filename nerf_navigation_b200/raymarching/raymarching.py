"""Host-side mirror of the reference's raymarching/raymarching.py (same Function.apply aliases and positional
argument order: raymarching.py:49,104,126,155,230,286,337,362) on top of the C ABI.  Citations are into the reference.

Differences by design: `march_rays_train` is deterministic (rays in ray order, offsets = exclusive scan) where the
reference's atomics give an arbitrary permutation (raymarching.cu:409-417); `compact_alive` replaces the boolean-mask
compaction of renderer.py:365.  `sph_from_ray` is out of scope (background model unused by nav).
"""
import torch
from torch.autograd import Function

from .. import _lib

__all__ = ["near_far_from_aabb", "morton3D", "morton3D_invert", "packbits", "march_rays_train",
           "composite_rays_train", "march_rays", "composite_rays", "compact_alive"]

_fwd32 = torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
_bwd = torch.amp.custom_bwd(device_type="cuda")


def _cuda(t):
    return t if t.is_cuda else t.cuda()  # the reference wrappers silently .cuda() CPU inputs (raymarching.py:34-35)


class _near_far_from_aabb(Function):
    @staticmethod
    @_fwd32
    def forward(ctx, rays_o, rays_d, aabb, min_near=0.2):
        rays_o = _cuda(rays_o).contiguous().view(-1, 3)
        rays_d = _cuda(rays_d).contiguous().view(-1, 3)
        aabb = _cuda(aabb).contiguous().float()
        N = rays_o.shape[0]
        nears = torch.empty(N, dtype=rays_o.dtype, device=rays_o.device)
        fars = torch.empty(N, dtype=rays_o.dtype, device=rays_o.device)
        with torch.cuda.device(rays_o.device):
            _lib.check(_lib.lib().ngp_near_far_from_aabb(_lib.ptr(rays_o), _lib.ptr(rays_d), _lib.ptr(aabb), N,
                                                        float(min_near), _lib.ptr(nears), _lib.ptr(fars), _lib.stream()),
                       "near_far_from_aabb")
        return nears, fars


near_far_from_aabb = _near_far_from_aabb.apply


class _morton3D(Function):
    @staticmethod
    def forward(ctx, coords):
        coords = _cuda(coords).int().contiguous()
        N = coords.shape[0]
        indices = torch.empty(N, dtype=torch.int32, device=coords.device)
        with torch.cuda.device(coords.device):
            _lib.check(_lib.lib().ngp_morton3D(_lib.ptr(coords), N, _lib.ptr(indices), _lib.stream()), "morton3D")
        return indices


morton3D = _morton3D.apply


class _morton3D_invert(Function):
    @staticmethod
    def forward(ctx, indices):
        indices = _cuda(indices).int().contiguous()
        N = indices.shape[0]
        coords = torch.empty(N, 3, dtype=torch.int32, device=indices.device)
        with torch.cuda.device(indices.device):
            _lib.check(_lib.lib().ngp_morton3D_invert(_lib.ptr(indices), N, _lib.ptr(coords), _lib.stream()),
                       "morton3D_invert")
        return coords


morton3D_invert = _morton3D_invert.apply


class _packbits(Function):
    @staticmethod
    @_fwd32
    def forward(ctx, grid, thresh, bitfield=None):
        grid = _cuda(grid).contiguous()
        C, H3 = grid.shape
        N = C * H3 // 8
        if bitfield is None:
            bitfield = torch.empty(N, dtype=torch.uint8, device=grid.device)
        with torch.cuda.device(grid.device):
            _lib.check(_lib.lib().ngp_packbits(_lib.ptr(grid), N, float(thresh), _lib.ptr(bitfield), _lib.stream()),
                       "packbits")
        return bitfield


packbits = _packbits.apply


class _march_rays_train(Function):
    """raymarching.py:161-230 (forward only, like the reference)."""

    @staticmethod
    @_fwd32
    def forward(ctx, rays_o, rays_d, bound, density_bitfield, C, H, nears, fars, step_counter=None, mean_count=-1,
                perturb=False, align=-1, force_all_rays=False, dt_gamma=0, max_steps=1024):
        rays_o = _cuda(rays_o).contiguous().view(-1, 3)
        rays_d = _cuda(rays_d).contiguous().view(-1, 3)
        density_bitfield = _cuda(density_bitfield).contiguous()
        N = rays_o.shape[0]
        M = N * max_steps
        if not force_all_rays and mean_count > 0:
            if align > 0:
                mean_count += align - mean_count % align
            M = mean_count
        dev = rays_o.device
        rays = torch.empty(N, 3, dtype=torch.int32, device=dev)
        if step_counter is None:
            step_counter = torch.zeros(2, dtype=torch.int32, device=dev)
        L = _lib.lib()
        ws_bytes = L.ngp_march_rays_train_workspace_bytes(N)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        exact = force_all_rays or mean_count <= 0
        with torch.cuda.device(dev):
            if exact:
                # the reference zero-fills N*max_steps samples and slices after a .item() (raymarching.py:209-224);
                # here the count pass runs first so only the samples that exist are allocated.  Same result.
                _lib.check(L.ngp_march_rays_train(_lib.ptr(rays_o), _lib.ptr(rays_d), _lib.ptr(density_bitfield),
                                                  float(bound), float(dt_gamma), int(max_steps), N, int(C), int(H), 0,
                                                  _lib.ptr(nears), _lib.ptr(fars), None, None, None, _lib.ptr(rays),
                                                  _lib.ptr(step_counter), int(perturb), _lib.ptr(ws), ws_bytes,
                                                  _lib.stream()), "march_rays_train")
                m = int(step_counter[0].item())
                if align > 0:
                    m += align - m % align
                M_alloc = m
            else:
                M_alloc = M
            xyzs = torch.zeros(M_alloc, 3, dtype=rays_o.dtype, device=dev)
            dirs = torch.zeros(M_alloc, 3, dtype=rays_o.dtype, device=dev)
            deltas = torch.zeros(M_alloc, 2, dtype=rays_o.dtype, device=dev)
            if exact:   # counts, offsets and the counter are in place: only the write pass is left
                _lib.check(L.ngp_march_rays_train_write(_lib.ptr(rays_o), _lib.ptr(rays_d), _lib.ptr(density_bitfield),
                                                        float(bound), float(dt_gamma), int(max_steps), N, int(C), int(H), M,
                                                        _lib.ptr(nears), _lib.ptr(fars), _lib.ptr(xyzs), _lib.ptr(dirs),
                                                        _lib.ptr(deltas), _lib.ptr(rays), int(perturb), _lib.ptr(ws),
                                                        _lib.stream()), "march_rays_train_write")
            else:
                _lib.check(L.ngp_march_rays_train(_lib.ptr(rays_o), _lib.ptr(rays_d), _lib.ptr(density_bitfield),
                                                  float(bound), float(dt_gamma), int(max_steps), N, int(C), int(H), M_alloc,
                                                  _lib.ptr(nears), _lib.ptr(fars), _lib.ptr(xyzs), _lib.ptr(dirs),
                                                  _lib.ptr(deltas), _lib.ptr(rays), _lib.ptr(step_counter), int(perturb),
                                                  _lib.ptr(ws), ws_bytes, _lib.stream()), "march_rays_train")
        if ctx.needs_input_grad[0] or ctx.needs_input_grad[1]:
            ctx.save_for_backward(rays_o, rays_d, nears, deltas, rays)
            ctx.cfg = (float(bound), int(max_steps), int(perturb), N, M if exact else M_alloc)
        ctx.mark_non_differentiable(deltas, rays)
        return xyzs, dirs, deltas, rays

    @staticmethod
    @_bwd
    def backward(ctx, grad_xyzs, grad_dirs, grad_deltas, grad_rays):
        """EXTENSION (the reference op has no backward): dL/drays_o, dL/drays_d with the sample abscissae held
        constant, i.e. what the reference's torch sampler differentiates (renderer.py:159-160)."""
        rays_o, rays_d, nears, deltas, rays = ctx.saved_tensors
        bound, max_steps, perturb, N, M = ctx.cfg
        dev = rays_o.device
        if grad_xyzs is None:
            grad_xyzs = torch.zeros(deltas.shape[0], 3, device=dev)
        grad_xyzs = grad_xyzs.contiguous().float()
        grad_dirs = None if grad_dirs is None else grad_dirs.contiguous().float()
        go = torch.empty(N, 3, device=dev); gd = torch.empty(N, 3, device=dev)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().ngp_march_rays_train_backward(
                _lib.ptr(grad_xyzs), _lib.ptr(grad_dirs), _lib.ptr(rays_o), _lib.ptr(rays_d), _lib.ptr(nears),
                _lib.ptr(deltas), _lib.ptr(rays), bound, max_steps, perturb, N, M, _lib.ptr(go), _lib.ptr(gd),
                _lib.stream()), "march_rays_train_backward")
        return (go, gd) + (None,) * 13


march_rays_train = _march_rays_train.apply


class _composite_rays_train(Function):
    """raymarching.py:233-286."""

    @staticmethod
    @_fwd32
    def forward(ctx, sigmas, rgbs, deltas, rays):
        sigmas = sigmas.contiguous(); rgbs = rgbs.contiguous()
        M, N = sigmas.shape[0], rays.shape[0]
        dev = sigmas.device
        weights_sum = torch.empty(N, dtype=sigmas.dtype, device=dev)
        depth = torch.empty(N, dtype=sigmas.dtype, device=dev)
        image = torch.empty(N, 3, dtype=sigmas.dtype, device=dev)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().ngp_composite_rays_train_forward(
                _lib.ptr(sigmas), _lib.ptr(rgbs), _lib.ptr(deltas), _lib.ptr(rays), M, N, _lib.ptr(weights_sum),
                _lib.ptr(depth), _lib.ptr(image), _lib.stream()), "composite_rays_train_forward")
        ctx.save_for_backward(sigmas, rgbs, deltas, rays, weights_sum, depth, image)
        ctx.dims = [M, N]
        return weights_sum, depth, image

    @staticmethod
    @_bwd
    def backward(ctx, grad_weights_sum, grad_depth, grad_image):
        grad_weights_sum = grad_weights_sum.contiguous()
        grad_image = grad_image.contiguous()
        sigmas, rgbs, deltas, rays, weights_sum, depth, image = ctx.saved_tensors
        M, N = ctx.dims
        grad_sigmas = torch.zeros_like(sigmas)
        grad_rgbs = torch.zeros_like(rgbs)
        with torch.cuda.device(sigmas.device):
            _lib.check(_lib.lib().ngp_composite_rays_train_backward(
                _lib.ptr(grad_weights_sum), _lib.ptr(grad_image), _lib.ptr(sigmas), _lib.ptr(rgbs), _lib.ptr(deltas),
                _lib.ptr(rays), _lib.ptr(weights_sum), _lib.ptr(image), M, N, _lib.ptr(grad_sigmas),
                _lib.ptr(grad_rgbs), _lib.stream()), "composite_rays_train_backward")
        return grad_sigmas, grad_rgbs, None, None


composite_rays_train = _composite_rays_train.apply


class _march_rays(Function):
    """raymarching.py:292-337."""

    @staticmethod
    @_fwd32
    def forward(ctx, n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, bound, density_bitfield, C, H, near, far,
                align=-1, perturb=False, dt_gamma=0, max_steps=1024):
        rays_o = _cuda(rays_o).contiguous().view(-1, 3)
        rays_d = _cuda(rays_d).contiguous().view(-1, 3)
        M = n_alive * n_step
        if align > 0:
            M += align - (M % align)
        dev = rays_o.device
        xyzs = torch.zeros(M, 3, dtype=rays_o.dtype, device=dev)
        dirs = torch.zeros(M, 3, dtype=rays_o.dtype, device=dev)
        deltas = torch.zeros(M, 2, dtype=rays_o.dtype, device=dev)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().ngp_march_rays(
                n_alive, n_step, _lib.ptr(rays_alive), _lib.ptr(rays_t), _lib.ptr(rays_o), _lib.ptr(rays_d), float(bound),
                float(dt_gamma), int(max_steps), int(C), int(H), _lib.ptr(density_bitfield), _lib.ptr(near),
                _lib.ptr(far), _lib.ptr(xyzs), _lib.ptr(dirs), _lib.ptr(deltas), int(perturb), _lib.stream()),
                "march_rays")
        return xyzs, dirs, deltas


march_rays = _march_rays.apply


class _composite_rays(Function):
    """raymarching.py:340-362 (in place on rays_alive, rays_t, weights_sum, depth, image)."""

    @staticmethod
    @_fwd32
    def forward(ctx, n_alive, n_step, rays_alive, rays_t, sigmas, rgbs, deltas, weights_sum, depth, image):
        sigmas = sigmas.contiguous(); rgbs = rgbs.contiguous()
        with torch.cuda.device(sigmas.device):
            _lib.check(_lib.lib().ngp_composite_rays(
                n_alive, n_step, _lib.ptr(rays_alive), _lib.ptr(rays_t), _lib.ptr(sigmas), _lib.ptr(rgbs),
                _lib.ptr(deltas), _lib.ptr(weights_sum), _lib.ptr(depth), _lib.ptr(image), _lib.stream()),
                "composite_rays")
        return tuple()


composite_rays = _composite_rays.apply


def compact_alive(rays_alive):
    """Stream compaction of the alive list (renderer.py:365: rays_alive[rays_alive >= 0])."""
    return rays_alive[rays_alive >= 0]
