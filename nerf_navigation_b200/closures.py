"""Pose -> rays -> photometric loss closure for batches of pose hypotheses.

The reference builds ALL H*W rays of one pose and then indexes the ~1k sampled pixels (nerf/utils.py:52-116 `get_rays`,
nav/estimator_helpers.py:293-327 `measurement_fn`); here only the sampled pixels' rays are generated, for a batch of
hypotheses at once, with the same conventions: pixel centres at +0.5, direction ((i-cx)/fx, (j-cy)/fy, 1) normalised,
rays_d = dir @ R^T, rays_o = t (cam2world pose [R|t]); loss = mse(render(rays).image, target) per hypothesis.
Everything is differentiable w.r.t. the poses through the march_rays_train pose-gradient extension (cuda_ray=True) or
the torch sampler (cuda_ray=False).
"""
import torch


def rays_from_poses(poses, pix, intrinsics, W):
    """poses [P,4,4] cam2world, pix [R] int64 flat pixel indices (row-major, j*W+i) shared by all hypotheses
    -> rays_o, rays_d [P,R,3]."""
    fx, fy, cx, cy = intrinsics
    i = (pix % W).to(poses.dtype) + 0.5
    j = torch.div(pix, W, rounding_mode="floor").to(poses.dtype) + 0.5
    d = torch.stack(((i - cx) / fx, (j - cy) / fy, torch.ones_like(i)), dim=-1)
    d = d / torch.norm(d, dim=-1, keepdim=True)
    rays_d = d.unsqueeze(0) @ poses[:, :3, :3].transpose(-1, -2)
    rays_o = poses[:, None, :3, 3].expand_as(rays_d)
    return rays_o, rays_d


def photometric_loss(model, poses, pix, target, intrinsics, W, **render_kwargs):
    """Per-hypothesis MSE between the render at `poses` [P,4,4] and `target` [P,R,3] (or [R,3]) at pixels `pix` [R].
    Returns losses [P]; call .sum().backward() to get d(loss_p)/d(pose_p) for every hypothesis in one pass."""
    P, R = poses.shape[0], pix.shape[0]
    rays_o, rays_d = rays_from_poses(poses, pix, intrinsics, W)
    out = model.render(rays_o.reshape(1, P * R, 3), rays_d.reshape(1, P * R, 3), **render_kwargs)
    rgb = out["image"].reshape(P, R, 3)
    return ((rgb - target) ** 2).mean(dim=(1, 2))


def look_at_pose(eye, target=(0.0, 0.0, 0.0), up=(0.0, 0.0, 1.0)):
    """cam2world [4,4] with the reference's camera convention (x right, y down, z forward)."""
    eye = torch.as_tensor(eye, dtype=torch.float32)
    fwd = torch.as_tensor(target, dtype=torch.float32) - eye
    fwd = fwd / fwd.norm()
    right = torch.linalg.cross(fwd, torch.as_tensor(up, dtype=torch.float32))
    right = right / right.norm()
    down = torch.linalg.cross(fwd, right)
    pose = torch.eye(4)
    pose[:3, 0], pose[:3, 1], pose[:3, 2], pose[:3, 3] = right, down, fwd, eye
    return pose


def perturb_poses(pose, n, sigma_t=0.02, sigma_r=0.02, generator=None):
    """n hypotheses = pose (+) N(0, diag(sigma_t m, sigma_r rad)); the first one is the pose itself."""
    dt = torch.randn(n, 3, generator=generator) * sigma_t
    w = torch.randn(n, 3, generator=generator) * sigma_r
    dt[0] = 0
    w[0] = 0
    th = w.norm(dim=-1, keepdim=True).clamp_min(1e-12)
    k = w / th
    K = torch.zeros(n, 3, 3)
    K[:, 0, 1], K[:, 0, 2], K[:, 1, 0], K[:, 1, 2], K[:, 2, 0], K[:, 2, 1] = -k[:, 2], k[:, 1], k[:, 2], -k[:, 0], -k[:, 1], k[:, 0]
    s, c = torch.sin(th)[..., None], torch.cos(th)[..., None]
    Rm = torch.eye(3) + s * K + (1 - c) * (K @ K)
    out = pose.repeat(n, 1, 1)
    out[:, :3, :3] = Rm @ pose[:3, :3]
    out[:, :3, 3] = pose[:3, 3] + dt
    return out


# NeRF training looks along +z, Blender along -z: the planner's points are rotated before the query (simulate.py:341)
NAV_ROT = ((0.0, 0.0, 1.0), (1.0, 0.0, 0.0), (0.0, 1.0, 0.0))


def make_density_fn(model, rot=NAV_ROT, single_pass=True):
    """The planner's closure (simulate.py:343): x [..., 3] -> sigma [...] = model.density(x @ rot)['sigma'].  With the
    model in eval() this is the fused field query (one kernel forward, one backward to x)."""
    def density_fn(x):
        r = torch.as_tensor(rot, dtype=torch.float32, device=x.device)
        p = x.reshape((-1, 3)) @ r
        if single_pass and p.is_cuda and not model.training and getattr(model, "fused_eval", False) and p.dtype == torch.float32:
            from .field import sigma_single_pass
            return sigma_single_pass(model, p).reshape(x.shape[:-1])      # value + Jacobian in one kernel (csrc/fieldvg.cu)
        return model.density(p)["sigma"].reshape(x.shape[:-1])
    return density_fn


@torch.no_grad()
def astar_occupancy(density_fn, side=100, kernel_size=5, thresh=0.3, lo=(-1.0, -1.0, -1.0), hi=(1.0, 1.0, 1.0),
                    device="cuda"):
    """The planner's A* occupancy grid (nav/quad_plot.py:64-83): density at a side^3 lattice over [lo, hi], max-pooled by
    kernel_size^3, thresholded -> bool [side//k]^3.  Three launches: lattice positions (torch.linspace's arithmetic, meshgrid
    'ij' order) -> the density closure (one fused field query of side^3 points, 10^6 at the defaults) -> max-pool + threshold
    (csrc/occgrid.cu)."""
    import ctypes

    from . import _lib
    dev = torch.device(device)
    L = _lib.lib()
    xyz = torch.empty(side, side, side, 3, device=dev)
    f3 = ctypes.c_float * 3
    g = side // kernel_size
    out = torch.empty(g, g, g, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _lib.check(L.ngp_lattice_positions(ctypes.cast(f3(*[float(v) for v in lo]), ctypes.c_void_p),
                                           ctypes.cast(f3(*[float(v) for v in hi]), ctypes.c_void_p), side, _lib.ptr(xyz),
                                           _lib.stream()), "lattice_positions")
        sigma = density_fn(xyz).float().contiguous()
        _lib.check(L.ngp_maxpool_threshold(_lib.ptr(sigma), side, kernel_size, float(thresh), _lib.ptr(out), _lib.stream()),
                   "maxpool_threshold")
    return out.bool()
