"""Pose -> rays -> photometric loss closure for batches of pose hypotheses.

The reference builds ALL H*W rays of one pose and then indexes the ~1k sampled pixels (nerf/utils.py:52-116 `get_rays`,
nav/estimator_helpers.py:293-327 `measurement_fn`); here only the sampled pixels' rays are generated, for a batch of
hypotheses at once, with the same conventions: pixel centres at +0.5, direction ((i-cx)/fx, (j-cy)/fy, 1) normalised,
rays_d = dir @ R^T, rays_o = t (cam2world pose [R|t]); loss = mse(render(rays).image, target) per hypothesis.
Everything is differentiable w.r.t. the poses through the march_rays_train pose-gradient extension (cuda_ray=True) or
the torch sampler (cuda_ray=False).
"""
import math

import torch


class _pose_rays(torch.autograd.Function):
    """rays_from_poses on the GPU: one kernel forward, one per-hypothesis reduction backward (csrc/rays.cu)."""

    @staticmethod
    def forward(ctx, poses, pix, intrinsics, W):
        from . import _lib
        P, R = poses.shape[0], pix.shape[0]
        poses = poses.contiguous().float()
        pix = pix.contiguous().long()
        rays_o = torch.empty(P, R, 3, device=poses.device)
        rays_d = torch.empty(P, R, 3, device=poses.device)
        fx, fy, cx, cy = (float(v) for v in intrinsics)
        with torch.cuda.device(poses.device):
            _lib.check(_lib.lib().ngp_pose_rays_forward(_lib.ptr(poses), _lib.ptr(pix), P, R, int(W), fx, fy, cx, cy,
                                                        _lib.ptr(rays_o), _lib.ptr(rays_d), _lib.stream()), "pose_rays_forward")
        ctx.save_for_backward(pix)
        ctx.cfg = (P, R, int(W), fx, fy, cx, cy)
        return rays_o, rays_d

    @staticmethod
    def backward(ctx, grad_o, grad_d):
        from . import _lib
        (pix,) = ctx.saved_tensors
        P, R, W, fx, fy, cx, cy = ctx.cfg
        go = None if grad_o is None else grad_o.contiguous().float()
        gd = None if grad_d is None else grad_d.contiguous().float()
        grad_poses = torch.empty(P, 4, 4, device=pix.device)
        with torch.cuda.device(pix.device):
            _lib.check(_lib.lib().ngp_pose_rays_backward(_lib.ptr(go), _lib.ptr(gd), _lib.ptr(pix), P, R, W, fx, fy, cx, cy,
                                                         _lib.ptr(grad_poses), _lib.stream()), "pose_rays_backward")
        return grad_poses, None, None, None


def rays_from_poses(poses, pix, intrinsics, W):
    """poses [P,4,4] cam2world, pix [R] int64 flat pixel indices (row-major, j*W+i) shared by all hypotheses
    -> rays_o, rays_d [P,R,3].  CUDA tensors take the fused kernels (forward + per-hypothesis gradient reduction); the
    torch formulation below is the CPU / reference-shaped form (and the one the golden vectors pin)."""
    if poses.is_cuda and poses.dtype == torch.float32:
        return _pose_rays.apply(poses, pix, tuple(intrinsics), W)
    fx, fy, cx, cy = intrinsics
    i = (pix % W).to(poses.dtype) + 0.5
    j = torch.div(pix, W, rounding_mode="floor").to(poses.dtype) + 0.5
    d = torch.stack(((i - cx) / fx, (j - cy) / fy, torch.ones_like(i)), dim=-1)
    d = d / torch.norm(d, dim=-1, keepdim=True)
    rays_d = d.unsqueeze(0) @ poses[:, :3, :3].transpose(-1, -2)
    rays_o = poses[:, None, :3, 3].expand_as(rays_d)
    return rays_o, rays_d


def rotvec_to_matrix(v):
    """Rodrigues' formula as nav/math_utils.py:159-174 writes it: axis = v / (1e-10 + |v|) (so v = 0 gives the identity and
    a zero gradient), R = I + sin(a) K + (1 - cos(a)) K K.  v [..., 3] -> [..., 3, 3]."""
    angle = torch.norm(v, dim=-1, keepdim=True)
    k = v / (1e-10 + angle)
    zero = torch.zeros_like(k[..., 0])
    K = torch.stack([torch.stack([zero, -k[..., 2], k[..., 1]], -1),
                     torch.stack([k[..., 2], zero, -k[..., 0]], -1),
                     torch.stack([-k[..., 1], k[..., 0], zero], -1)], -2)
    a = angle[..., None]
    return torch.eye(3, dtype=v.dtype, device=v.device) + torch.sin(a) * K + (1 - torch.cos(a)) * (K @ K)


class _state_to_pose(torch.autograd.Function):
    """state_to_pose on the GPU: one kernel forward, one backward (csrc/rays.cu), instead of ~65 eager ops per step — at a few
    dozen hypotheses per GPU the estimator step is bound by the host's launch rate, not by the kernels."""

    @staticmethod
    def forward(ctx, states):
        from . import _lib
        states = states.contiguous()
        P = states.shape[0]
        poses = torch.empty(P, 4, 4, device=states.device)
        c, s = _half_pi_cos_sin()
        with torch.cuda.device(states.device):
            _lib.check(_lib.lib().ngp_state_to_pose_forward(_lib.ptr(states), P, c, s, _lib.ptr(poses), _lib.stream()),
                       "state_to_pose_forward")
        ctx.save_for_backward(states)
        return poses

    @staticmethod
    def backward(ctx, grad_poses):
        from . import _lib
        (states,) = ctx.saved_tensors
        P = states.shape[0]
        grad_states = torch.empty(P, 12, device=states.device)
        c, s = _half_pi_cos_sin()
        with torch.cuda.device(states.device):
            _lib.check(_lib.lib().ngp_state_to_pose_backward(_lib.ptr(states), _lib.ptr(grad_poses.contiguous().float()), P, c, s,
                                                             _lib.ptr(grad_states), _lib.stream()), "state_to_pose_backward")
        return grad_states


def _half_pi_cos_sin():
    half_pi = torch.tensor(math.pi / 2, dtype=torch.float32)           # the reference evaluates cos / sin of a float32 pi / 2
    return float(torch.cos(half_pi)), float(torch.sin(half_pi))


def state_to_pose(states):
    """The estimator's 12-state (position 3, velocity 3, rotation vector 3, angular velocity 3) -> the cam2world pose the
    renderer sees (nav/estimator_helpers.py:293-327 measurement_fn): R = vec_to_rot_matrix(state[6:9]); the camera looks
    along the body y axis, hence rot_x(pi/2) @ R; then Blender -> NGP axes (nav/math_utils.py:26-38): flip_yz @ rot @ neg_yz
    and flip_yz @ position.  states [P,12] -> poses [P,4,4], differentiable (dL/d(12-state) by autograd: velocity and
    angular velocity take no gradient from the measurement)."""
    if states.is_cuda and states.dtype == torch.float32 and states.dim() == 2 and states.shape[1] == 12:
        return _state_to_pose.apply(states)      # the torch formulation below is the host / reference-shaped form (golden-pinned)
    dt, dev = states.dtype, states.device
    c, s = _half_pi_cos_sin()
    rot_x = torch.tensor([[1.0, 0.0, 0.0], [0.0, c, -s], [0.0, s, c]], dtype=dt, device=dev)
    flip_yz = torch.tensor([[0.0, 1.0, 0.0], [0.0, 0.0, 1.0], [1.0, 0.0, 0.0]], dtype=dt, device=dev)
    neg_yz = torch.tensor([[1.0, 0.0, 0.0], [0.0, -1.0, 0.0], [0.0, 0.0, -1.0]], dtype=dt, device=dev)
    R = rotvec_to_matrix(states[:, 6:9])
    rot = flip_yz @ (rot_x @ R) @ neg_yz
    trans = states[:, :3] @ flip_yz.t()
    top = torch.cat([rot, trans[..., None]], -1)
    bottom = torch.tensor([0.0, 0.0, 0.0, 1.0], dtype=dt, device=dev).expand(states.shape[0], 1, 4)
    return torch.cat([top, bottom], -2)


def pose_to_state(pose, velocity=None, angular_velocity=None):
    """Inverse of state_to_pose for one cam2world pose [4,4] (rotation angle in (0, pi)): the 12-state whose measurement
    sees that pose — used to seed hypothesis batches from a camera pose."""
    dt = pose.dtype
    half_pi = torch.tensor(math.pi / 2, dtype=torch.float32)
    c, s = float(torch.cos(half_pi)), float(torch.sin(half_pi))
    rot_x = torch.tensor([[1.0, 0.0, 0.0], [0.0, c, -s], [0.0, s, c]], dtype=dt)
    flip_yz = torch.tensor([[0.0, 1.0, 0.0], [0.0, 0.0, 1.0], [1.0, 0.0, 0.0]], dtype=dt)
    neg_yz = torch.tensor([[1.0, 0.0, 0.0], [0.0, -1.0, 0.0], [0.0, 0.0, -1.0]], dtype=dt)
    pose = pose.detach().cpu()
    R = rot_x.t() @ flip_yz.t() @ pose[:3, :3] @ neg_yz          # neg_yz is its own inverse
    angle = torch.acos(((torch.trace(R) - 1) / 2).clamp(-1, 1))
    axis = torch.stack([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]]) / (2 * torch.sin(angle)).clamp_min(1e-12)
    z3 = torch.zeros(3, dtype=dt)
    return torch.cat([flip_yz.t() @ pose[:3, 3], z3 if velocity is None else velocity, angle * axis,
                      z3 if angular_velocity is None else angular_velocity])


def measurement_loss(model, states, pix, target, intrinsics, W, start_states=None, sig_inv=None, **render_kwargs):
    """measurement_fn for a batch of state hypotheses [P,12]: mse(render(rays(state_to_pose(state))), target) per hypothesis
    (+ the Mahalanobis process term delta^T sig_inv delta when start_states / sig_inv are given).  Returns losses [P];
    .sum().backward() gives dL_p / d state_p for every hypothesis in one pass."""
    loss = photometric_loss(model, state_to_pose(states), pix, target, intrinsics, W, **render_kwargs)
    if start_states is not None and sig_inv is not None:
        delta = states - start_states
        loss = loss + torch.einsum("pi,ij,pj->p", delta, sig_inv, delta)
    return loss


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
