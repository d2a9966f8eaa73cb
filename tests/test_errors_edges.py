"""Error conventions and edge cases of the drop-in boundary: argument validation returns a non-zero status +
ngp_last_error() (the reference raises RuntimeError through TORCH_CHECK / std::runtime_error, gridencoder.cu:15-18, 355,
372; ffmlp.cu:658) — checked on CPU where the check fires before any launch; empty and ragged inputs on the GPU."""
import ctypes as C

import numpy as np
import pytest
import torch


def test_argument_validation_without_a_gpu(ngp):
    """Null pointers / unsupported shapes are rejected before any CUDA call, with a message naming the entry point."""
    L = ngp.lib()
    prev = L.ngp_set_device_check(0)      # the host scratch pointers below are deliberate: argument checks come first here
    rc = L.ngp_near_far_from_aabb(None, None, None, 4, 0.2, None, None, None)
    assert rc == 1 and b"ngp_near_far_from_aabb" in L.ngp_last_error()
    rc = L.ngp_field_query_forward(None, 2.0, None, None, 0.5, 16, 0, 0, None, None, None, None, None, 8, None)
    assert rc != 0 and b"field_query" in L.ngp_last_error()
    rc = L.ngp_render_rays_infer(None, None, 8, None, 0.2, None, 2.0, 0.0, 1024, 2, 128, None, None, 0.5, 16, 0, 0, None, None,
                                 None, None, None, 1.0, None, 0, None, None, None, None, None, None)
    assert rc == 1 and b"ngp_render_rays_infer" in L.ngp_last_error()
    buf = (C.c_float * 64)()
    p = C.addressof(buf)
    rc = L.ngp_linear_wgrad(p, p, p, 16, 65, 64, p, 1 << 20, None)      # n_out > 64
    assert rc == 3 and b"n_out" in L.ngp_last_error()
    rc = L.ngp_adam_step(p, p, p, p, None, 16, 1e-2, 0.9, 0.99, 1e-15, 0, 1.0, 0.0, None, None, None)   # step counts from 1
    assert rc == 1 and b"step" in L.ngp_last_error()
    assert L.ngp_grid_backward_table_workspace_bytes(1000, 3, 16) > 4 * 1000 * 16 * 8 * 4
    assert L.ngp_march_rays_train_workspace_bytes(1000) >= 2 * 1000 * 4
    assert L.ngp_render_rays_infer_workspace_bytes() >= 16 and L.ngp_linear_wgrad_workspace_bytes(1 << 20) > 0
    assert L.ngp_occ_grid_partial_workspace_bytes(128) > 4 * 128 ** 3
    rc = L.ngp_occ_grid_mark_untrained(None, 3, 1.0, 1.0, 1.0, 1.0, 2, 128, 2.0, None, None)
    assert rc == 1 and b"mark_untrained" in L.ngp_last_error()
    # round-2 entry points
    rargs = (None, None, 8, None, 0.2, None, 2.0, 0.0, 1024, 2, 128, None, None, 0.5, 16, 0, 0, None, None, None, None, None, 1.0,
             None, 0, None, None, None, None)
    rc = L.ngp_render_rays_infer_segments(*rargs, 32, None, 0, None)
    assert rc == 1 and b"ngp_render_rays_infer_segments" in L.ngp_last_error()
    assert L.ngp_render_rays_infer_segments_workspace_bytes(1000, 32) == 512 + 4000 + 4000 + 1000 * 32 * 8
    assert L.ngp_render_rays_infer_segments_workspace_bytes(1000, 0) < 0
    rc = L.ngp_render_rays_infer_planned(*rargs, None, None, None, 0, None, None)
    assert rc == 1 and b"ngp_render_rays_infer_planned" in L.ngp_last_error()
    rc = L.ngp_render_plan(None, None, 8, None, 0.2, None, 2.0, 0.0, 1024, 2, 128, None, None, None, 0, None)
    assert rc == 1 and b"ngp_render_plan" in L.ngp_last_error()
    assert L.ngp_render_plan_workspace_bytes(1000) >= 256
    rc = L.ngp_field_query_forward_jacobian(None, 2.0, None, None, 0.5, 16, 0, 0, None, None, None, None, None, None, 8, None)
    assert rc == 1 and b"forward_jacobian" in L.ngp_last_error()
    rc = L.ngp_field_query_ff_forward(None, 2.0, None, None, 0.5, 16, 0, 0, None, 2, 0, None, 8, None)
    assert rc == 1 and b"ngp_field_query_ff_forward" in L.ngp_last_error()
    rc = L.ngp_grid_encode_backward_table_fixed(None, 1, 0, None, None, 100, None, 8, 3, 2, 16, 0.5, 16, 0, 0, None, 0, None)
    assert rc == 1 and b"backward_table_fixed" in L.ngp_last_error()
    assert L.ngp_grid_backward_table_fixed_workspace_bytes(1000, 2) == 256 + 1000 * 2 * 8
    rc = L.ngp_linear_wgrad_static(p, p, p, 65537, 16, 32, None)        # beyond the static scratch
    assert rc == 3 and b"static scratch limit" in L.ngp_last_error()
    L.ngp_set_device_check(prev)


@pytest.mark.gpu
def test_unsupported_shapes_raise_runtime_error():
    from nerf_navigation_b200.gridencoder import GridEncoder
    enc = GridEncoder(input_dim=3, num_levels=4, level_dim=3, log2_hashmap_size=10, desired_resolution=64).cuda()
    with pytest.raises(RuntimeError, match="C must be"):       # gridencoder.cu:355: C in {1, 2, 4, 8}
        enc(torch.rand(8, 3, device="cuda"))
    enc4 = GridEncoder(input_dim=4, num_levels=4, level_dim=2, log2_hashmap_size=10, desired_resolution=64).cuda()
    with pytest.raises(RuntimeError, match="D must be"):       # this library covers D in {2, 3} (the reference: 2..5)
        enc4(torch.rand(8, 4, device="cuda"))
    from nerf_navigation_b200 import raymarching as rm
    # CPU tensors: the raymarching wrappers move them to the GPU like the reference's (raymarching.py:29-30)
    nears, fars = rm.near_far_from_aabb(torch.zeros(4, 3), torch.ones(4, 3), torch.tensor([-1.0] * 3 + [1.0] * 3).cuda(), 0.2)
    assert nears.is_cuda and fars.is_cuda
    # ... the encoder does not (grid.py has no such move): loud failure, no CPU fallback
    enc2 = GridEncoder(input_dim=3, num_levels=4, level_dim=2, log2_hashmap_size=10, desired_resolution=64).cuda()
    with pytest.raises(RuntimeError):
        enc2(torch.rand(8, 3))


@pytest.mark.gpu
def test_empty_inputs():
    """B = 0 / N = 0 everywhere: empty outputs, no launch failure."""
    from nerf_navigation_b200.nerf.network import NeRFNetwork
    from nerf_navigation_b200 import raymarching as rm
    from nerf_navigation_b200.field import render_rays_infer, field_value_grad
    m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True).cuda().eval()
    e = torch.empty(0, 3, device="cuda")
    d = m.density(e)
    assert d["sigma"].shape == (0,) and d["geo_feat"].shape == (0, 15)
    s, c = m(e, e)
    assert s.shape == (0,) and c.shape == (0, 3)
    assert m.encoder(e, bound=2).shape == (0, 32)
    sg, jac = field_value_grad(m, e)
    assert sg.shape == (0,) and jac.shape == (0, 3)
    nears, fars = rm.near_far_from_aabb(e, e, m.aabb_infer, 0.2)
    assert nears.shape == (0,)
    img, dep, ws = render_rays_infer(m, e, e, m.aabb_infer, 1.0)
    assert img.shape == (0, 3) and dep.shape == (0,)
    xyzs, dirs, deltas, rays = rm.march_rays_train(e, e, 2.0, m.density_bitfield, 2, 128, nears, fars, None, -1, False, 128, True)
    assert rays.shape == (0, 3)
    ws, depth, image = rm.composite_rays_train(torch.empty(0, device="cuda"), torch.empty(0, 3, device="cuda"),
                                               torch.empty(0, 2, device="cuda"), torch.empty(0, 3, dtype=torch.int32, device="cuda"))
    assert image.shape == (0, 3)


@pytest.mark.gpu
def test_empty_scene_renders_background():
    """An all-empty occupancy grid: no samples anywhere, every pixel is the background, through both render paths."""
    from nerf_navigation_b200.nerf.network import NeRFNetwork
    from util import orbit_rays
    m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True).cuda().eval()
    o, d = orbit_rays(1000, 0)
    ro, rd = torch.from_numpy(o).cuda().view(1, -1, 3), torch.from_numpy(d).cuda().view(1, -1, 3)
    with torch.no_grad():
        a = m.render(ro, rd, bg_color=0.5, perturb=False)
        m.fused_render = False
        b = m.render(ro, rd, bg_color=0.5, perturb=False)
        m.fused_render = True
    assert (a["image"] == 0.5).all() and (b["image"] == 0.5).all()
    assert torch.equal(a["depth"], b["depth"])
