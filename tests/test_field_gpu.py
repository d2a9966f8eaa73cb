"""GPU parity: fused field query (density fwd + dL/dx bwd) vs the CPU oracle."""
import numpy as np
import pytest
import torch

from util import hashgrid_config, sigma_net_weights, rel_err, rel_err_cols

pytestmark = pytest.mark.gpu


def _run(ngp, cfg, W0, W1, x, gs, gg=None):
    B = x.shape[0]
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    xt, tt, ot, w0, w1, gst = dev(x), dev(cfg["table"]), dev(cfg["offsets"]), dev(W0), dev(W1), dev(gs)
    ggt = None if gg is None else dev(gg)
    sigma = torch.empty(B, device="cuda"); geo = torch.empty(B, 15, device="cuda")
    mask = torch.empty(B, device="cuda", dtype=torch.int64); gx = torch.empty(B, 3, device="cuda")
    L = ngp.lib()
    ngp.check(L.ngp_field_query_forward(xt.data_ptr(), cfg["bound"], tt.data_ptr(), ot.data_ptr(), cfg["S"], cfg["H"], 0, 0,
                                        w0.data_ptr(), w1.data_ptr(), sigma.data_ptr(), geo.data_ptr(), mask.data_ptr(),
                                        B, ngp.stream()), "fwd")
    ngp.check(L.ngp_field_query_backward(xt.data_ptr(), cfg["bound"], tt.data_ptr(), ot.data_ptr(), cfg["S"], cfg["H"], 0, 0,
                                         w0.data_ptr(), w1.data_ptr(), sigma.data_ptr(), mask.data_ptr(), gst.data_ptr(),
                                         ngp.ptr(ggt), gx.data_ptr(), B, ngp.stream()), "bwd")
    torch.cuda.synchronize()
    return sigma.cpu().numpy(), geo.cpu().numpy(), gx.cpu().numpy()


@pytest.mark.parametrize("B,with_geo", [(1, False), (4097, False), (30000, True)])
def test_field_query_vs_oracle(ngp, oracle, B, with_geo):
    cfg = hashgrid_config(bound=2.0, table_std=0.5)
    W0, W1 = sigma_net_weights()
    rng = np.random.default_rng(B)
    x = rng.uniform(-1, 1, (B, 3)).astype(np.float32)
    if B > 10:
        x[1] = [2.0, -2.0, 0.0]      # on the boundary: still in range
        x[2] = [2.5, 0.0, 0.0]       # outside: features zero -> sigma = exp(0) = 1, zero gradient
    gs = rng.standard_normal(B).astype(np.float32)
    gg = rng.standard_normal((B, 15)).astype(np.float32) if with_geo else None
    s = torch.empty(16, device="cuda")
    ngp.check(ngp.lib().ngp_grid_level_scales(s.data_ptr(), 16, cfg["S"], cfg["H"], ngp.stream()))
    scales = s.cpu().numpy()
    rs, rg, rgx = oracle.field_query(x, cfg["bound"], cfg["table"], cfg["offsets"], scales, W0, W1, gs, gg)
    sigma, geo, gx = _run(ngp, cfg, W0, W1, x, gs, gg)
    assert rel_err(sigma, rs) < 1e-5
    assert rel_err(geo, rg) < 1e-5
    assert rel_err(gx, rgx) < 1e-5
    if B > 10:
        # every geo_feat channel / gradient axis against its OWN scale (a small channel cannot hide behind a large one)
        assert rel_err_cols(geo, rg) < 1e-5 and rel_err_cols(gx, rgx) < 1e-5
    if B > 10:
        assert sigma[2] == 1.0 and np.all(gx[2] == 0.0)


def test_host_field_query_pipeline_matches_device_path():
    """HostFieldQuery (chunked H2D / kernels / D2H overlap) == NeRFNetwork.density + autograd on device tensors, incl. a
    ragged last chunk and repeated use of the slots."""
    import torch
    from nerf_navigation_b200.nerf.network import NeRFNetwork
    from nerf_navigation_b200.field import HostFieldQuery
    torch.manual_seed(0)
    m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=False).cuda().eval()
    m.encoder.embeddings.data.uniform_(-0.5, 0.5)
    B = 100003
    xh = (torch.rand(B, 3) * 2 - 1).pin_memory()
    sh = torch.empty(B).pin_memory(); gh = torch.empty(B, 3).pin_memory()
    q = HostFieldQuery(m, B, chunk=8192)
    for _ in range(2):
        q(xh, lambda s: s * (2.0 / B), sh, gh)
    xd = xh.cuda().requires_grad_(True)
    s = m.density(xd)["sigma"]
    (s ** 2).mean().backward()
    # the host query runs the single-pass value+grad kernel (mma.sync), density() the tcgen05 kernels: two 3xTF32 orders
    assert rel_err(sh.numpy(), s.detach().cpu().numpy()) < 1e-6
    assert rel_err(gh.numpy(), xd.grad.cpu().numpy()) < 1e-6
    # ramped chunk schedule (shorter first / last chunks) covers the batch exactly, in order
    for ramp in (False, True, 3):
        qr = HostFieldQuery(m, B, chunk=8192, ramp=ramp)
        sizes = qr._schedule(B)
        assert sum(sizes) == B and max(sizes) <= 8192 and min(sizes) > 0
        if ramp:
            assert sizes[0] < 8192 and sizes[-1] < 8192 and sizes[0] == sizes[-1]
    # affine upstream (a0 + b0 sigma) = one kernel per chunk; sigma_hook sees every chunk on the compute stream
    sh2 = torch.empty(B).pin_memory(); gh2 = torch.empty(B, 3).pin_memory()
    acc = torch.zeros((), device="cuda", dtype=torch.float64)
    qr(xh, (0.0, 2.0 / B), sh2, gh2, sigma_hook=lambda sig: acc.add_((sig.double() ** 2).sum()))
    assert rel_err(sh2.numpy(), sh.numpy()) < 1e-6     # single-pass kernel vs forward kernel (see above)
    assert rel_err(gh2.numpy(), gh.numpy()) < 1e-5
    assert abs(float(acc) / B - float((s.detach().double() ** 2).mean())) < 1e-7 * max(1.0, float(acc) / B)


@pytest.mark.parametrize("B,mode", [(1, "jac"), (191, "quad"), (4097, "coef"), (30000, "quad")])
def test_field_value_grad_vs_oracle_and_two_pass(ngp, oracle, B, mode):
    """ngp_field_query_value_grad (one table pass, corners cached in shared memory) == forward + backward kernels and
    == the CPU oracle, for the three upstream forms: Jacobian rows (u = 1), mean(sigma^2) (u = 2 sigma / B) and per-point
    collision-cost coefficients (u = a_i + b_i sigma_i)."""
    cfg = hashgrid_config(bound=2.0, table_std=0.5)
    W0, W1 = sigma_net_weights()
    rng = np.random.default_rng(B)
    x = rng.uniform(-1, 1, (B, 3)).astype(np.float32)
    if B > 10:
        x[1] = [2.0, -2.0, 0.0]
        x[2] = [2.5, 0.0, 0.0]
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    xt, tt, ot, w0, w1 = dev(x), dev(cfg["table"]), dev(cfg["offsets"]), dev(W0), dev(W1)
    ca = cb = None
    a0, b0 = 1.0, 0.0
    if mode == "quad":
        a0, b0 = 0.0, 2.0 / B
    elif mode == "coef":
        ca, cb = dev(rng.standard_normal(B).astype(np.float32)), dev(rng.uniform(0, 2, B).astype(np.float32))
    sigma = torch.empty(B, device="cuda"); geo = torch.empty(B, 15, device="cuda"); gx = torch.empty(B, 3, device="cuda")
    ngp.check(ngp.lib().ngp_field_query_value_grad(xt.data_ptr(), cfg["bound"], tt.data_ptr(), ot.data_ptr(), cfg["S"], cfg["H"],
                                                   0, 0, w0.data_ptr(), w1.data_ptr(), ngp.ptr(ca), ngp.ptr(cb), a0, b0,
                                                   sigma.data_ptr(), geo.data_ptr(), gx.data_ptr(), B, ngp.stream()), "vg")
    u = (ca if ca is not None else a0) + (cb if cb is not None else b0) * sigma
    s2, g2, gx2 = _run(ngp, cfg, W0, W1, x, u.cpu().numpy())
    assert np.array_equal(sigma.cpu().numpy(), s2) and np.array_equal(geo.cpu().numpy(), g2)   # same forward arithmetic
    assert rel_err(gx.cpu().numpy(), gx2) < 1e-6
    s16 = torch.empty(16, device="cuda")
    ngp.check(ngp.lib().ngp_grid_level_scales(s16.data_ptr(), 16, cfg["S"], cfg["H"], ngp.stream()))
    rs, rg, rgx = oracle.field_query(x, cfg["bound"], cfg["table"], cfg["offsets"], s16.cpu().numpy(), W0, W1, u.cpu().numpy())
    assert rel_err(sigma.cpu().numpy(), rs) < 1e-5 and rel_err(geo.cpu().numpy(), rg) < 1e-5
    assert rel_err(gx.cpu().numpy(), rgx) < 1e-5
    if B > 10:
        assert sigma[2] == 1.0 and (gx[2] == 0).all()


def test_density_on_pinned_host_points_matches_device_path():
    """model.density(x) with x in PINNED HOST memory (the drop-in call, copies pipelined inside field._field_query_host):
    sigma / geo_feat on the device and x.grad on the host equal the device-tensor path; ragged sizes around the chunk
    schedule; a pageable host tensor is refused (no CPU path)."""
    from nerf_navigation_b200.nerf.network import NeRFNetwork
    torch.manual_seed(0)
    m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=False).cuda().eval()
    m.encoder.embeddings.data.uniform_(-0.5, 0.5)
    for p in m.parameters():
        p.requires_grad_(False)
    for B in (1, 5000, (1 << 20) + 17, 3 * (1 << 20) + 5):
        xh = (torch.rand(B, 3) * 4.2 - 2.1).pin_memory().requires_grad_(True)
        out = m.density(xh)
        assert out["sigma"].is_cuda and out["sigma"].shape == (B,) and out["geo_feat"].shape == (B, 15)
        ((out["sigma"] ** 2).mean() + (out["geo_feat"] * 0.01).sum()).backward()
        xd = xh.detach().cuda().requires_grad_(True)
        ref = m.density(xd)
        ((ref["sigma"] ** 2).mean() + (ref["geo_feat"] * 0.01).sum()).backward()
        assert not xh.grad.is_cuda and xh.grad.is_pinned()
        assert torch.equal(out["sigma"], ref["sigma"]) and torch.equal(out["geo_feat"], ref["geo_feat"])
        assert torch.equal(xh.grad, xd.grad.cpu())
    with pytest.raises(RuntimeError):
        m.density(torch.rand(8, 3))


@pytest.mark.parametrize("impl", ["tc5", "mma"])
@pytest.mark.parametrize("B", [1, 4097, 30001])
def test_forward_jacobian_vs_oracle_and_two_kernels(ngp, oracle, B, impl):
    """ngp_field_query_forward_jacobian (what density(x) runs when x requires a gradient): sigma / geo_feat / the saved ReLU
    bits equal the plain forward's, jac_i * grad_sigma_i (ngp_field_query_backward_jacobian) equals the second-table-pass
    backward and the CPU oracle, and the saved bits drive ngp_field_query_backward when a gradient arrives through geo_feat."""
    L = ngp.lib()
    prev = L.ngp_set_field_impl(1 if impl == "tc5" else 0)
    try:
        cfg = hashgrid_config(bound=2.0, table_std=0.5)
        W0, W1 = sigma_net_weights()
        rng = np.random.default_rng(B)
        x = rng.uniform(-1, 1, (B, 3)).astype(np.float32)
        if B > 10:
            x[1] = [2.0, -2.0, 0.0]
            x[2] = [2.5, 0.0, 0.0]
        gs = rng.standard_normal(B).astype(np.float32)
        gg = rng.standard_normal((B, 15)).astype(np.float32)
        dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
        xt, tt, ot, w0, w1, gst, ggt = dev(x), dev(cfg["table"]), dev(cfg["offsets"]), dev(W0), dev(W1), dev(gs), dev(gg)
        sigma = torch.empty(B, device="cuda"); geo = torch.empty(B, 15, device="cuda"); jac = torch.empty(B, 3, device="cuda")
        mask = torch.empty(B, device="cuda", dtype=torch.int64); gx = torch.empty(B, 3, device="cuda")
        ngp.check(L.ngp_field_query_forward_jacobian(xt.data_ptr(), cfg["bound"], tt.data_ptr(), ot.data_ptr(), cfg["S"], cfg["H"],
                                                     0, 0, w0.data_ptr(), w1.data_ptr(), sigma.data_ptr(), geo.data_ptr(),
                                                     mask.data_ptr(), jac.data_ptr(), B, ngp.stream()), "fwd_jac")
        ngp.check(L.ngp_field_query_backward_jacobian(jac.data_ptr(), gst.data_ptr(), gx.data_ptr(), B, ngp.stream()), "bwd_jac")
        s2, g2, gx2 = _run(ngp, cfg, W0, W1, x, gs)
        assert np.array_equal(sigma.cpu().numpy(), s2) and np.array_equal(geo.cpu().numpy(), g2)      # same forward arithmetic
        assert rel_err(gx.cpu().numpy(), gx2) < 1e-6
        s16 = torch.empty(16, device="cuda")
        ngp.check(L.ngp_grid_level_scales(s16.data_ptr(), 16, cfg["S"], cfg["H"], ngp.stream()))
        scales = s16.cpu().numpy()
        rs, rg, rgx = oracle.field_query(x, cfg["bound"], cfg["table"], cfg["offsets"], scales, W0, W1, gs)
        assert rel_err(sigma.cpu().numpy(), rs) < 1e-5 and rel_err(geo.cpu().numpy(), rg) < 1e-5
        assert rel_err(gx.cpu().numpy(), rgx) < 1e-5
        if B > 10:
            assert rel_err_cols(gx.cpu().numpy(), rgx) < 1e-5
            assert sigma[2] == 1.0 and (jac[2] == 0).all()
        # gradient through geo_feat as well: the second table pass from the bits the Jacobian forward saved
        gx3 = torch.empty(B, 3, device="cuda")
        ngp.check(L.ngp_field_query_backward(xt.data_ptr(), cfg["bound"], tt.data_ptr(), ot.data_ptr(), cfg["S"], cfg["H"], 0, 0,
                                             w0.data_ptr(), w1.data_ptr(), sigma.data_ptr(), mask.data_ptr(), gst.data_ptr(),
                                             ggt.data_ptr(), gx3.data_ptr(), B, ngp.stream()), "bwd")
        _, _, rgx3 = oracle.field_query(x, cfg["bound"], cfg["table"], cfg["offsets"], scales, W0, W1, gs, gg)
        assert rel_err(gx3.cpu().numpy(), rgx3) < 1e-5
    finally:
        L.ngp_set_field_impl(prev)


def test_density_autograd_takes_the_jacobian_forward_and_handles_geo_gradients():
    """model.density(x) with x.requires_grad: a loss on sigma alone (Jacobian rows, no second table pass) and a loss that also
    uses geo_feat (second pass from the saved bits) both match the unfused GridEncoder -> Linear -> trunc_exp autograd path;
    an unused output gives a zero gradient."""
    from nerf_navigation_b200.nerf.network import NeRFNetwork
    torch.manual_seed(1)
    m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=False).cuda().eval()
    m.encoder.embeddings.data.uniform_(-0.5, 0.5)
    for p in m.parameters():
        p.requires_grad_(False)
    x = torch.rand(20011, 3, device="cuda") * 2 - 1
    w = torch.randn(15, device="cuda")
    for use_geo in (False, True):
        grads = []
        for fused in (True, False):
            m.fused_eval = fused
            xi = x.clone().requires_grad_(True)
            out = m.density(xi)
            loss = (out["sigma"] ** 2).sum() + ((out["geo_feat"] * w).sum() if use_geo else 0.0)
            loss.backward()
            grads.append(xi.grad.clone())
        assert rel_err(grads[0].cpu().numpy(), grads[1].cpu().numpy()) < 1e-5, use_geo
    m.fused_eval = True
    xi = x.clone().requires_grad_(True)
    out = m.density(xi)
    (out["geo_feat"] * w).sum().backward()        # sigma unused: grad_sigma arrives as None
    xj = x.clone().requires_grad_(True)
    m.fused_eval = False
    (m.density(xj)["geo_feat"] * w).sum().backward()
    assert rel_err(xi.grad.cpu().numpy(), xj.grad.cpu().numpy()) < 1e-5
