"""GPU: the training configuration (BASELINE config 5) — fused Adam against torch.optim.Adam, gradient averaging over
emulated ranks against the single-process gradient of the concatenated batch, and a few optimiser steps end to end."""
import copy

import numpy as np
import pytest
import torch

from util import stonehenge_grid, orbit_rays, rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n", [1, 7, 4096, 100003])
def test_fused_adam_matches_torch_adam(n):
    from nerf_navigation_b200.optim import FusedAdam
    torch.manual_seed(n)
    p0 = torch.randn(n, device="cuda")
    pa = torch.nn.Parameter(p0.clone()); pb = torch.nn.Parameter(p0.clone())
    oa = FusedAdam([pa], lr=1e-2, betas=(0.9, 0.99), eps=1e-15, ema_decay=0.95)
    ob = torch.optim.Adam([pb], lr=1e-2, betas=(0.9, 0.99), eps=1e-15)
    shadow = p0.clone()
    for step in range(6):
        g = torch.randn(n, device="cuda") * (10.0 ** (step - 3))
        pa.grad = g.clone() * 4.0          # as if 4 ranks' gradients had been summed
        oa.grad_mult = 0.25
        pb.grad = g.clone()
        oa.step(); ob.step()
        decay = min(0.95, (1 + (step + 1)) / (10 + (step + 1)))     # torch_ema, use_num_updates=True (the reference's default)
        shadow = shadow - (1 - decay) * (shadow - pb.detach())      # torch_ema update (nerf/utils.py:814)
    assert rel_err(pa.detach().cpu().numpy(), pb.detach().cpu().numpy()) < 1e-6
    sa, sb = oa.state[pa], ob.state[pb]
    assert rel_err(sa["exp_avg"].cpu().numpy(), sb["exp_avg"].cpu().numpy()) < 1e-6
    assert rel_err(sa["exp_avg_sq"].cpu().numpy(), sb["exp_avg_sq"].cpu().numpy()) < 1e-6
    assert rel_err(oa.ema_shadow(pa).cpu().numpy(), shadow.cpu().numpy()) < 1e-6
    assert sa["step"] == 6


def test_fused_adam_under_grad_scaler():
    """The reference's fp16 recipe (nerf/utils.py:784-791): scaler.scale(loss).backward(); scaler.step(opt); scaler.update().
    FusedAdam takes GradScaler's device-side scale / found_inf like torch's fused optimisers: finite gradients are unscaled
    inside the kernel (== torch.optim.Adam under the same scaler), an inf gradient skips the step — parameters, moments,
    step count and EMA untouched — and the scaler backs off."""
    from nerf_navigation_b200.optim import FusedAdam
    torch.manual_seed(3)
    p0 = torch.randn(5000, device="cuda")
    pa = torch.nn.Parameter(p0.clone()); pb = torch.nn.Parameter(p0.clone())
    oa = FusedAdam([pa], lr=1e-2, betas=(0.9, 0.99), eps=1e-15, ema_decay=0.95)
    ob = torch.optim.Adam([pb], lr=1e-2, betas=(0.9, 0.99), eps=1e-15)
    sa = torch.amp.GradScaler("cuda", init_scale=1024.0); sb = torch.amp.GradScaler("cuda", init_scale=1024.0)
    w = torch.randn(5000, device="cuda")
    for step in range(5):
        for p, o, sc in ((pa, oa, sa), (pb, ob, sb)):
            o.zero_grad()
            loss = ((p * w) ** 2).sum() * (float("inf") if step == 2 else 1.0)
            sc.scale(loss).backward()
            sc.step(o)
            sc.update()
        if step == 2:   # the overflow step changed nothing and halved the scale
            assert oa.state[pa]["step"] == 2 and sa.get_scale() == 512.0 and sb.get_scale() == 512.0
    assert rel_err(pa.detach().cpu().numpy(), pb.detach().cpu().numpy()) < 1e-6
    assert oa.state[pa]["step"] == 4 and int(ob.state[pb]["step"]) == 4 and oa.ema_updates == 4
    assert torch.isfinite(pa).all() and torch.isfinite(oa.ema_shadow(pa)).all()


@pytest.fixture(scope="module")
def model():
    from nerf_navigation_b200.nerf.network import NeRFNetwork
    from nerf_navigation_b200 import raymarching as rm
    torch.manual_seed(0)
    m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10).cuda()
    m.encoder.embeddings.data.uniform_(-0.5, 0.5)
    m.density_grid.copy_(torch.from_numpy(stonehenge_grid()))
    m.density_bitfield = rm.packbits(m.density_grid, 10.0, m.density_bitfield)
    return m


def _grads(model, ro, rd, target):
    model.train()
    model.zero_grad(set_to_none=True)
    out = model.render(ro.unsqueeze(0), rd.unsqueeze(0), bg_color=1.0, perturb=False, force_all_rays=True, dt_gamma=0,
                       max_steps=1024)
    ((out["image"].reshape(-1, 3) - target) ** 2).mean().backward()
    return [p.grad.clone() for p in model.parameters() if p.requires_grad]


def test_rank_mean_gradient_equals_concatenated_batch(model):
    """the reduced (mean over ranks) table / MLP gradients equal a single-process computation on the
    concatenated batch (equal shard sizes, loss = mean over rays).  fp32, different summation order: 1e-5."""
    o, d = orbit_rays(2048, 4)
    ro, rd = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    target = torch.rand(2048, 3, device="cuda")
    full = _grads(model, ro, rd, target)
    parts = [_grads(model, ro[a:a + 1024], rd[a:a + 1024], target[a:a + 1024]) for a in (0, 1024)]
    for gf, g0, g1 in zip(full, parts[0], parts[1]):
        assert rel_err(((g0 + g1) * 0.5).cpu().numpy(), gf.cpu().numpy()) < 1e-5
    assert (full[0] != 0).any()


def test_train_steps_reduce_loss(model):
    from nerf_navigation_b200.training import make_optimizer, train_step
    m = copy.deepcopy(model)
    opt = make_optimizer(m, lr=1e-2, ema_decay=0.95)
    assert len(opt.param_groups) == 3      # encoder, sigma_net, color_net (the SH encoder has no parameters)
    o, d = orbit_rays(4096, 6)
    ro, rd = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    target = torch.full((4096, 3), 0.3, device="cuda")
    before = m.encoder.embeddings.detach().clone()
    losses = [float(train_step(m, opt, ro, rd, target, world=1, perturb=False, force_all_rays=True, dt_gamma=0,
                               max_steps=1024)) for _ in range(8)]
    assert losses[-1] < losses[0] and np.isfinite(losses).all()
    assert (m.encoder.embeddings.detach() != before).any()
    assert opt.state[m.encoder.embeddings]["step"] == 8


@pytest.mark.parametrize("M,n_out,n_in", [(1, 3, 64), (7, 16, 64), (1000, 64, 31), (12345, 64, 32), (327680, 64, 64)])
def test_linear_wgrad_kernel(ngp, M, n_out, n_in):
    """csrc/wgrad.cu: dW = dY^T X at fp32 accuracy (3xTF32) against an fp64 product, and bit-reproducible."""
    g = torch.Generator(device="cuda").manual_seed(M)
    dY = torch.randn(M, n_out, device="cuda", generator=g)
    X = torch.randn(M, n_in, device="cuda", generator=g)
    L = ngp.lib()
    nb = L.ngp_linear_wgrad_workspace_bytes(M)
    ws = torch.empty(nb, dtype=torch.uint8, device="cuda")
    outs = []
    for _ in range(2):
        dW = torch.full((n_out, n_in), float("nan"), device="cuda")
        ngp.check(L.ngp_linear_wgrad(dY.data_ptr(), X.data_ptr(), dW.data_ptr(), M, n_out, n_in, ws.data_ptr(), nb, ngp.stream()))
        outs.append(dW)
    ref = (dY.double().t() @ X.double()).cpu().numpy()
    sgemm = (dY.t() @ X).double().cpu().numpy()
    assert torch.equal(outs[0], outs[1])
    err, err_sgemm = rel_err(outs[0].cpu().numpy(), ref), rel_err(sgemm, ref)
    print(f"wgrad M={M} {n_out}x{n_in}: rel err {err:.2e} (torch sgemm {err_sgemm:.2e})")
    assert err < 1e-5 and err < 4 * err_sgemm + 2e-6, (err, err_sgemm)   # fp32 bar of north_star; on par with the library GEMM


@pytest.mark.parametrize("M,n_out,n_in", [(1, 3, 64), (17, 16, 64), (1000, 64, 31), (12345, 64, 32), (327683, 64, 64), (4096, 3, 64), (999, 15, 7)])
def test_linear_small_kernel(ngp, M, n_out, n_in):
    """csrc/linear.cu: y = x W^T and dx = dy W at fp32 accuracy (3xTF32) against fp64 products, ragged M / odd widths."""
    g = torch.Generator(device="cuda").manual_seed(M + n_out)
    W = torch.randn(n_out, n_in, device="cuda", generator=g) / n_in ** 0.5
    X = torch.randn(M, n_in, device="cuda", generator=g)
    dY = torch.randn(M, n_out, device="cuda", generator=g)
    L = ngp.lib()
    Y = torch.full((M, n_out), float("nan"), device="cuda"); dX = torch.full((M, n_in), float("nan"), device="cuda")
    ngp.check(L.ngp_linear_small(X.data_ptr(), W.data_ptr(), Y.data_ptr(), M, n_out, n_in, 0, ngp.stream()))
    ngp.check(L.ngp_linear_small(dY.data_ptr(), W.data_ptr(), dX.data_ptr(), M, n_out, n_in, 1, ngp.stream()))
    ref_y = (X.double() @ W.double().t()).cpu().numpy(); ref_dx = (dY.double() @ W.double()).cpu().numpy()
    ey, ex = rel_err(Y.cpu().numpy(), ref_y), rel_err(dX.cpu().numpy(), ref_dx)
    sy = rel_err((X @ W.t()).double().cpu().numpy(), ref_y)
    print(f"linear M={M} {n_out}x{n_in}: rel err y {ey:.2e} dx {ex:.2e} (torch sgemm {sy:.2e})")
    assert ey < 2e-6 and ex < 2e-6


def test_linear_nobias_autograd_matches_nn_linear():
    from nerf_navigation_b200.field import linear_nobias
    torch.manual_seed(0)
    layer = torch.nn.Linear(31, 64, bias=False).cuda()
    x = torch.randn(5000, 31, device="cuda", requires_grad=True)
    w = torch.randn(5000, 64, device="cuda")
    (linear_nobias(layer, x) * w).sum().backward()
    ga, gwa = x.grad.clone(), layer.weight.grad.clone()
    x.grad = None; layer.weight.grad = None
    (layer(x) * w).sum().backward()
    assert rel_err(ga.cpu().numpy(), x.grad.cpu().numpy()) < 1e-6
    # two fp32 summation orders over 5000 samples (ours is checked against fp64 in test_linear_wgrad_kernel)
    assert rel_err(gwa.cpu().numpy(), layer.weight.grad.cpu().numpy()) < 1e-5
