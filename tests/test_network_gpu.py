"""GPU: NeRFNetwork / NeRFRenderer host mirror — fused vs unfused density, run() and run_cuda() against an oracle pipeline."""
import numpy as np
import pytest
import torch

from util import stonehenge_grid, orbit_rays, rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def model():
    from nerf_navigation_b200.nerf.network import NeRFNetwork
    torch.manual_seed(0)
    m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10).cuda()
    m.encoder.embeddings.data.uniform_(-0.5, 0.5)
    m.density_grid.copy_(torch.from_numpy(stonehenge_grid()))
    from nerf_navigation_b200 import raymarching as rm
    m.density_bitfield = rm.packbits(m.density_grid, 10.0, m.density_bitfield)
    return m


def test_state_dict_keys_match_reference(model):
    keys = set(model.state_dict().keys())
    want = {"encoder.embeddings", "encoder.offsets", "sigma_net.0.weight", "sigma_net.1.weight", "color_net.0.weight",
            "color_net.1.weight", "color_net.2.weight", "aabb_train", "aabb_infer", "density_grid", "density_bitfield",
            "step_counter"}
    assert want <= keys, want - keys
    assert tuple(model.sigma_net[0].weight.shape) == (64, 32) and tuple(model.color_net[0].weight.shape) == (64, 31)


def test_fused_density_equals_unfused(model):
    x = (torch.rand(20000, 3, device="cuda") * 4 - 2)
    model.eval()
    xa = x.clone().requires_grad_()
    da = model.density(xa)
    (da["sigma"] ** 2).mean().add((da["geo_feat"] * 0.01).sum()).backward()
    model.fused_eval = False
    xb = x.clone().requires_grad_()
    db = model.density(xb)
    (db["sigma"] ** 2).mean().add((db["geo_feat"] * 0.01).sum()).backward()
    model.fused_eval = True
    assert rel_err(da["sigma"].detach().cpu().numpy(), db["sigma"].detach().cpu().numpy()) < 1e-5
    assert rel_err(da["geo_feat"].detach().cpu().numpy(), db["geo_feat"].detach().cpu().numpy()) < 1e-5
    assert rel_err(xa.grad.cpu().numpy(), xb.grad.cpu().numpy()) < 1e-5


def test_hessian_does_not_raise(model):
    """estimator_helpers.py:384 calls autograd.functional.hessian through the encoder: must not error."""
    model.eval()
    f = lambda p: (model.density(p.view(1, 3) + torch.linspace(0, 0.1, 8, device="cuda").view(8, 1))["sigma"]).sum()
    h = torch.autograd.functional.hessian(f, torch.zeros(3, device="cuda"))
    assert h.shape == (3, 3) and torch.isfinite(h).all()


def test_run_torch_sampler_fwd_bwd_to_pose(model):
    """cuda_ray=False path (the one simulate.py actually uses): 512 samples/ray, gradient reaches rays_o/rays_d."""
    model.eval()
    o, d = orbit_rays(256, 1)
    ro = torch.from_numpy(o).cuda().view(1, -1, 3).requires_grad_(); rd = torch.from_numpy(d).cuda().view(1, -1, 3).requires_grad_()
    model.cuda_ray = False
    try:
        out = model.render(ro, rd, staged=True, bg_color=1.0, perturb=False, num_steps=512, upsample_steps=0, max_ray_batch=4096)
        model.fused_eval = False
        ref = model.render(ro, rd, staged=True, bg_color=1.0, perturb=False, num_steps=512, upsample_steps=0, max_ray_batch=4096)
        model.fused_eval = True
    finally:
        model.cuda_ray = True
    assert out["image"].shape == (1, 256, 3) and out["depth"].shape == (1, 256)
    assert rel_err(out["image"].detach().cpu().numpy(), ref["image"].detach().cpu().numpy()) < 1e-5
    tgt = torch.rand(1, 256, 3, device="cuda")
    ((out["image"] - tgt) ** 2).mean().backward()
    ga, gb = ro.grad.clone(), rd.grad.clone()
    ro.grad = None; rd.grad = None
    ((ref["image"] - tgt) ** 2).mean().backward()
    assert ga.abs().max() > 0
    # fused vs unfused are two fp32 evaluation orders of a ReLU net summed over 131072 samples with a 1e-4-scale table:
    # a pre-activation within rounding of 0 flips its ReLU bit and moves dL/dx of that sample by a whole term, so the
    # ray-summed pose gradient agrees to ~1e-4, not 1e-5 (the per-point 1e-5 bar is test_field_gpu.py vs the oracle).
    assert rel_err(ga.cpu().numpy(), ro.grad.cpu().numpy()) < 5e-4 and rel_err(gb.cpu().numpy(), rd.grad.cpu().numpy()) < 5e-4


def test_ray_gradient_bar_is_the_fp32_floor(model, ngp):
    """The pose gradient (dL/d rays_o, dL/d rays_d of the photometric loss, 512 samples per ray) against a FLOAT64 oracle, on
    a 64-ray subset (32,768 samples): util.run_f64 evaluates the reference's definition of the field — sample depths,
    box clipping, grid cells and interpolation weights taken at the fp32 positions the reference itself computes
    (gridencoder.cu:134-138), everything after that exact — with torch autograd in double.  Three fp32 evaluations are
    measured against it: ours (fused kernels), the unfused op kernels, and the reference's own CUDA build driven as its
    Python does (oracle/ref_pipeline.ref_run, when oracle/_ref is in the tree).  Asserted: ours meets north_star's 1e-5
    against fp64, and is not further from it than 1.5x the reference build / the unfused path.  (The looser 5e-4 bars of the
    fused-vs-unfused and golden comparisons in this file are between TWO fp32 evaluation orders over 4x as many rays, where
    a ReLU or cell-face tie moves one sample's gradient by a whole term: this test is the pinned statement.)"""
    from util import run_f64, load_ref
    model.eval()
    n = 64
    o, d = orbit_rays(n, 9)
    ro32, rd32 = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    tgt = torch.rand(n, 3, device="cuda")
    from nerf_navigation_b200 import raymarching as rm
    near, far = rm.near_far_from_aabb(ro32, rd32, model.aabb_infer, model.min_near)
    s = torch.empty(16, device="cuda")
    ngp.check(ngp.lib().ngp_grid_level_scales(s.data_ptr(), 16, float(np.log2(model.encoder.per_level_scale)), 16, ngp.stream()))
    a, b = ro32.double().requires_grad_(), rd32.double().requires_grad_()
    img64 = run_f64(model, a, b, near, far, s.cpu().numpy().astype(np.float64))
    ((img64 - tgt.double()) ** 2).mean().backward()
    g64 = torch.cat([a.grad, b.grad], -1)

    def grad32(render):
        x, y = ro32.clone().requires_grad_(), rd32.clone().requires_grad_()
        img = render(x, y)
        ((img - tgt) ** 2).mean().backward()
        return torch.cat([x.grad, y.grad], -1).double(), img.detach()

    def dist(g):
        return float((g - g64).abs().max() / g64.abs().max())

    kw = dict(staged=True, bg_color=1.0, perturb=False, num_steps=512, upsample_steps=0, max_ray_batch=4096)
    model.cuda_ray = False
    try:
        g_ours, img_ours = grad32(lambda x, y: model.render(x[None], y[None], **kw)["image"][0])
        model.fused_eval = False
        g_unf, _ = grad32(lambda x, y: model.render(x[None], y[None], **kw)["image"][0])
    finally:
        model.fused_eval = True
        model.cuda_ray = True
    assert rel_err(img_ours.cpu().numpy(), img64.detach().cpu().numpy()) < 1e-5          # the VALUES meet the fp32 bar
    e_ours, e_unf = dist(g_ours), dist(g_unf)
    others = [e_unf]
    if load_ref("gridencoder") is not None and load_ref("shencoder") is not None and load_ref("raymarching") is not None:
        from oracle import ref_pipeline as RP
        field = RP.RefField(model.encoder.embeddings.detach(), model.encoder.offsets, model.encoder.per_level_scale,
                            [l.weight.detach() for l in model.sigma_net], [l.weight.detach() for l in model.color_net], 2.0)
        g_ref, _ = grad32(lambda x, y: RP.ref_run(field, x, y, model.aabb_infer, model.min_near, num_steps=512)[0])
        others.append(dist(g_ref))
    print(f"ray-gradient distance from fp64: ours {e_ours:.2e}, unfused {e_unf:.2e}" +
          (f", reference CUDA build {others[-1]:.2e}" if len(others) > 1 else ""))
    assert e_ours < 1e-5
    assert e_ours <= 1.5 * max(others) + 2e-6


def test_run_cuda_train_and_eval_agree(model):
    """Same rays through the training branch (march_rays_train + composite_rays_train) and the inference loop
    (march_rays / composite_rays): identical samples => images agree to fp32 rounding."""
    o, d = orbit_rays(4096, 2)
    ro = torch.from_numpy(o).cuda().view(1, -1, 3); rd = torch.from_numpy(d).cuda().view(1, -1, 3)
    model.train()
    with torch.no_grad():
        tr = model.render(ro, rd, bg_color=1.0, perturb=False, force_all_rays=True, dt_gamma=0, max_steps=1024)
    model.eval()
    with torch.no_grad():
        ev = model.render(ro, rd, bg_color=1.0, perturb=False, dt_gamma=0, max_steps=1024)
    assert rel_err(tr["image"].cpu().numpy(), ev["image"].cpu().numpy()) < 1e-5
    # depth is NOT compared: the reference's training compositor integrates t from 0 (raymarching.cu:552) while the
    # inference one starts at rays_t = near (:855), so the two branches differ by design.
    assert (ev["image"] < 0.999).any()          # the scene is actually hit


def test_training_step_param_grads(model):
    """run_cuda train branch + MSE backward reaches table and MLP parameters (nerf/utils.py:446-450)."""
    o, d = orbit_rays(1024, 3)
    ro = torch.from_numpy(o).cuda().view(1, -1, 3); rd = torch.from_numpy(d).cuda().view(1, -1, 3)
    model.train()
    model.zero_grad()
    out = model.render(ro, rd, bg_color=1.0, perturb=True, force_all_rays=False, dt_gamma=0, max_steps=1024)
    ((out["image"] - 0.5) ** 2).mean().backward()
    model.eval()
    g = model.encoder.embeddings.grad
    assert g is not None and torch.isfinite(g).all() and (g != 0).any()
    assert all(p.grad is not None and (p.grad != 0).any() for p in list(model.sigma_net.parameters()) + list(model.color_net.parameters()))


def test_update_extra_state_runs(model):
    import copy
    m = copy.deepcopy(model)
    m.reset_extra_state()
    m.update_extra_state()
    assert m.iter_density == 1 and m.density_bitfield.dtype == torch.uint8 and m.mean_density > 0


def test_network_ff_variant():
    """nerf/network_ff.py mirror: FFMLP parameters with the reference names/sizes, forward/density/color contracts, and
    agreement with a torch restatement of the same fp16 network (2e-3 bar)."""
    from nerf_navigation_b200.nerf.network_ff import NeRFNetwork as NetFF
    m = NetFF(encoding="hashgrid", bound=2, cuda_ray=False).cuda().eval()
    m.encoder.embeddings.data.uniform_(-0.5, 0.5)
    keys = set(m.state_dict().keys())
    assert {"encoder.embeddings", "encoder.offsets", "sigma_net.weights", "color_net.weights"} <= keys
    assert m.sigma_net.weights.numel() == 64 * (32 + 64 + 16) and m.color_net.weights.numel() == 64 * (32 + 128 + 16)
    x = torch.rand(5000, 3, device="cuda") * 2 - 1
    d = torch.nn.functional.normalize(torch.randn(5000, 3, device="cuda"), dim=-1)
    with torch.no_grad():
        sigma, rgb = m(x, d)
        dens = m.density(x)
        rgb2 = m.color(x, d, geo_feat=dens["geo_feat"])
        # torch restatement
        enc = m.encoder(x, bound=2).half()
        Ws = m.sigma_net.weights.half()
        h = torch.relu(enc.float() @ Ws[:2048].view(64, 32).float().t()).half()
        h = torch.relu(h.float() @ Ws[2048:6144].view(64, 64).float().t()).half()
        o = (h.float() @ Ws[6144:].view(16, 64).float().t()).half()
        sig_ref = torch.exp(o[:, 0].float())
        cin = torch.cat([m.encoder_dir(d).half(), o[:, 1:], torch.zeros_like(o[:, :1])], -1)
        Wc = m.color_net.weights.half()
        c = torch.relu(cin.float() @ Wc[:2048].view(64, 32).float().t()).half()
        c = torch.relu(c.float() @ Wc[2048:6144].view(64, 64).float().t()).half()
        c = torch.relu(c.float() @ Wc[6144:10240].view(64, 64).float().t()).half()
        rgb_ref = torch.sigmoid((c.float() @ Wc[10240:].view(16, 64).float().t()).half()[:, :3])
    assert sigma.shape == (5000,) and rgb.shape == (5000, 3)
    assert rel_err(sigma.float().cpu().numpy(), sig_ref.cpu().numpy()) < 2e-3
    assert rel_err(rgb.float().cpu().numpy(), rgb_ref.float().cpu().numpy()) < 2e-3
    assert rel_err(dens["sigma"].float().cpu().numpy(), sigma.float().cpu().numpy()) < 1e-6
    assert rel_err(rgb2.float().cpu().numpy(), rgb.float().cpu().numpy()) < 1e-6


def _sample_points(n, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.rand(n, 3, device="cuda", generator=g) * 4.4 - 2.2          # some points fall outside bound=2
    d = torch.nn.functional.normalize(torch.randn(n, 3, device="cuda", generator=g), dim=-1)
    return x, d


def _relu_tie_samples(model, x, d, tol=1e-5):
    """Samples with a hidden pre-activation within `tol` (relative to the layer's largest) of zero, computed with plain
    torch matmuls on the unfused encoder outputs: the sign of such a unit — hence the sample's gradient — can differ between
    two fp32 evaluation orders.  They are named and excluded from the 1e-5 gradient bar (and counted)."""
    with torch.no_grad():
        feat = model.encoder(x, bound=model.bound)
        z0 = feat @ model.sigma_net[0].weight.t()
        o = torch.relu(z0) @ model.sigma_net[1].weight.t()
        cin = torch.cat([model.encoder_dir(d), o[:, 1:]], -1)
        z1 = cin @ model.color_net[0].weight.t()
        z2 = torch.relu(z1) @ model.color_net[1].weight.t()
        tie = torch.zeros(x.shape[0], dtype=torch.bool, device=x.device)
        for z in (z0, z1, z2):
            tie |= ((z.abs() < tol * z.abs().max()) & (z != 0)).any(-1)   # exact zeros (points outside the box) are not ties
    return tie


@pytest.mark.parametrize("impl", ["tc5", "mma"])
@pytest.mark.parametrize("n", [1, 191, 5000, 40001])
def test_fused_forward_xd_equals_unfused(model, ngp, n, impl):
    """Fused sample kernels (encode + sigma MLP + SH + colour MLP: tcgen05 csrc/cfield5.cu, mma.sync csrc/renderfield.cu)
    == the unfused path (GridEncoder / nn.Linear / SHEncoder kernels, themselves pinned to the oracle + reference CUDA
    build): values and the gradients w.r.t. sample positions and view directions (1e-5 bar for the fp32 path; samples
    with a ReLU pre-activation at a tie are named by _relu_tie_samples and excluded from the gradient bar)."""
    model.eval()
    x, d = _sample_points(n, n)
    w_s = torch.randn(n, device="cuda"); w_c = torch.randn(n, 3, device="cuda")
    outs = []
    prev = ngp.lib().ngp_set_field_impl(1 if impl == "tc5" else 0)
    try:
        for fused in (True, False):
            model.fused_eval = fused
            xa, da = x.clone().requires_grad_(), d.clone().requires_grad_()
            sigma, rgb = model(xa, da)
            ((sigma * w_s).sum() + (rgb * w_c).sum()).backward()
            outs.append((sigma.detach(), rgb.detach(), xa.grad.clone(), da.grad.clone()))
    finally:
        model.fused_eval = True
        ngp.lib().ngp_set_field_impl(prev)
    (s1, c1, gx1, gd1), (s0, c0, gx0, gd0) = outs
    assert rel_err(s1.cpu().numpy(), s0.cpu().numpy()) < 1e-5
    assert rel_err(c1.cpu().numpy(), c0.cpu().numpy()) < 1e-5
    keep = (~_relu_tie_samples(model, x, d)).cpu().numpy()
    assert keep.mean() > 0.99
    for a, b in ((gx1, gx0), (gd1, gd0)):
        err = (a - b).abs().amax(-1).cpu().numpy() / max(float(b.abs().max()), 1e-30)
        assert err[keep].max(initial=0.0) < 1e-5, (impl, float(err[keep].max(initial=0.0)), int((err > 1e-5).sum()))


def test_fused_forward_xd_vs_oracle(model, oracle, ngp):
    """The fused sample kernel against the CPU oracle chain (field_query -> sh_encode -> mlp_forward -> sigmoid)."""
    model.eval()
    n = 3000
    x, d = _sample_points(n, 7)
    with torch.no_grad():
        sigma, rgb = model(x, d)
    enc = model.encoder
    s = torch.empty(16, device="cuda")
    ngp.check(ngp.lib().ngp_grid_level_scales(s.data_ptr(), 16, float(np.log2(enc.per_level_scale)), 16, ngp.stream()))
    W = [l.weight.detach().cpu().numpy() for l in model.sigma_net] + [l.weight.detach().cpu().numpy() for l in model.color_net]
    rs, rg, _ = oracle.field_query(x.cpu().numpy(), 2.0, enc.embeddings.detach().cpu().numpy(), enc.offsets.cpu().numpy(),
                                   s.cpu().numpy(), W[0], W[1], np.zeros(n, np.float32))
    sh, _ = oracle.sh_encode(d.cpu().numpy(), 4)
    h = np.concatenate([sh, rg], -1).astype(np.float32)
    for i, w in enumerate(W[2:]):
        h = h @ w.T
        if i < 2:
            h = np.maximum(h, 0)
    ref_rgb = 1.0 / (1.0 + np.exp(-h.astype(np.float64)))
    assert rel_err(sigma.cpu().numpy(), rs) < 1e-5
    assert rel_err(rgb.cpu().numpy(), ref_rgb) < 1e-5


@pytest.mark.parametrize("impl", ["tc5", "mma"])
def test_fused_backward_xd_vs_oracle(model, oracle, ngp, impl):
    """k_cfield5_backward / k_render_field_backward (dL/dx, dL/dd of the fused sample kernel) directly against the CPU
    oracle chain: orc_field_query backward + orc_sh_encode Jacobian + the colour-MLP chain rule in float64
    (oracle.field_color).  1e-5 on every sample except ones the oracle itself names as a ReLU tie: a hidden
    pre-activation within 1e-5 of zero relative to the layer's scale, whose sign two fp32 evaluation orders may disagree on."""
    model.eval()
    n = 6000
    x, d = _sample_points(n, 23)
    g = torch.Generator(device="cuda").manual_seed(5)
    w_s = torch.randn(n, device="cuda", generator=g)
    w_c = torch.randn(n, 3, device="cuda", generator=g)
    prev = ngp.lib().ngp_set_field_impl(1 if impl == "tc5" else 0)
    try:
        xa, da = x.clone().requires_grad_(), d.clone().requires_grad_()
        sigma, rgb = model(xa, da)
        torch.autograd.backward([sigma, rgb], [w_s, w_c])
        torch.cuda.synchronize()
    finally:
        ngp.lib().ngp_set_field_impl(prev)
    enc = model.encoder
    s = torch.empty(16, device="cuda")
    ngp.check(ngp.lib().ngp_grid_level_scales(s.data_ptr(), 16, float(np.log2(enc.per_level_scale)), 16, ngp.stream()))
    Ws = [l.weight.detach().cpu().numpy() for l in model.sigma_net]
    Wc = [l.weight.detach().cpu().numpy() for l in model.color_net]
    ref = oracle.field_color(x.cpu().numpy(), d.cpu().numpy(), 2.0, enc.embeddings.detach().cpu().numpy(),
                             enc.offsets.cpu().numpy(), s.cpu().numpy(), Ws, Wc, w_s.cpu().numpy(), w_c.cpu().numpy())
    assert rel_err(sigma.detach().cpu().numpy(), ref["sigma"]) < 1e-5
    assert rel_err(rgb.detach().cpu().numpy(), ref["rgb"]) < 1e-5
    tie = np.zeros(n, bool)
    for z in ref["z"]:
        tie |= (np.abs(z) < 1e-5 * np.abs(z).max()).any(-1)
    for ours, want in ((xa.grad, ref["grad_x"]), (da.grad, ref["grad_d"])):
        err = np.abs(ours.cpu().numpy().astype(np.float64) - want).max(-1) / np.abs(want).max()
        bad = err > 1e-5
        assert not (bad & ~tie).any(), (impl, float(err[~tie].max()), int(bad.sum()))
    assert tie.sum() < n // 100


def test_cfg1_frequency_network_vs_reference_golden():
    """BASELINE config 1 through the host mirror ON THE GPU (NeRFNetwork(frequency), cuda_ray=False, render staged ->
    run -> density / masked color) against tests/golden/cfg1.npz, which scripts/make_golden.py produced by importing the
    real reference.  Pins the host logic of N3/N5 (renderer.py:125-254, 542-575) and near_far.  fp32: 1e-5 on values;
    5e-4 on the ray gradients (|g| ~ 1e-5 summed over 524,288 samples of sin/cos(32 x) terms, CPU libm vs GPU libdevice)."""
    import os
    from nerf_navigation_b200.nerf.network import NeRFNetwork
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cfg1.npz"))
    m = NeRFNetwork(encoding="frequency", encoding_dir="frequency", bound=2, cuda_ray=False, density_scale=1, min_near=0.2,
                    density_thresh=10, bg_radius=-1)
    sd = {k[3:].replace("__", "."): torch.from_numpy(g[k]) for k in g.files if k.startswith("sd_")}
    missing = m.load_state_dict(sd, strict=True)
    m = m.cuda().eval()
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        x = torch.from_numpy(g["x"]).cuda().requires_grad_(True)
        sigma = m.density(x.reshape(-1, 3))["sigma"]
        (sigma ** 2).mean().backward()
        assert rel_err(sigma.detach().cpu().numpy(), g["sigma"]) < 1e-5
        assert rel_err(x.grad.cpu().numpy(), g["dx"]) < 1e-5
        ro = torch.from_numpy(g["rays_o"]).cuda().requires_grad_(True)
        rd = torch.from_numpy(g["rays_d"]).cuda().requires_grad_(True)
        out = m.render(ro, rd, staged=True, bg_color=1.0, perturb=False, num_steps=512, upsample_steps=0, max_ray_batch=4096)
        loss = ((out["image"] - torch.from_numpy(g["target"]).cuda()) ** 2).mean()
        loss.backward()
        assert rel_err(out["image"].detach().cpu().numpy(), g["image"]) < 1e-5
        assert rel_err(out["depth"].detach().cpu().numpy(), g["depth"]) < 1e-5
        assert abs(loss.item() - float(g["loss"])) < 1e-5 * float(g["loss"])
        assert rel_err(ro.grad.cpu().numpy(), g["d_rays_o"]) < 5e-4 and rel_err(rd.grad.cpu().numpy(), g["d_rays_d"]) < 5e-4
        with torch.no_grad():
            up = m.render(ro[:, :64], rd[:, :64], staged=True, bg_color=1.0, perturb=False, num_steps=64, upsample_steps=32,
                          max_ray_batch=4096)
        assert rel_err(up["image"].cpu().numpy(), g["image_up"]) < 1e-4 and rel_err(up["depth"].cpu().numpy(), g["depth_up"]) < 1e-4
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


def test_planner_closures(model):
    """simulate.py:343 density closure and quad_plot.py:64-83 A* occupancy: fused query == unfused query."""
    from nerf_navigation_b200.closures import make_density_fn, astar_occupancy
    model.eval()
    fn = make_density_fn(model)
    occ = astar_occupancy(fn, side=40, kernel_size=5, thresh=1.0)
    assert occ.shape == (8, 8, 8) and occ.dtype == torch.bool
    model.fused_eval = False
    ref = astar_occupancy(fn, side=40, kernel_size=5, thresh=1.0)
    model.fused_eval = True
    assert (occ != ref).float().mean() < 0.01       # a cell whose max sits within rounding of the threshold may flip
    x = torch.rand(7, 5, 3, device="cuda", requires_grad=True)
    s = fn(x)
    s.sum().backward()
    assert s.shape == (7, 5) and x.grad.abs().sum() > 0


def test_update_extra_state_uses_fused_query_in_training_mode(model):
    """update_extra_state runs under no_grad while the model trains: the 2 x 128^3 sweep must take the fused kernel."""
    import copy
    m = copy.deepcopy(model).train()
    x = torch.zeros(4, 3, device="cuda")
    assert not m._use_fused_density(x)          # parameters require grad: autograd path
    with torch.no_grad():
        assert m._use_fused_density(x)


def test_reference_checkpoint_layout_loads(model):
    """nerf/utils.py:938-1059 checkpoint format: {'model': state_dict, ...} with the reference's keys/shapes loads strictly;
    a missing density grid (checkpoints saved with cuda_ray=False) loads non-strictly like the reference (:1006)."""
    from nerf_navigation_b200.nerf.network import NeRFNetwork
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    ckpt = {"epoch": 3, "global_step": 1234, "model": sd}
    m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10).cuda()
    missing, unexpected = m.load_state_dict(ckpt["model"], strict=False)
    assert not missing and not unexpected
    assert torch.equal(m.encoder.embeddings, model.encoder.embeddings) and torch.equal(m.density_bitfield, model.density_bitfield)
    sd2 = {k: v for k, v in sd.items() if not k.startswith("density") and k != "step_counter"}
    missing, unexpected = m.load_state_dict(sd2, strict=False)
    assert set(missing) == {"density_grid", "density_bitfield", "step_counter"} and not unexpected


def test_single_pass_density_closure(model):
    """closures.make_density_fn(single_pass=True): sigma and dL/dx from ONE kernel == the two-kernel autograd path, for
    the planner's cost shape (quad_plot.py:224-250: mean over the body cloud of sigma^2 * distance)."""
    from nerf_navigation_b200.closures import make_density_fn
    model.eval()
    x = (torch.rand(20, 500, 3, device="cuda") * 1.6 - 0.8)
    dist = torch.rand(20, 1, device="cuda") + 0.5
    outs = []
    for sp in (True, False):
        xi = x.clone().requires_grad_(True)
        sig = make_density_fn(model, single_pass=sp)(xi)
        cost = (torch.mean(sig ** 2 * dist, dim=-1) * 1e6).mean()
        cost.backward()
        outs.append((sig.detach(), xi.grad.clone(), cost.item()))
    assert torch.equal(outs[0][0], outs[1][0])
    assert rel_err(outs[0][1].cpu().numpy(), outs[1][1].cpu().numpy()) < 1e-5
    assert abs(outs[0][2] - outs[1][2]) <= 1e-6 * abs(outs[1][2])


def test_fused_kernels_serve_autocast_regions():
    """Inside torch.autocast (the reference Trainer's fp16 eval) the fused kernels still run — in fp32, more accurate than
    the half GridEncoder + half Linear path they replace — and return the dtypes that path returns: sigma fp32
    (activation.py:7), geo_feat / rgb in the autocast dtype.  Values against the unfused autocast path at the fp16 bar."""
    from nerf_navigation_b200.nerf.network import NeRFNetwork
    from nerf_navigation_b200.nerf.network_ff import NeRFNetwork as NetFF
    torch.manual_seed(3)
    x = torch.rand(20000, 3, device="cuda") * 2 - 1
    d = torch.nn.functional.normalize(torch.randn(20000, 3, device="cuda"), dim=-1)
    m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=False).cuda().eval()
    m.encoder.embeddings.data.uniform_(-0.5, 0.5)
    outs = {}
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.float16):
        for fused in (True, False):
            m.fused_eval = fused
            dens = m.density(x)
            sigma, rgb = m(x, d)
            outs[fused] = (dens["sigma"], dens["geo_feat"], sigma, rgb)
    for a, b in zip(outs[True], outs[False]):
        assert a.dtype == b.dtype and a.shape == b.shape
        assert rel_err(a.float().cpu().numpy(), b.float().cpu().numpy()) < 5e-3
    assert outs[True][0].dtype == torch.float32 and outs[True][1].dtype == torch.float16 and outs[True][3].dtype == torch.float16
    # and against the fp32 result (what autocast approximates): the fused path is the closer of the two
    with torch.no_grad():
        m.fused_eval = True
        ref = m.density(x)["geo_feat"]
    e_f = rel_err(outs[True][1].float().cpu().numpy(), ref.cpu().numpy())
    e_u = rel_err(outs[False][1].float().cpu().numpy(), ref.cpu().numpy())
    assert e_f <= e_u + 1e-6 and e_f < 1e-3
    mf = NetFF(encoding="hashgrid", bound=2, cuda_ray=False).cuda().eval()
    mf.encoder.embeddings.data.uniform_(-0.5, 0.5)
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.float16):
        mf.fused_eval = True
        a = mf.density(x)
        mf.fused_eval = False
        b = mf.density(x)
    assert a["sigma"].dtype == b["sigma"].dtype and a["geo_feat"].dtype == b["geo_feat"].dtype
    assert rel_err(a["sigma"].float().cpu().numpy(), b["sigma"].float().cpu().numpy()) < 5e-3
    assert rel_err(a["geo_feat"].float().cpu().numpy(), b["geo_feat"].float().cpu().numpy()) < 5e-3
