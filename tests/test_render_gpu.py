"""GPU: fused inference renderer (ngp_render_rays_infer: tcgen05 csrc/render5.cu, mma.sync csrc/render.cu) == the reference's eval-branch loop
(nerf/renderer.py:325-374: march_rays -> forward(x, d) -> composite_rays -> compaction), checked against
 (a) the op-by-op loop of the host mirror (whose ops are pinned to the reference CUDA build), and
 (b) the CPU oracle chain (orc_march_rays / field / orc_composite_rays) run as that same loop."""
import numpy as np
import pytest
import torch

from util import stonehenge_grid, orbit_rays, rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def model():
    from nerf_navigation_b200.nerf.network import NeRFNetwork
    from nerf_navigation_b200 import raymarching as rm
    torch.manual_seed(0)
    # density_scale 25 and a wide log-sigma range: a good part of the rays reaches T < 1e-4 well before the end of its march
    # (early termination), the rest runs to the far plane
    m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=25, min_near=0.2, density_thresh=10).cuda()
    m.encoder.embeddings.data.uniform_(-0.5, 0.5)
    m.sigma_net[1].weight.data[0].abs_().mul_(6.0)
    m.density_grid.copy_(torch.from_numpy(stonehenge_grid()))
    m.density_bitfield = rm.packbits(m.density_grid, 10.0, m.density_bitfield)
    return m.eval()


@pytest.fixture(params=["tc5", "tc5-overflow", "tc5-lists", "tc5-inkernel-march", "mma"])
def impl(request, ngp):
    """Every test runs on all implementations: the tcgen05 kernel fed by pre-marched runs of samples (default), the same
    with room for ONE run per ray (so most rays take the overflow path: the marcher-inside kernel over a ray list, in the
    same call), the tcgen05 kernel fed by explicit sample lists, the tcgen05 kernel with the marcher inside, and the
    mma.sync kernel."""
    from nerf_navigation_b200 import field
    prev = ngp.lib().ngp_set_field_impl(0 if request.param == "mma" else 1)
    prev_pm, prev_cap = field.PREMARCH, field.SEGMENT_CAPACITY
    field.PREMARCH = {"tc5": "segments", "tc5-overflow": "segments", "tc5-lists": "lists"}.get(request.param, False)
    if request.param == "tc5-overflow":
        field.SEGMENT_CAPACITY = 1
    yield request.param
    field.PREMARCH, field.SEGMENT_CAPACITY = prev_pm, prev_cap
    ngp.lib().ngp_set_field_impl(prev)


def _one_chunk(model, ro, rd, max_steps=1024, bg=1.0):
    """The reference loop with ONE chunk of max_steps samples per ray, on the op kernels (march_rays -> forward(x, d) ->
    composite_rays; each pinned to the reference CUDA build): the form of the loop whose t has no n_step-dependent
    re-summation wobble.  Returns image, depth, weights_sum, and per ray: samples marched, samples composited (float64
    transmittance, first T < 1e-4 inclusive — raymarching.cu:880), and `tie` = a transmittance within 0.4 % of the 1e-4
    threshold: the kernels evaluate T = 1 - weights_sum in fp32, whose rounding near weights_sum = 0.9999 is 6e-8 absolute,
    i.e. ~1e-3 of the threshold, so only such rays can have their count moved by rounding."""
    from nerf_navigation_b200 import raymarching as rm
    n = ro.shape[0]
    dev = ro.device
    with torch.no_grad():
        nears, fars = rm.near_far_from_aabb(ro, rd, model.aabb_infer, model.min_near)
        alive = torch.arange(n, dtype=torch.int32, device=dev)
        rays_t = nears.clone()
        xyzs, dirs, deltas = rm.march_rays(n, max_steps, alive, rays_t, ro, rd, model.bound, model.density_bitfield,
                                           model.cascade, model.grid_size, nears, fars, -1, False, 0, max_steps)
        sig = torch.zeros(n * max_steps, device=dev); rgb = torch.zeros(n * max_steps, 3, device=dev)
        used = deltas[:, 0] > 0
        model.fused_eval = False
        try:
            s_u, c_u = model(xyzs[used], dirs[used])
        finally:
            model.fused_eval = True
        sig[used] = s_u * model.density_scale; rgb[used] = c_u
        ws = torch.zeros(n, device=dev); dep = torch.zeros(n, device=dev); img = torch.zeros(n, 3, device=dev)
        rm.composite_rays(n, max_steps, alive, rays_t, sig, rgb, deltas, ws, dep, img)
        img = img + (1 - ws).unsqueeze(-1) * bg
        dep = torch.clamp(dep - nears, min=0) / (fars - nears)
        marched = used.view(n, max_steps).sum(-1)
        a = 1.0 - torch.exp(-(sig.double() * deltas[:, 0].double()).view(n, max_steps))
        T = torch.cumprod(torch.cat([torch.ones(n, 1, device=dev, dtype=torch.float64), 1.0 - a], -1), -1)[:, :-1]  # before sample k
        k = torch.arange(max_steps, device=dev).unsqueeze(0)
        live = k < marched.unsqueeze(-1)
        dead_at = torch.where((T < 1e-4) & live, k, torch.full_like(k, max_steps)).amin(-1)   # first sample that sees T < 1e-4
        comp = torch.minimum(marched, dead_at + 1)
        tie = (((T / 1e-4 - 1.0).abs() < 4e-3) & live).any(-1)
    return img, dep, ws, marched.int(), comp.int(), tie


def _loop(model, ro, rd, **kw):
    model.fused_render = False
    try:
        with torch.no_grad():
            return model.render(ro, rd, **kw)
    finally:
        model.fused_render = True


@pytest.mark.parametrize("n,seed", [(1, 0), (37, 1), (4096, 2), (50000, 3)])
def test_fused_render_equals_loop(model, impl, n, seed):
    o, d = orbit_rays(n, seed)
    ro = torch.from_numpy(o).cuda().view(1, -1, 3); rd = torch.from_numpy(d).cuda().view(1, -1, 3)
    kw = dict(bg_color=1.0, perturb=False, dt_gamma=0, max_steps=1024)
    with torch.no_grad():
        out = model.render(ro, rd, **kw)
    ref = _loop(model, ro, rd, **kw)
    assert out["image"].shape == (1, n, 3) and out["depth"].shape == (1, n)
    img, rimg = out["image"].cpu().numpy(), ref["image"].cpu().numpy()
    dep, rdep = out["depth"].cpu().numpy(), ref["depth"].cpu().numpy()
    # tolerance: fp32 path, 1e-5 relative (north_star) on every ray the reference loop itself renders consistently.  The
    # loop re-sums t from the deltas of its n_step chunks (raymarching.cu:876), an ulp-level wobble that can move a sample
    # across a cell boundary; such rays are NAMED: they are the ones on which the loop differs from its own one-chunk form
    # (no re-summation), and only they may differ from the fused kernel (which is held to 1e-5 against the one-chunk form
    # on all rays in test_fused_render_counts_and_one_chunk).
    oimg, odep, _, _, _, _ = _one_chunk(model, ro.view(-1, 3), rd.view(-1, 3))
    oimg, odep = oimg.cpu().numpy(), odep.cpu().numpy()
    scale = max(1.0, np.abs(rimg).max())
    wobble = np.abs(oimg - rimg[0]).max(-1) > 1e-6 * scale
    with np.errstate(invalid="ignore"):
        wobble |= np.nan_to_num(np.abs(odep - rdep[0]), nan=0.0) > 1e-6 * max(1.0, np.nanmax(np.abs(rdep)))
    bad = np.abs(img - rimg)[0].max(-1) > 1e-5 * scale
    assert not (bad & ~wobble).any(), (int(bad.sum()), int(wobble.sum()))
    assert wobble.mean() < 5e-3, wobble.mean()
    ok = np.isfinite(rdep)
    assert (np.isfinite(dep) == ok).all()
    badd = np.zeros_like(bad)
    badd[ok[0]] = np.abs(dep[ok] - rdep[ok]) > 1e-5 * max(1.0, np.abs(rdep[ok]).max())
    assert not (badd & ~wobble).any(), (int(badd.sum()), int(wobble.sum()))
    if n >= 4096:
        assert (img < 0.999).any()           # the scene is hit
        assert (img > 0.99999).all(-1).any()  # and some rays see only background


def test_fused_render_miss_and_background(model, impl):
    """Rays that miss the AABB: image = bg, weights_sum = 0, depth = 0/0 = NaN exactly like renderer.py:372-373."""
    from nerf_navigation_b200.field import render_rays_infer
    o = torch.tensor([[5.0, 5.0, 5.0], [0.0, 0.0, 1.5]], device="cuda")
    d = torch.nn.functional.normalize(torch.tensor([[1.0, -1.0, 0.0], [0.0, 0.0, -1.0]], device="cuda"), dim=-1)
    bg = torch.tensor([0.25, 0.5, 0.75], device="cuda")
    image, depth, ws, stats = render_rays_infer(model, o, d, model.aabb_infer, bg, return_stats=True)
    assert torch.equal(image[0], bg) and ws[0] == 0 and torch.isnan(depth[0])
    assert ws[1] > 0 and torch.isfinite(depth[1]) and int(stats[1].item()) > 0
    # per-ray background
    bgn = torch.rand(2, 3, device="cuda")
    image2, _, ws2 = render_rays_infer(model, o, d, model.aabb_infer, bgn)
    assert torch.equal(image2[0], bgn[0]) and torch.equal(ws2, ws)


def test_fused_render_vs_oracle(model, oracle, ngp, impl):
    """Against the CPU oracle run as the reference loop with one chunk of max_steps samples per ray."""
    n, max_steps = 600, 1024
    o, d = orbit_rays(n, 11)
    from nerf_navigation_b200.field import render_rays_infer
    ro, rd = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    image, depth, ws, stats, counts = render_rays_infer(model, ro, rd, model.aabb_infer, 1.0, max_steps=max_steps,
                                                        return_stats=True, return_counts=True)
    enc = model.encoder
    s = torch.empty(16, device="cuda")
    ngp.check(ngp.lib().ngp_grid_level_scales(s.data_ptr(), 16, float(np.log2(enc.per_level_scale)), 16, ngp.stream()))
    W = [l.weight.detach().cpu().numpy() for l in model.sigma_net] + [l.weight.detach().cpu().numpy() for l in model.color_net]
    bits = model.density_bitfield.cpu().numpy()
    nears, fars = oracle.near_far_from_aabb(o, d, model.aabb_infer.cpu().numpy(), 0.2)
    alive = np.arange(n, dtype=np.int32)
    rays_t = nears.copy()
    xyzs, dirs, deltas = oracle.march_rays(n, max_steps, alive, rays_t, o, d, 2.0, bits, 2, 128, nears, fars, 0.0, max_steps)
    used = deltas[:, 0] > 0
    rs = np.zeros(len(xyzs), np.float32); rgb = np.zeros((len(xyzs), 3), np.float32)
    sig, geo, _ = oracle.field_query(xyzs[used], 2.0, enc.embeddings.detach().cpu().numpy(), enc.offsets.cpu().numpy(),
                                     s.cpu().numpy(), W[0], W[1], np.zeros(int(used.sum()), np.float32))
    sh, _ = oracle.sh_encode(dirs[used], 4)
    h = np.concatenate([sh, geo], -1).astype(np.float32)
    for i, w in enumerate(W[2:]):
        h = h @ w.T
        if i < 2:
            h = np.maximum(h, 0)
    rs[used] = sig * float(model.density_scale)      # renderer.py:361
    rgb[used] = (1.0 / (1.0 + np.exp(-h.astype(np.float64)))).astype(np.float32)
    wsum = np.zeros(n, np.float32); dep = np.zeros(n, np.float32); img = np.zeros((n, 3), np.float32)
    oracle.composite_rays(n, max_steps, alive, rays_t, rs, rgb, deltas, wsum, dep, img)
    img = img + (1 - wsum)[:, None] * 1.0
    with np.errstate(invalid="ignore", divide="ignore"):
        dep = np.clip(dep - nears, 0, None) / (fars - nears)
    assert rel_err(ws.cpu().numpy(), wsum) < 1e-5
    assert rel_err(image.cpu().numpy(), img) < 1e-5
    ok = np.isfinite(dep)
    assert rel_err(depth.cpu().numpy()[ok], dep[ok]) < 1e-5
    assert 0 < int(stats[1].item()) <= int(used.sum())      # early termination evaluates no more samples than marched
    # the integer output: samples composited per ray, bit-exact against the oracle's samples walked by the compositing rule
    # (fp32, in order, stop after the first sample that sees T = 1 - weights_sum < 1e-4: raymarching.cu:860-893)
    want = np.zeros(n, np.int32)
    tie = np.zeros(n, bool)
    dl = deltas.reshape(n, max_steps, 2); sg = rs.reshape(n, max_steps)
    for i in range(n):
        acc = np.float32(0.0)
        for k in range(max_steps):
            dt = dl[i, k, 0]
            if dt == 0:
                break
            T = np.float32(1.0) - acc
            acc = np.float32(acc + np.float32(np.float32(1.0) - np.exp(np.float32(-sg[i, k] * dt))) * T)
            want[i] += 1
            tie[i] |= abs(float(T) / 1e-4 - 1.0) < 4e-3
            if float(T) < 1e-4:
                break
    got = counts.cpu().numpy()
    assert int(stats[1].item()) == int(got.sum())
    assert np.array_equal(got[~tie], want[~tie]) and tie.sum() <= n // 10, (int((got != want).sum()), int(tie.sum()))
    assert (want < used.reshape(n, max_steps).sum(-1)).any()      # early termination is exercised


def test_fused_render_counts_and_one_chunk(model, impl):
    """20,000 rays against the op kernels run as the reference loop in its one-chunk form: image / depth / weights_sum at 1e-5
    on EVERY ray (no allowance), per-ray sample counts bit-exact (except rays whose transmittance sits on the 1e-4
    threshold, named by _one_chunk), and equal to march_rays_train's counts for rays that do not terminate early."""
    from nerf_navigation_b200.field import render_rays_infer
    from nerf_navigation_b200 import raymarching as rm
    n = 20000
    o, d = orbit_rays(n, 21)
    ro, rd = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    image, depth, ws, stats, counts = render_rays_infer(model, ro, rd, model.aabb_infer, 1.0, return_stats=True,
                                                        return_counts=True)
    oimg, odep, ows, marched, comp, tie = _one_chunk(model, ro, rd)
    assert rel_err(image.cpu().numpy(), oimg.cpu().numpy()) < 1e-5
    assert rel_err(ws.cpu().numpy(), ows.cpu().numpy()) < 1e-5
    ok = torch.isfinite(odep)
    assert torch.equal(torch.isfinite(depth), ok)
    assert rel_err(depth[ok].cpu().numpy(), odep[ok].cpu().numpy()) < 1e-5
    assert torch.equal(counts[~tie], comp[~tie]), int((counts != comp).sum())
    assert int(tie.sum()) < n // 10
    assert int(stats[1].item()) == int(counts.sum())
    # rays that never see T < 1e-4 take exactly the samples the training marcher emits for them
    nears, fars = rm.near_far_from_aabb(ro, rd, model.aabb_infer, model.min_near)
    cnt = torch.zeros(2, dtype=torch.int32, device="cuda")
    _, _, _, rays = rm.march_rays_train(ro, rd, model.bound, model.density_bitfield, model.cascade, model.grid_size, nears,
                                        fars, cnt, 0, False, 128, True, 0, 1024)
    train_counts = torch.zeros(n, dtype=torch.int32, device="cuda")
    train_counts[rays[:, 0].long()] = rays[:, 2]
    full = (comp == marched) & ~tie
    assert full.sum() > n // 10 and (~full).sum() > n // 100
    assert torch.equal(counts[full], train_counts[full])


def test_fused_render_sample_cap(model, impl):
    """max_steps caps the samples a ray takes (and changes dt_min): the kernel must stop there and still write every ray."""
    from nerf_navigation_b200.field import render_rays_infer
    o, d = orbit_rays(2048, 5)
    ro, rd = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    image, depth, ws, stats, counts = render_rays_infer(model, ro, rd, model.aabb_infer, 1.0, max_steps=16, return_stats=True,
                                                        return_counts=True)
    assert torch.isfinite(image).all() and int(stats[1].item()) <= 16 * 2048 and int(stats[0].item()) >= 2048
    assert int(counts.max()) <= 16 and int(counts.sum()) == int(stats[1].item())
