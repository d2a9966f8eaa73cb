"""GPU: fused inference renderer (csrc/render.cu, ngp_render_rays_infer) == the reference's eval-branch loop
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
    m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10).cuda()
    m.encoder.embeddings.data.uniform_(-0.5, 0.5)
    # denser field than the random init (sigma ~ e^3) so that early termination (T < 1e-4) is exercised
    m.sigma_net[1].weight.data[0].abs_().mul_(6.0)
    m.density_grid.copy_(torch.from_numpy(stonehenge_grid()))
    m.density_bitfield = rm.packbits(m.density_grid, 10.0, m.density_bitfield)
    return m.eval()


def _loop(model, ro, rd, **kw):
    model.fused_render = False
    try:
        with torch.no_grad():
            return model.render(ro, rd, **kw)
    finally:
        model.fused_render = True


@pytest.mark.parametrize("n,seed", [(1, 0), (37, 1), (4096, 2), (50000, 3)])
def test_fused_render_equals_loop(model, n, seed):
    o, d = orbit_rays(n, seed)
    ro = torch.from_numpy(o).cuda().view(1, -1, 3); rd = torch.from_numpy(d).cuda().view(1, -1, 3)
    kw = dict(bg_color=1.0, perturb=False, dt_gamma=0, max_steps=1024)
    with torch.no_grad():
        out = model.render(ro, rd, **kw)
    ref = _loop(model, ro, rd, **kw)
    assert out["image"].shape == (1, n, 3) and out["depth"].shape == (1, n)
    img, rimg = out["image"].cpu().numpy(), ref["image"].cpu().numpy()
    dep, rdep = out["depth"].cpu().numpy(), ref["depth"].cpu().numpy()
    # tolerance: fp32 path, 1e-5 relative (north_star); the reference loop's own t wobbles by an ulp with its n_step
    # chunking (render.cu header), which can move a sample across a cell boundary for a handful of rays
    bad = np.abs(img - rimg).max(-1) > 1e-5 * max(1.0, np.abs(rimg).max())
    assert bad.mean() < 2e-3, bad.mean()
    assert rel_err(img[~bad], rimg[~bad]) < 1e-5
    ok = np.isfinite(rdep)
    assert (np.isfinite(dep) == ok).all()
    badd = np.abs(dep[ok] - rdep[ok]) > 1e-5 * max(1.0, np.abs(rdep[ok]).max())
    assert badd.mean() < 2e-3, badd.mean()
    if n >= 4096:
        assert (img < 0.999).any()           # the scene is hit
        assert (img > 0.99999).all(-1).any()  # and some rays see only background


def test_fused_render_miss_and_background(model):
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


def test_fused_render_vs_oracle(model, oracle, ngp):
    """Against the CPU oracle run as the reference loop with one chunk of max_steps samples per ray."""
    n, max_steps = 600, 1024
    o, d = orbit_rays(n, 11)
    from nerf_navigation_b200.field import render_rays_infer
    ro, rd = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    image, depth, ws, stats = render_rays_infer(model, ro, rd, model.aabb_infer, 1.0, max_steps=max_steps,
                                                return_stats=True)
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
    rs[used] = sig
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


def test_fused_render_sample_cap(model):
    """max_steps caps the samples a ray takes (and changes dt_min): the kernel must stop there and still write every ray."""
    from nerf_navigation_b200.field import render_rays_infer
    o, d = orbit_rays(2048, 5)
    ro, rd = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    image, depth, ws, stats = render_rays_infer(model, ro, rd, model.aabb_infer, 1.0, max_steps=16, return_stats=True)
    assert torch.isfinite(image).all() and int(stats[1].item()) <= 16 * 2048 and int(stats[0].item()) >= 2048
