"""GPU: BASELINE.json's FULL sizes, checked through size-independent properties (the oracle cannot run them in seconds)
plus oracle parity on random subsamples.
  config 2: 4,194,304 points, uniform and trajectory-clustered: permutation equivariance and chunk
            consistency (bit-exact: a point's result must not depend on its neighbours in the batch), backward linearity,
            out-of-range points, 4096-point subsample vs the CPU oracle.
  config 3: 1,048,576 rays: offsets are the exclusive scan of the counts (ray order), sum = sample count, composited
            weights in [0, 1], 2048-ray subsample vs the CPU oracle (bit-exact counts).
  config 4: one 800x800 frame (640,000 rays) through the fused renderer: run-to-run and implementation-to-implementation
            bit equality (pre-marched runs, runs with overflow, explicit lists, marcher inside the kernel), permutation
            equivariance (a ray's pixel must not depend on which rays share its tiles), counts summing to the samples
            evaluated, and the pre-pass plan against the composited counts."""
import numpy as np
import pytest
import torch

from util import hashgrid_config, sigma_net_weights, rel_err, stonehenge_grid, orbit_rays

pytestmark = pytest.mark.gpu

B_FULL = 1 << 22


def _points(kind, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    if kind == "uniform":
        return torch.rand(B_FULL, 3, device="cuda", generator=g) * 2 - 1
    # trajectory-clustered: 8192 waypoints on a helix (radius 0.6, z in [0.1, 0.6]), each with the 8x8x8 body cloud
    # linspace(+-0.05) x linspace(+-0.05) x linspace(+-0.02) (SURVEY §8d cfg 2 (ii); body cloud of quad_plot.py:224-250)
    k = torch.arange(8192, device="cuda", dtype=torch.float32) / 8192
    way = torch.stack([0.6 * torch.cos(6 * np.pi * k), 0.6 * torch.sin(6 * np.pi * k), 0.1 + 0.5 * k], -1)
    lx = torch.linspace(-0.05, 0.05, 8, device="cuda"); lz = torch.linspace(-0.02, 0.02, 8, device="cuda")
    body = torch.stack(torch.meshgrid(lx, lx, lz, indexing="ij"), -1).reshape(-1, 3)
    return (way[:, None, :] + body[None, :, :]).reshape(-1, 3).contiguous()


class _Field:
    def __init__(self, ngp):
        self.ngp, self.L = ngp, ngp.lib()
        self.cfg = hashgrid_config(bound=2.0, table_std=0.5)
        self.W0n, self.W1n = sigma_net_weights()
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
        self.tab, self.off, self.W0, self.W1 = t(self.cfg["table"]), t(self.cfg["offsets"]), t(self.W0n), t(self.W1n)

    def fwd(self, x):
        B = x.shape[0]
        sigma = torch.empty(B, device="cuda"); geo = torch.empty(B, 15, device="cuda")
        mask = torch.empty(B, device="cuda", dtype=torch.int64)
        c = self.cfg
        self.ngp.check(self.L.ngp_field_query_forward(x.data_ptr(), c["bound"], self.tab.data_ptr(), self.off.data_ptr(), c["S"],
                                                      c["H"], 0, 0, self.W0.data_ptr(), self.W1.data_ptr(), sigma.data_ptr(),
                                                      geo.data_ptr(), mask.data_ptr(), B, self.ngp.stream()), "fwd")
        return sigma, geo, mask

    def bwd(self, x, sigma, mask, gs):
        B = x.shape[0]
        gx = torch.empty(B, 3, device="cuda")
        c = self.cfg
        self.ngp.check(self.L.ngp_field_query_backward(x.data_ptr(), c["bound"], self.tab.data_ptr(), self.off.data_ptr(), c["S"],
                                                       c["H"], 0, 0, self.W0.data_ptr(), self.W1.data_ptr(), sigma.data_ptr(),
                                                       mask.data_ptr(), gs.data_ptr(), None, gx.data_ptr(), B,
                                                       self.ngp.stream()), "bwd")
        return gx


    def fwd_jac(self, x):
        B = x.shape[0]
        sigma = torch.empty(B, device="cuda"); geo = torch.empty(B, 15, device="cuda"); jac = torch.empty(B, 3, device="cuda")
        mask = torch.empty(B, device="cuda", dtype=torch.int64)
        c = self.cfg
        self.ngp.check(self.L.ngp_field_query_forward_jacobian(x.data_ptr(), c["bound"], self.tab.data_ptr(), self.off.data_ptr(),
                                                               c["S"], c["H"], 0, 0, self.W0.data_ptr(), self.W1.data_ptr(),
                                                               sigma.data_ptr(), geo.data_ptr(), mask.data_ptr(), jac.data_ptr(), B,
                                                               self.ngp.stream()), "fwd_jac")
        return sigma, geo, mask, jac


@pytest.fixture(scope="module")
def field(ngp):
    return _Field(ngp)


@pytest.mark.parametrize("kind", ["uniform", "clustered"])
def test_config2_single_table_pass_full_size(field, kind):
    """The headline kernel (sigma, geo_feat, ReLU bits and d sigma / dx from ONE table pass) at 4,194,304 points: same
    forward bits as the forward kernel, Jacobian rows x dL/dsigma == the second-table-pass backward (1e-6), permutation
    equivariance and ragged-chunk consistency of the Jacobian (bit-exact), and the saved bits drive the backward kernel."""
    x = _points(kind)
    sigma, geo, mask = field.fwd(x)
    sj, gj, mj, jac = field.fwd_jac(x)
    assert torch.equal(sj, sigma) and torch.equal(gj, geo) and torch.equal(mj, mask)
    gs = sigma * (2.0 / B_FULL)
    gx = field.bwd(x, sigma, mask, gs)
    assert rel_err((jac * gs[:, None]).cpu().numpy(), gx.cpu().numpy()) < 1e-6
    perm = torch.randperm(B_FULL, device="cuda", generator=torch.Generator(device="cuda").manual_seed(4))
    assert torch.equal(field.fwd_jac(x[perm].contiguous())[3], jac[perm])
    a, n = 333_333, 2_000_001
    assert torch.equal(field.fwd_jac(x[a:a + n].contiguous())[3], jac[a:a + n])
    assert torch.equal(field.bwd(x, sj, mj, gs), gx)


@pytest.mark.parametrize("kind", ["uniform", "clustered"])
def test_config2_full_size_properties(field, oracle, ngp, kind):
    x = _points(kind)
    assert x.shape == (B_FULL, 3)
    sigma, geo, mask = field.fwd(x)
    gs = sigma * (2.0 / B_FULL)
    gx = field.bwd(x, sigma, mask, gs)
    assert torch.isfinite(sigma).all() and torch.isfinite(gx).all() and (sigma > 0).all()
    # permutation equivariance, bit-exact
    perm = torch.randperm(B_FULL, device="cuda", generator=torch.Generator(device="cuda").manual_seed(1))
    xp = x[perm].contiguous()
    sp, gp, mp = field.fwd(xp)
    assert torch.equal(sp, sigma[perm]) and torch.equal(gp, geo[perm])
    assert torch.equal(field.bwd(xp, sp, mp, gs[perm].contiguous()), gx[perm])
    # chunk consistency (ragged chunk that does not align with the kernel's 64-point tiles), bit-exact
    a, n = 777_777, 1_000_003
    sc, gc, mc = field.fwd(x[a:a + n].contiguous())
    assert torch.equal(sc, sigma[a:a + n]) and torch.equal(gc, geo[a:a + n])
    assert torch.equal(field.bwd(x[a:a + n].contiguous(), sc, mc, gs[a:a + n].contiguous()), gx[a:a + n])
    # backward is linear in the upstream gradient: exact for a power-of-two scale, 1e-5 for a sum
    assert torch.equal(field.bwd(x, sigma, mask, gs * 4.0), gx * 4.0)
    g2 = torch.randn(B_FULL, device="cuda", generator=torch.Generator(device="cuda").manual_seed(2)) / B_FULL
    both = field.bwd(x, sigma, mask, gs + g2)
    parts = gx + field.bwd(x, sigma, mask, g2)
    assert rel_err(both.cpu().numpy(), parts.cpu().numpy()) < 1e-5
    # subsample vs the CPU oracle
    idx = torch.randint(0, B_FULL, (4096,), device="cuda", generator=torch.Generator(device="cuda").manual_seed(3))
    s16 = torch.empty(16, device="cuda")
    c = field.cfg
    ngp.check(ngp.lib().ngp_grid_level_scales(s16.data_ptr(), 16, c["S"], c["H"], ngp.stream()))
    rs, rg, rgx = oracle.field_query(x[idx].cpu().numpy(), c["bound"], c["table"], c["offsets"], s16.cpu().numpy(), field.W0n,
                                     field.W1n, gs[idx].cpu().numpy())
    assert rel_err(sigma[idx].cpu().numpy(), rs) < 1e-5 and rel_err(geo[idx].cpu().numpy(), rg) < 1e-5
    assert rel_err(gx[idx].cpu().numpy(), rgx) < 1e-5


def test_config2_out_of_range_points(field):
    """gridencoder.cu:99-123: outside [0,1]^3 the features are zero -> sigma = exp(0), geo = 0, zero gradient; the
    boundary itself is inside."""
    x = torch.tensor([[2.0, 2.0, 2.0], [-2.0, -2.0, -2.0], [2.0001, 0, 0], [0, -3.0, 0], [0, 0, 1e9], [0.1, 0.2, 0.3]], device="cuda")
    sigma, geo, mask = field.fwd(x)
    gx = field.bwd(x, sigma, mask, torch.ones(6, device="cuda"))
    assert (sigma[2:5] == 1.0).all() and (geo[2:5] == 0).all() and (gx[2:5] == 0).all()
    assert (sigma[[0, 1, 5]] != 1.0).all() and gx[5].abs().sum() > 0


def test_config3_full_size_march_and_composite(ngp, oracle):
    from nerf_navigation_b200 import raymarching as rm
    BOUND, CAS, H = 2.0, 2, 128
    bits_np = oracle.packbits(stonehenge_grid(CAS, H, BOUND), 10.0)
    bits = torch.from_numpy(bits_np).cuda()
    N = 1 << 20
    o, d = orbit_rays(4096, 5)
    # 256 jittered copies of 4096 pixel rays (the hypotheses of config 3 share pixels, differ in pose)
    g = torch.Generator(device="cuda").manual_seed(0)
    ro = torch.from_numpy(o).cuda().repeat(256, 1) + 0.02 * torch.randn(N, 3, device="cuda", generator=g)
    rd = torch.nn.functional.normalize(torch.from_numpy(d).cuda().repeat(256, 1) + 0.02 * torch.randn(N, 3, device="cuda", generator=g), dim=-1)
    aabb = torch.tensor([-BOUND] * 3 + [BOUND] * 3, device="cuda")
    nears, fars = rm.near_far_from_aabb(ro, rd, aabb, 0.2)
    counter = torch.zeros(2, dtype=torch.int32, device="cuda")
    xyzs, dirs, deltas, rays = rm.march_rays_train(ro, rd, BOUND, bits, CAS, H, nears, fars, counter, -1, False, 128, True, 0, 1024)
    cnt = rays[:, 2].long(); offs = rays[:, 1].long()
    m = int(counter[0])
    assert torch.equal(rays[:, 0].long(), torch.arange(N, device="cuda"))            # ray order
    assert torch.equal(offs, torch.cumsum(cnt, 0) - cnt) and int(cnt.sum()) == m       # exclusive scan, total
    assert int(counter[1]) == N and (cnt <= 1024).all() and m > 20 * N
    assert (xyzs[:m].abs() <= BOUND).all() and (deltas[:m, 0] > 0).all() and not xyzs[m:].any()
    # compositing: weights in [0, 1], image in [0, 1] for colours in [0, 1]
    M = xyzs.shape[0]
    sig = torch.rand(M, device="cuda", generator=g) * 30; rgb = torch.rand(M, 3, device="cuda", generator=g)
    ws, depth, image = rm.composite_rays_train(sig, rgb, deltas, rays)
    assert (ws >= 0).all() and (ws <= 1 + 1e-5).all() and (image >= 0).all() and (image <= 1 + 1e-5).all()
    assert (ws[cnt == 0] == 0).all()
    # subsample vs the oracle: bit-exact counts and positions
    sel = torch.randint(0, N, (2048,), device="cuda", generator=g)
    so, sd = ro[sel].cpu().numpy(), rd[sel].cpu().numpy()
    rx, _, rdl, rrays, rcnt = oracle.march_rays_train(so, sd, BOUND, bits_np, CAS, H, nears[sel].cpu().numpy(),
                                                      fars[sel].cpu().numpy(), 2048 * 1024)
    assert np.array_equal(rrays[:, 2], cnt[sel].cpu().numpy())
    for j in (0, 17, 2047):
        a, n = int(offs[sel[j]]), int(cnt[sel[j]])
        ra = int(rrays[j, 1])
        assert np.array_equal(xyzs[a:a + n].cpu().numpy(), rx[ra:ra + n])
        assert np.array_equal(deltas[a:a + n].cpu().numpy(), rdl[ra:ra + n])


def test_config4_full_frame_properties(ngp):
    from nerf_navigation_b200 import field, raymarching as rm
    from nerf_navigation_b200.nerf.network import NeRFNetwork
    torch.manual_seed(0)
    m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=25, min_near=0.2, density_thresh=10).cuda().eval()
    m.encoder.embeddings.data.uniform_(-0.5, 0.5)
    m.density_grid.copy_(torch.from_numpy(stonehenge_grid()))
    m.density_bitfield = rm.packbits(m.density_grid, 10.0, m.density_bitfield)
    N = 800 * 800
    o, d = orbit_rays(N, 3)
    ro, rd = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
    ren = lambda a, b, mode: field.render_rays_infer(m, a, b, m.aabb_infer, 1.0, premarch=mode, return_stats=True, return_counts=True)
    img, dep, ws, stats, cnt = ren(ro, rd, "segments")
    assert torch.isfinite(img).all() and (ws >= 0).all() and (ws <= 1 + 1e-5).all() and (cnt >= 0).all() and (cnt <= 1024).all()
    assert int(stats[1]) == int(cnt.sum()) and int(stats[26]) == 0             # samples evaluated; no ray needed > 32 runs
    assert (ws > 0.9999).float().mean() > 0.05                                  # density_scale 25: early termination is exercised
    # the same frame again, and through the other three feeds of the same kernel core: bit-identical
    prev_cap = field.SEGMENT_CAPACITY
    try:
        for mode, cap in (("segments", 32), ("segments", 2), ("lists", 32), (False, 32)):
            field.SEGMENT_CAPACITY = cap
            i2, d2, w2, st2, c2 = ren(ro, rd, mode)
            assert torch.equal(i2, img) and torch.equal(d2, dep) and torch.equal(w2, ws) and torch.equal(c2, cnt), (mode, cap)
            if mode == "segments" and cap == 2:
                assert int(st2[26]) > 1000                                      # the overflow path really carried rays
    finally:
        field.SEGMENT_CAPACITY = prev_cap
    # permutation equivariance
    perm = torch.randperm(N, device="cuda")
    ip, dp, wp, _, cp = ren(ro[perm].contiguous(), rd[perm].contiguous(), "segments")
    assert torch.equal(ip, img[perm]) and torch.equal(dp, dep[perm]) and torch.equal(cp, cnt[perm])
    # the pre-pass plan: samples marched per ray >= samples composited, equal where the ray did not terminate early
    L = ngp.lib()
    plan = torch.empty(N, device="cuda", dtype=torch.int32); ends = torch.empty(N, device="cuda", dtype=torch.int64)
    wb = L.ngp_render_plan_workspace_bytes(N); wsb = torch.empty(wb, device="cuda", dtype=torch.uint8)
    ngp.check(L.ngp_render_plan(ro.data_ptr(), rd.data_ptr(), N, m.aabb_infer.data_ptr(), float(m.min_near),
                                m.density_bitfield.data_ptr(), 2.0, 0.0, 1024, 2, 128, plan.data_ptr(), ends.data_ptr(),
                                wsb.data_ptr(), wb, ngp.stream()), "plan")
    assert torch.equal(ends, torch.cumsum(plan.long(), 0)) and (plan >= cnt).all()
    open_rays = ws < 0.999
    assert torch.equal(plan[open_rays], cnt[open_rays])
