"""GPU parity: ray ops (C ABI via the raymarching mirror) vs the CPU oracle and vs the reference's CUDA build."""
import numpy as np
import pytest
import torch

from util import stonehenge_grid, orbit_rays, rel_err, load_ref

pytestmark = pytest.mark.gpu

BOUND, CAS, H = 2.0, 2, 128


@pytest.fixture(scope="module")
def scene(oracle):
    grid = stonehenge_grid(CAS, H, BOUND)
    bits = oracle.packbits(grid, 10.0)
    return grid, bits


def _rays(n, seed=0):
    o, d = orbit_rays(n, seed)
    return o, d


def test_packbits_morton_bit_exact(ngp, oracle, scene):
    from nerf_navigation_b200 import raymarching as rm
    grid, bits = scene
    out = rm.packbits(torch.from_numpy(grid).cuda(), 10.0)
    assert np.array_equal(out.cpu().numpy(), bits)
    rng = np.random.default_rng(0)
    coords = rng.integers(0, 1024, (5000, 3)).astype(np.int32)
    idx = rm.morton3D(torch.from_numpy(coords).cuda())
    assert np.array_equal(idx.cpu().numpy(), oracle.morton3D(coords))
    back = rm.morton3D_invert(idx)
    assert np.array_equal(back.cpu().numpy(), coords)


def test_near_far_vs_oracle(ngp, oracle):
    from nerf_navigation_b200 import raymarching as rm
    o, d = _rays(20000, 1)
    d[:5] = [[0, 0, 1], [1, 0, 0], [0, -1, 0], [0.6, 0.8, 0], [0, 0, -1]]   # axis-parallel: 1/0 = inf paths
    o[5:10] = [5.0, 5.0, 5.0]                                                # outside, most miss
    aabb = np.array([-BOUND] * 3 + [BOUND] * 3, np.float32)
    rn, rf = oracle.near_far_from_aabb(o, d, aabb, 0.2)
    n, f = rm.near_far_from_aabb(torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda(), torch.from_numpy(aabb).cuda(), 0.2)
    assert np.array_equal(n.cpu().numpy(), rn, equal_nan=True) and np.array_equal(f.cpu().numpy(), rf, equal_nan=True)


@pytest.mark.parametrize("perturb", [False, True])
def test_march_rays_train_vs_oracle(ngp, oracle, scene, perturb):
    from nerf_navigation_b200 import raymarching as rm
    _, bits = scene
    N = 8192
    o, d = _rays(N, 2)
    aabb = np.array([-BOUND] * 3 + [BOUND] * 3, np.float32)
    nears, fars = oracle.near_far_from_aabb(o, d, aabb, 0.2)
    M = N * 1024
    rx, rd, rdl, rrays, rcnt = oracle.march_rays_train(o, d, BOUND, bits, CAS, H, nears, fars, M, perturb=perturb)
    t = lambda a: torch.from_numpy(a).cuda()
    counter = torch.zeros(2, dtype=torch.int32, device="cuda")
    xyzs, dirs, deltas, rays = rm.march_rays_train(t(o), t(d), BOUND, t(bits), CAS, H, t(nears), t(fars), counter, -1,
                                                   perturb, 128, True, 0, 1024)
    assert np.array_equal(rays.cpu().numpy(), rrays)                 # ids, offsets, counts bit-exact
    assert np.array_equal(counter.cpu().numpy(), rcnt)
    m = int(rcnt[0])
    assert m > 1000 and xyzs.shape[0] == m + (128 - m % 128)
    assert np.array_equal(xyzs[:m].cpu().numpy(), rx[:m])
    assert np.array_equal(dirs[:m].cpu().numpy(), rd[:m])
    assert np.array_equal(deltas[:m].cpu().numpy(), rdl[:m])
    assert not xyzs[m:].any()


@pytest.mark.parametrize("N,dt_gamma,max_steps", [(3000, 1.0 / 128, 1024), (8192, 1.0 / 64, 512), (20000, 1.0 / 128, 1024),
                                                  (20000, 0.0, 1024), (33, 0.0, 16)])
def test_march_rays_train_both_kernels_and_cone_steps(ngp, oracle, scene, N, dt_gamma, max_steps):
    """The marcher has two forms — one warp per ray for N <= 8192 (32 consecutive steps probed at once, the reference's
    decisions replayed on the probes) and one thread per ray above — and both must reproduce the reference's samples bit for
    bit, also with dt_gamma > 0 where the step grows along the ray (raymarching.cu:369) and with a tight max_steps."""
    from nerf_navigation_b200 import raymarching as rm
    _, bits = scene
    o, d = _rays(N, 7)
    aabb = np.array([-BOUND] * 3 + [BOUND] * 3, np.float32)
    nears, fars = oracle.near_far_from_aabb(o, d, aabb, 0.2)
    M = N * max_steps
    rx, rd, rdl, rrays, rcnt = oracle.march_rays_train(o, d, BOUND, bits, CAS, H, nears, fars, M, dt_gamma=dt_gamma,
                                                       max_steps=max_steps, perturb=True)
    t = lambda a: torch.from_numpy(a).cuda()
    counter = torch.zeros(2, dtype=torch.int32, device="cuda")
    xyzs, dirs, deltas, rays = rm.march_rays_train(t(o), t(d), BOUND, t(bits), CAS, H, t(nears), t(fars), counter, -1,
                                                   True, 128, True, dt_gamma, max_steps)
    assert np.array_equal(rays.cpu().numpy(), rrays) and np.array_equal(counter.cpu().numpy(), rcnt)
    m = int(rcnt[0])
    assert m > 0 and (rrays[:, 2] <= max_steps).all()
    assert np.array_equal(xyzs[:m].cpu().numpy(), rx[:m]) and np.array_equal(deltas[:m].cpu().numpy(), rdl[:m])
    if N == 20000:   # the first 8192 rays through the warp-per-ray kernels: same counts and samples as the thread form
        c2 = torch.zeros(2, dtype=torch.int32, device="cuda")
        n2 = 8192
        x2, _, dl2, r2 = rm.march_rays_train(t(o[:n2]), t(d[:n2]), BOUND, t(bits), CAS, H, t(nears[:n2]), t(fars[:n2]), c2, -1,
                                             True, 128, True, dt_gamma, max_steps)
        assert torch.equal(r2, rays[:n2])
        m2 = int(c2[0])
        assert torch.equal(x2[:m2], xyzs[:m2]) and torch.equal(dl2[:m2], deltas[:m2])


@pytest.mark.parametrize("grid_kind,dt_gamma,max_steps", [("scene", 1.0 / 128, 256), ("scene", 0.0, 256), ("noise", 0.0, 256),
                                                          ("noise", 1.0 / 64, 256), ("full", 0.0, 3000), ("noise", 0.0, 3000)])
def test_march_write_pass_from_runs(ngp, oracle, scene, grid_kind, dt_gamma, max_steps, monkeypatch):
    """Thread-per-ray form (N > 8192): the count pass leaves each ray's runs of consecutive samples and the write pass
    (k_march_replay) rebuilds the samples from them without reading the grid.  Bit for bit the oracle's samples, and the
    samples of the write pass that marches again (NGP_B200_MARCH_WRITE=remarch), directions included.  A random
    occupancy grid gives most rays more runs than MARCH_SEG_CAP (16): those are marched again by their warp (march_ray_warp).  The
    one-call form with a sample budget drops the same rays as the reference (raymarching.cu:419-420).  max_steps = 3000
    gives rays with more samples than the kernel's 1024-entry shared-memory strip (a fully occupied grid: one long run
    per ray; the random grid: the warp's own march)."""
    from nerf_navigation_b200 import raymarching as rm
    if grid_kind == "scene":
        bits = scene[1]
    elif grid_kind == "full":
        bits = np.full(CAS * H ** 3 // 8, 255, np.uint8)
    else:
        bits = np.random.default_rng(5).integers(0, 256, size=CAS * H ** 3 // 8, dtype=np.uint8)
    N = 9000
    o, d = _rays(N, 11)
    aabb = np.array([-BOUND] * 3 + [BOUND] * 3, np.float32)
    nears, fars = oracle.near_far_from_aabb(o, d, aabb, 0.2)
    rx, rd, rdl, rrays, rcnt = oracle.march_rays_train(o, d, BOUND, bits, CAS, H, nears, fars, N * max_steps, dt_gamma=dt_gamma,
                                                       max_steps=max_steps, perturb=True)
    m = int(rcnt[0])
    assert m > 0
    t = lambda a: torch.from_numpy(a).cuda()

    def run(mean_count, force_all):
        counter = torch.zeros(2, dtype=torch.int32, device="cuda")
        out = rm.march_rays_train(t(o), t(d), BOUND, t(bits), CAS, H, t(nears), t(fars), counter, mean_count, True, 128,
                                  force_all, dt_gamma, max_steps)
        return out, counter

    (xyzs, dirs, deltas, rays), counter = run(-1, True)
    assert np.array_equal(rays.cpu().numpy(), rrays) and np.array_equal(counter.cpu().numpy(), rcnt)
    assert np.array_equal(xyzs[:m].cpu().numpy(), rx[:m]) and np.array_equal(deltas[:m].cpu().numpy(), rdl[:m])
    assert np.array_equal(dirs[:m].cpu().numpy(), rd[:m])
    if grid_kind == "noise":      # the point of the random grid: rays with more than 16 runs
        cnt = rrays[:, 2]
        gaps = np.array([np.count_nonzero(rdl[a:a + c, 1] > 1.5 * rdl[a:a + c, 0]) for _, a, c in rrays[:200]])   # run starts
        assert cnt.max() > 16 and gaps.max() > 16
    if max_steps > 1024:
        assert rrays[:, 2].max() > 1024
    budget = (m // 2) - (m // 2) % 128
    (bx, bd, bdl, brays), _ = run(budget, False)          # one call, count + scan + write, M = budget + 128
    monkeypatch.setenv("NGP_B200_MARCH_WRITE", "remarch")
    (x2, d2, dl2, r2), _ = run(-1, True)
    (bx2, bd2, bdl2, brays2), _ = run(budget, False)
    monkeypatch.delenv("NGP_B200_MARCH_WRITE")
    assert torch.equal(r2, rays) and torch.equal(x2, xyzs) and torch.equal(d2, dirs) and torch.equal(dl2, deltas)
    assert torch.equal(brays2, brays) and torch.equal(bx2, bx) and torch.equal(bd2, bd) and torch.equal(bdl2, bdl)
    M = bx.shape[0]
    kept = (rrays[:, 1].astype(np.int64) + rrays[:, 2] < M) & (rrays[:, 2] > 0)
    assert 0 < kept.sum() < N
    last = rrays[kept][-1]
    end = int(last[1] + last[2])
    assert np.array_equal(bx[:end].cpu().numpy(), rx[:end]) and float(bx[end:].abs().max()) == 0.0


def test_march_budget_drops_rays_like_reference(ngp, oracle, scene):
    """mean_count budget: rays whose offset+count >= M write nothing (raymarching.cu:420)."""
    from nerf_navigation_b200 import raymarching as rm
    _, bits = scene
    N = 2048
    o, d = _rays(N, 3)
    aabb = np.array([-BOUND] * 3 + [BOUND] * 3, np.float32)
    nears, fars = oracle.near_far_from_aabb(o, d, aabb, 0.2)
    M = 4096
    rx, rd, rdl, rrays, rcnt = oracle.march_rays_train(o, d, BOUND, bits, CAS, H, nears, fars, M)
    assert rcnt[0] > M
    t = lambda a: torch.from_numpy(a).cuda()
    counter = torch.zeros(2, dtype=torch.int32, device="cuda")
    xyzs, dirs, deltas, rays = rm.march_rays_train(t(o), t(d), BOUND, t(bits), CAS, H, t(nears), t(fars), counter, M - 128,
                                                   False, 128, False, 0, 1024)
    assert xyzs.shape[0] == M
    assert np.array_equal(rays.cpu().numpy(), rrays)
    assert np.array_equal(xyzs.cpu().numpy(), rx) and np.array_equal(deltas.cpu().numpy(), rdl)


def test_composite_train_fwd_bwd_vs_oracle(ngp, oracle, scene):
    from nerf_navigation_b200 import raymarching as rm
    _, bits = scene
    N = 4096
    o, d = _rays(N, 4)
    aabb = np.array([-BOUND] * 3 + [BOUND] * 3, np.float32)
    nears, fars = oracle.near_far_from_aabb(o, d, aabb, 0.2)
    xyzs, dirs, deltas, rays, cnt = oracle.march_rays_train(o, d, BOUND, bits, CAS, H, nears, fars, N * 1024)
    m = int(cnt[0]); m_al = m + (128 - m % 128)
    deltas = deltas[:m_al]
    rng = np.random.default_rng(0)
    sig = (rng.uniform(0, 1, m_al) ** 4 * 300).astype(np.float32)    # some rays saturate -> early termination path
    rgb = rng.uniform(0, 1, (m_al, 3)).astype(np.float32)
    ws, dep, img = oracle.composite_rays_train_forward(sig, rgb, deltas, rays)
    gws = rng.standard_normal(N).astype(np.float32); gim = rng.standard_normal((N, 3)).astype(np.float32)
    rgs, rgc = oracle.composite_rays_train_backward(gws, gim, sig, rgb, deltas, rays, ws, img)
    t = lambda a: torch.from_numpy(a).cuda()
    st, rt = t(sig).requires_grad_(), t(rgb).requires_grad_()
    w2, d2, i2 = rm.composite_rays_train(st, rt, t(deltas), t(rays))
    (w2 * t(gws)).sum().add((i2 * t(gim)).sum()).backward()
    # __expf (device, like the reference CUDA) vs expf (oracle): 2 ulp on alpha
    assert rel_err(w2.detach().cpu().numpy(), ws) < 1e-5
    assert rel_err(d2.detach().cpu().numpy(), dep) < 1e-5
    assert rel_err(i2.detach().cpu().numpy(), img) < 1e-5
    assert rel_err(st.grad.cpu().numpy(), rgs) < 2e-5
    assert rel_err(rt.grad.cpu().numpy(), rgc) < 1e-5


def test_inference_loop_vs_oracle(ngp, oracle, scene):
    """march_rays / composite_rays alive-ray loop (renderer.py:343-369) with an analytic field."""
    from nerf_navigation_b200 import raymarching as rm
    _, bits = scene
    N = 3000
    o, d = _rays(N, 5)
    aabb = np.array([-BOUND] * 3 + [BOUND] * 3, np.float32)
    nears, fars = oracle.near_far_from_aabb(o, d, aabb, 0.2)
    field_s = lambda p: (40.0 * (0.5 + 0.5 * np.sin(7 * p[:, 0] + 3 * p[:, 1]))).astype(np.float32)
    field_c = lambda p: np.ascontiguousarray((0.5 + 0.5 * np.cos(p * 5.0)).astype(np.float32))
    t = lambda a: torch.from_numpy(a).cuda()
    # oracle loop
    alive = np.arange(N, dtype=np.int32); rt = nears.copy()
    ws = np.zeros(N, np.float32); dep = np.zeros(N, np.float32); img = np.zeros((N, 3), np.float32)
    # device loop
    alive_d = torch.arange(N, dtype=torch.int32, device="cuda"); rt_d = t(nears).clone()
    ws_d = torch.zeros(N, device="cuda"); dep_d = torch.zeros(N, device="cuda"); img_d = torch.zeros(N, 3, device="cuda")
    step = 0
    while step < 1024 and alive.shape[0] > 0:
        n_alive = alive.shape[0]
        assert alive_d.shape[0] == n_alive
        n_step = max(min(N // n_alive, 8), 1)
        x, dd, dl = oracle.march_rays(n_alive, n_step, alive, rt, o, d, BOUND, bits, CAS, H, nears, fars)
        oracle.composite_rays(n_alive, n_step, alive, rt, field_s(x), field_c(x), dl, ws, dep, img)
        xd, ddd, dld = rm.march_rays(n_alive, n_step, alive_d, rt_d, t(o), t(d), BOUND, t(bits), CAS, H, t(nears), t(fars),
                                     -1, False, 0, 1024)
        assert np.array_equal(xd.cpu().numpy(), x) and np.array_equal(dld.cpu().numpy(), dl)
        rm.composite_rays(n_alive, n_step, alive_d, rt_d, t(field_s(x)), t(field_c(x)), dld, ws_d, dep_d, img_d)
        assert np.array_equal(alive_d.cpu().numpy(), alive)          # same rays die at the same step
        alive = alive[alive >= 0]; alive_d = rm.compact_alive(alive_d)
        step += n_step
    assert rel_err(ws_d.cpu().numpy(), ws) < 1e-5 and rel_err(img_d.cpu().numpy(), img) < 1e-5
    assert rel_err(dep_d.cpu().numpy(), dep) < 1e-5


def test_sh_vs_oracle(ngp, oracle):
    from nerf_navigation_b200.shencoder import SHEncoder
    _, d = _rays(10000, 6)
    enc = SHEncoder(3, 4)
    dt = torch.from_numpy(d).cuda().requires_grad_()
    out = enc(dt)
    ro, rj = oracle.sh_encode(d, 4, True)
    assert rel_err(out.detach().cpu().numpy(), ro) < 1e-6
    g = torch.randn(10000, 16, device="cuda", generator=torch.Generator(device="cuda").manual_seed(0))
    out.backward(g)
    ref = np.einsum("bc,bdc->bd", g.cpu().numpy().astype(np.float64), rj.astype(np.float64))
    assert rel_err(dt.grad.cpu().numpy(), ref) < 1e-5


def test_march_rays_train_pose_gradient(ngp):
    """EXTENSION: dL/drays_o, dL/drays_d through march_rays_train equals torch autograd through
    x_k = clamp(o + t_k d), dirs_k = d with the marcher's own (detached) abscissae t_k."""
    from nerf_navigation_b200 import raymarching as rm
    N = 3000
    grid = torch.from_numpy(stonehenge_grid()).cuda()
    bits = rm.packbits(grid, 10.0)
    o, d = orbit_rays(N, 5)
    o = torch.from_numpy(o).cuda().requires_grad_(True); d = torch.from_numpy(d).cuda().requires_grad_(True)
    aabb = torch.tensor([-2.0] * 3 + [2.0] * 3, device="cuda")
    nears, fars = rm.near_far_from_aabb(o.detach(), d.detach(), aabb, 0.2)
    xyzs, dirs, deltas, rays = rm.march_rays_train(o, d, 2.0, bits, 2, 128, nears, fars, None, -1, False, -1, True, 0, 1024)
    M = xyzs.shape[0]
    assert M > N and xyzs.requires_grad
    w1 = torch.randn(M, 3, device="cuda"); w2 = torch.randn(M, 3, device="cuda")
    ((xyzs * w1).sum() + (dirs * w2).sum()).backward()
    go, gd = o.grad.clone(), d.grad.clone()
    # torch restatement on the same abscissae
    r = rays.long()
    cnt, off = r[:, 2], r[:, 1]
    ray_of = torch.repeat_interleave(r[:, 0], cnt)
    assert ray_of.shape[0] <= M
    m = ray_of.shape[0]
    csum = torch.cumsum(deltas[:m, 1].double(), 0)
    start = torch.repeat_interleave(torch.cat([torch.zeros(1, device="cuda", dtype=torch.double), csum])[off], cnt)
    t = (nears.double()[ray_of] + (csum - start) - deltas[:m, 0].double()).float()
    o2 = o.detach().clone().requires_grad_(True); d2 = d.detach().clone().requires_grad_(True)
    x2 = torch.clamp(o2[ray_of] + t[:, None] * d2[ray_of], -2.0, 2.0)
    assert rel_err(x2.detach().cpu().numpy(), xyzs[:m].detach().cpu().numpy()) < 1e-6
    ((x2 * w1[:m]).sum() + (d2[ray_of] * w2[:m]).sum()).backward()
    assert rel_err(go.cpu().numpy(), o2.grad.cpu().numpy()) < 1e-5
    assert rel_err(gd.cpu().numpy(), d2.grad.cpu().numpy()) < 1e-5


def test_pose_rays_kernels_vs_torch():
    """closures.rays_from_poses on the GPU (one kernel + a per-hypothesis reduction backward) == the torch formulation that
    the get_rays golden pins, values and the gradient w.r.t. the poses; and the full measurement (12-state -> pose -> rays ->
    render -> mse) back-propagates to the states."""
    from nerf_navigation_b200 import closures
    g = torch.Generator().manual_seed(4)
    gt = closures.look_at_pose((1.2, 0.0, 0.4))
    poses = closures.perturb_poses(gt, 5, generator=g)
    pix = torch.randint(0, 640000, (777,), generator=g)
    intr = (1111.1, 1111.1, 400.0, 400.0)
    wo, wd = torch.randn(5, 777, 3, generator=g), torch.randn(5, 777, 3, generator=g)
    pc = poses.clone().requires_grad_(True)
    ro_c, rd_c = closures.rays_from_poses(pc, pix, intr, 800)                 # CPU tensors: the torch formulation
    ((ro_c * wo).sum() + (rd_c * wd).sum()).backward()
    pg = poses.cuda().requires_grad_(True)
    ro_g, rd_g = closures.rays_from_poses(pg, pix.cuda(), intr, 800)          # CUDA tensors: the kernels
    ((ro_g * wo.cuda()).sum() + (rd_g * wd.cuda()).sum()).backward()
    assert rel_err(ro_g.detach().cpu().numpy(), ro_c.detach().numpy()) == 0.0
    assert rel_err(rd_g.detach().cpu().numpy(), rd_c.detach().numpy()) < 1e-6
    assert rel_err(pg.grad.cpu().numpy(), pc.grad.numpy()) < 1e-5
    assert (pg.grad[:, 3] == 0).all()


def test_measurement_loss_backpropagates_to_states():
    """nav/estimator_helpers.py:293-327 for a batch of state hypotheses: the hypothesis that equals the observed state has
    (near) zero photometric loss, perturbed ones do not, and every hypothesis gets a finite gradient w.r.t. its position and
    rotation vector (none w.r.t. its velocities)."""
    from nerf_navigation_b200 import closures, synthetic
    from nerf_navigation_b200.nerf.network import NeRFNetwork
    torch.manual_seed(0)
    model = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10).cuda()
    model.encoder.embeddings.data.uniform_(-0.5, 0.5)
    synthetic.install_stonehenge(model)
    for p in model.parameters():
        p.requires_grad_(False)
    model.train()
    gt = closures.look_at_pose((1.2, 0.0, 0.4))
    st0 = closures.pose_to_state(gt)
    assert float((closures.state_to_pose(st0[None])[0] - gt).abs().max()) < 1e-5       # round trip
    g = torch.Generator().manual_seed(1)
    noise = torch.zeros(6, 12)
    noise[1:, 0:3] = torch.randn(5, 3, generator=g) * 0.02
    noise[1:, 6:9] = torch.randn(5, 3, generator=g) * 0.02
    states = (st0[None] + noise).cuda().requires_grad_(True)
    pix = torch.randint(0, 640000, (1024,), generator=g).cuda()
    intr = (1111.1, 1111.1, 400.0, 400.0)
    kw = dict(bg_color=1.0, perturb=False, dt_gamma=0, max_steps=1024, force_all_rays=True)
    with torch.no_grad():
        ro, rd = closures.rays_from_poses(gt[None].cuda(), pix, intr, 800)
        target = model.render(ro.reshape(1, -1, 3), rd.reshape(1, -1, 3), **kw)["image"].reshape(-1, 3)
    losses = closures.measurement_loss(model, states, pix, target, intr, 800, **kw)
    losses.sum().backward()
    assert losses.shape == (6,) and float(losses[0]) < 1e-8 and float(losses[1:].min()) > 1e-6
    gs = states.grad
    assert torch.isfinite(gs).all() and (gs[:, 3:6] == 0).all() and (gs[:, 9:] == 0).all()
    assert (gs[1:, :3].abs().sum(-1) > 0).all() and (gs[1:, 6:9].abs().sum(-1) > 0).all()
    # the process term adds the Mahalanobis distance and its gradient
    sig_inv = torch.eye(12, device="cuda") * 3.0
    s2 = states.detach().clone().requires_grad_(True)
    l2 = closures.measurement_loss(model, s2, pix, target, intr, 800, start_states=st0[None].cuda(), sig_inv=sig_inv, **kw)
    assert torch.allclose(l2.detach() - losses.detach(), 3.0 * (noise.cuda() ** 2).sum(-1), atol=1e-7)


def test_state_to_pose_kernels_vs_torch_formulation():
    """ngp_state_to_pose_forward / backward (one launch each) against the torch formulation of the same algebra in float64
    (the one tests/golden/pose_algebra.npz pins to nav/math_utils.py): poses 1e-6, dL/dstate 1e-5, incl. the zero rotation
    vector (identity, zero gradient), tiny and near-pi angles; velocity / angular-velocity gradients are exactly zero."""
    from nerf_navigation_b200 import closures as C
    g = torch.Generator().manual_seed(5)
    P = 257
    st = torch.randn(P, 12, generator=g)
    st[:, 6:9] *= torch.rand(P, 1, generator=g) * 3.0
    st[0, 6:9] = 0.0
    st[1, 6:9] = torch.tensor([1e-6, -2e-6, 5e-7])
    st[2, 6:9] = torch.tensor([3.1, 0.3, -0.2])
    G = torch.randn(P, 4, 4, generator=g)
    sd = st.double().requires_grad_(True)
    pose_ref = C.state_to_pose(sd)                       # CPU float64: the host formulation
    (pose_ref * G.double()).sum().backward()
    sc = st.cuda().requires_grad_(True)
    pose = C.state_to_pose(sc)                           # CUDA float32: the kernels
    (pose * G.cuda()).sum().backward()
    assert pose.shape == (P, 4, 4)
    assert float((pose.cpu().double() - pose_ref.detach()).abs().max()) < 1e-6
    gr, gk = sd.grad, sc.grad.cpu().double()
    assert float((gk - gr).abs().max() / gr.abs().max()) < 1e-5
    assert (gk[:, 3:6] == 0).all() and (gk[:, 9:12] == 0).all() and (gk[0, 6:9] == 0).all()
    rows = (gk[:, 6:9] - gr[:, 6:9]).abs().amax(-1) / gr[:, 6:9].abs().amax(-1).clamp_min(1e-12)
    assert float(rows[1:].max()) < 1e-4                  # per hypothesis, against its own scale
