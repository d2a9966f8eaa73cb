"""GPU: occupancy-grid upkeep (csrc/occgrid.cu behind NeRFRenderer.update_extra_state / mark_untrained_grid) and the
planner's A* occupancy (closures.astar_occupancy) against the numpy restatement of the reference's host logic
(oracle/occupancy.py: nerf/renderer.py:381-539, nav/quad_plot.py:64-83).  Bitfields and index outputs bit-exact."""
import copy

import numpy as np
import pytest
import torch

from util import stonehenge_grid, rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def model():
    from nerf_navigation_b200.nerf.network import NeRFNetwork
    torch.manual_seed(0)
    m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10).cuda()
    m.encoder.embeddings.data.uniform_(-0.5, 0.5)
    return m.eval()


def _density_np(model, xyz):
    with torch.no_grad():
        return model.density(torch.from_numpy(np.ascontiguousarray(xyz)).cuda())["sigma"].reshape(-1).cpu().numpy()


def test_full_sweep_vs_oracle(model, ngp):
    """Two full refreshes without jitter (the reference's torch.rand_like noise cannot be reproduced; jitter = 0 in both):
    cell positions, EMA-max incl. the decay branch, untrained (-1) cells, mean density, threshold and packed bits."""
    from oracle import occupancy as OC
    m = copy.deepcopy(model)
    m.reset_extra_state()
    m.density_grid[0, 1000:1100] = -1.0
    grid = m.density_grid.cpu().numpy().copy()
    G, C = m.grid_size, m.cascade
    for it in range(2):
        m._update_extra_state_sweep(0.95, jitter=0.0)
        tmp = np.stack([_density_np(model, OC.cell_centers(cas, G, m.bound)[0]) * np.float32(m.density_scale) for cas in range(C)])
        grid, mean = OC.ema_update(grid, tmp, 0.95)
        assert np.array_equal(m.density_grid.cpu().numpy(), grid)
        assert abs(m.mean_density - mean) <= 1e-6 * abs(mean)
        bits = OC.bitfield(grid, mean, m.density_thresh)
        assert np.array_equal(m.density_bitfield.cpu().numpy(), bits)
    assert (m.density_grid[0, 1000:1100] == -1).all() and m.iter_density == 2
    # cell-centre positions bit-exact; jittered positions stay inside their cell, unbiased
    for cas in range(C):
        ctr = torch.empty(G ** 3, 3, device="cuda")
        ngp.check(ngp.lib().ngp_occ_grid_positions(cas, G, float(m.bound), 0.0, 7, ctr.data_ptr(), ngp.stream()))
        want, half = OC.cell_centers(cas, G, m.bound)
        assert np.array_equal(ctr.cpu().numpy(), want)
        jit = torch.empty_like(ctr)
        ngp.check(ngp.lib().ngp_occ_grid_positions(cas, G, float(m.bound), 1.0, 7, jit.data_ptr(), ngp.stream()))
        dev = (jit - ctr)
        assert float(dev.abs().max()) <= half and float(dev.abs().mean()) > 0.3 * half and abs(float(dev.mean())) < 1e-2 * half


def test_partial_refresh_vs_oracle(model, ngp):
    """The iter_density >= 16 branch (renderer.py:488-517): the drawn cells are G^3/4 uniform cells + G^3/4 draws from the
    occupied set; given the draw (seeded, reproducible) the grid / mean / bitfield follow the oracle bit for bit."""
    from oracle import occupancy as OC
    m = copy.deepcopy(model)
    m.reset_extra_state()
    m.density_grid.copy_(torch.from_numpy(stonehenge_grid()))
    m.density_grid[1, 5000:5200] = -1.0
    m.iter_density = 16
    grid0 = m.density_grid.cpu().numpy().copy()
    G, C, cells = m.grid_size, m.cascade, m.grid_size ** 3
    N = cells // 4
    sums = torch.empty(C, ngp.lib().ngp_occ_grid_update_blocks(cells), device="cuda")
    m2 = copy.deepcopy(m)
    with torch.no_grad():
        picked = m._refresh_partial(0.95, sums, jitter=0.0).cpu().numpy()
    assert picked.shape == (C, 2 * N)
    grid = grid0.copy()
    for cas in range(C):
        idx = picked[cas]
        assert idx.min() >= 0 and idx.max() < cells
        occ = np.flatnonzero(grid0[cas] > 0)
        assert np.isin(idx[N:], occ).all()                                  # second half: occupied cells only ...
        assert len(np.unique(idx[N:])) > 0.2 * min(N, len(occ))             # ... spread over the occupied set
        cnt = np.bincount(idx[:N] >> 18, minlength=8)                       # first half: uniform over the grid (octants)
        assert cnt.min() > 0.9 * N / 8 and cnt.max() < 1.1 * N / 8
        coords = OC.cell_coords(G)[idx]
        xyz, _ = OC.cell_centers(cas, G, m.bound, coords)
        tmp = OC.scatter_queries(cells, idx, _density_np(model, xyz), m.density_scale)
        grid[cas], _ = OC.ema_update(grid[cas], tmp, 0.95)
    assert np.array_equal(m.density_grid.cpu().numpy(), grid)
    mean = float(np.mean(np.clip(grid, 0, None).astype(np.float64)))
    assert abs(float(sums.double().sum().item() / (C * cells)) - mean) <= 1e-6 * mean
    # the public entry point takes the same branch, advances the counters and packs the bits; same seed -> same draw
    m2.update_extra_state()
    assert m2.iter_density == 17 and abs(m2.mean_density - mean) < 0.05 * mean
    assert np.array_equal(m2.density_bitfield.cpu().numpy(), OC.bitfield(m2.density_grid.cpu().numpy(), m2.mean_density, m2.density_thresh))
    assert (m2.density_grid[1, 5000:5200] == -1).all()
    m3 = copy.deepcopy(m); m3.density_grid.copy_(torch.from_numpy(grid0)); m3.iter_density = 16
    m4 = copy.deepcopy(m3)
    m3.update_extra_state(); m4.update_extra_state()
    assert torch.equal(m3.density_grid, m4.density_grid)                    # reproducible run to run
    # nothing occupied: the random half still refreshes, the occupied half issues no query
    m5 = copy.deepcopy(model); m5.reset_extra_state(); m5.iter_density = 16
    m5.update_extra_state()
    assert m5.mean_density > 0


def test_mark_untrained_grid_vs_oracle(model):
    from oracle import occupancy as OC
    from nerf_navigation_b200 import closures
    m = copy.deepcopy(model)
    m.reset_extra_state()
    m.density_grid.fill_(0.5)
    poses = torch.stack([closures.look_at_pose((1.2 * np.cos(a), 1.2 * np.sin(a), 0.4)) for a in np.linspace(0, 1.5, 7)])
    intr = (1111.1, 1111.1, 400.0, 400.0)
    want, edge = OC.mark_untrained(m.density_grid.cpu().numpy(), poses.numpy(), intr, m.grid_size, m.bound)
    m.mark_untrained_grid(poses.numpy(), intr)          # ndarray poses as Trainer passes them (nerf/utils.py:563)
    got = m.density_grid.cpu().numpy()
    diff = got != want
    assert not (diff & ~edge).any(), int(diff.sum())    # only cells within rounding of a frustum plane may differ
    assert diff.sum() <= 1e-4 * got.size
    assert (got == -1).mean() > 0.2 and (got == 0.5).mean() > 0.2
    # marked cells are never refreshed and never occupied
    m.update_extra_state()
    after = m.density_grid.cpu().numpy()
    assert (after[got == -1] == -1).all()
    bits = np.unpackbits(m.density_bitfield.cpu().numpy(), bitorder="little").reshape(got.shape)
    assert not bits[got == -1].any()


def test_astar_occupancy_vs_oracle(model):
    """quad_plot.py:64-83 — lattice positions bit-exact with torch.linspace / meshgrid, pooled grid vs the oracle."""
    from oracle import occupancy as OC
    from nerf_navigation_b200.closures import make_density_fn, astar_occupancy
    fn = make_density_fn(model)
    seen = {}

    def spy(x):
        seen["x"] = x.detach().clone()
        seen["s"] = fn(x)
        return seen["s"]

    side, k, th = 40, 5, 1.0
    occ = astar_occupancy(spy, side=side, kernel_size=k, thresh=th)
    assert occ.shape == (8, 8, 8) and occ.dtype == torch.bool
    lin = torch.linspace(-1, 1, side)
    want = torch.stack(torch.meshgrid(lin, lin, lin, indexing="ij"), -1)
    assert torch.equal(seen["x"].cpu(), want)
    assert np.array_equal(seen["x"].cpu().numpy().reshape(-1, 3), OC.astar_lattice((-1, -1, -1), (1, 1, 1), side))
    ref = OC.astar_occupied(seen["s"].cpu().numpy(), side, k, th)
    assert np.array_equal(occ.cpu().numpy(), ref)
    assert torch.equal(occ, torch.nn.functional.max_pool3d(seen["s"][None, None], k)[0, 0] > th)
    # rectangular extents (the CHURCH branch, quad_plot.py:67-72)
    occ2 = astar_occupancy(spy, side=20, kernel_size=5, thresh=th, lo=(-2, -1, 0), hi=(-1, 0, 1))
    ax = [torch.linspace(a, b, 20) for a, b in ((-2, -1), (-1, 0), (0, 1))]
    assert torch.equal(seen["x"].cpu(), torch.stack(torch.meshgrid(*ax, indexing="ij"), -1)) and occ2.shape == (4, 4, 4)
