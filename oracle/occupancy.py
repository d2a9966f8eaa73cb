"""TEST INFRASTRUCTURE ONLY — numpy restatement of the reference's occupancy-grid host logic, for checking the occupancy
kernels (csrc/occgrid.cu) and their host driver (nerf_navigation_b200/nerf/renderer.py).  The product never imports this.

Restated from /root/reference:
  nerf/renderer.py:381-444   mark_untrained_grid   (camera-frustum test per cell)
  nerf/renderer.py:446-539   update_extra_state    (cell positions, EMA-max, mean density, threshold, packbits;
                                                    full sweep :457-486 and partial draw :488-517)
  nav/quad_plot.py:64-83     a_star_init           (linspace lattice, MaxPool3d, threshold)
The reference draws its jitter / cell choices from torch's RNG, which nothing else can reproduce: the functions here take the
drawn cells (and jitter-free or given positions) as inputs and restate everything that is deterministic given them.
parity: unpinned by reference tests (it ships none); pinned through the kernels' other checks and by construction below.
"""
import numpy as np

from . import oracle as O


def cell_coords(G):
    """Integer cell coordinates of every Morton index 0..G^3-1 ([G^3, 3] int32): the inverse of raymarching.morton3D."""
    return O.morton3D_invert(np.arange(G ** 3, dtype=np.int32))


def cell_centers(cas, G, bound, coords=None):
    """renderer.py:470-479 without the noise: xyzs = 2 c / (G - 1) - 1, scaled to the cascade: xyzs * (b - b / G)."""
    c = cell_coords(G) if coords is None else coords
    # torch divides a float tensor by a python int as a multiplication by the float32 reciprocal
    xyz = (np.float32(2.0) * c.astype(np.float32)) * np.float32(1.0 / np.float32(G - 1)) - np.float32(1.0)
    b = np.float32(min(2 ** cas, bound))
    half = b / np.float32(G)
    return xyz * (b - half), half


def ema_update(grid, tmp, decay):
    """renderer.py:523-526: where both are >= 0, grid = max(grid * decay, tmp); returns the new grid and its mean density
    (clamped at 0, all cascades)."""
    grid = grid.astype(np.float32).copy()
    valid = (grid >= 0) & (tmp >= 0)
    grid[valid] = np.maximum(grid[valid] * np.float32(decay), tmp[valid])
    return grid, float(np.mean(np.clip(grid, 0, None).astype(np.float64)))


def scatter_queries(n_cells, cell_idx, sigma, density_scale):
    """renderer.py:513-516 `tmp_grid[cas, indices] = sigmas * density_scale` on a grid of -1.  With duplicate indices torch's
    index_put keeps an unspecified one of the values; the kernels keep the largest (one of them, and deterministic), which
    is what this restates."""
    tmp = np.full(n_cells, -1.0, np.float32)
    ok = cell_idx >= 0
    s = (sigma.astype(np.float32) * np.float32(density_scale))[ok]
    np.maximum.at(tmp, cell_idx[ok], np.where(s >= 0, s, np.float32(-1.0)))
    return tmp


def bitfield(grid, mean_density, density_thresh):
    """renderer.py:529-531: threshold = min(mean, density_thresh); packbits (raymarching.cu:269-291)."""
    return O.packbits(grid, min(mean_density, density_thresh))


def mark_untrained(grid, poses, intrinsic, G, bound):
    """renderer.py:381-444 in float32: a cell is kept iff at least one camera has it in front (z > 0) and inside the
    frustum widened by one cell: |x| < cx / fx * z + 2 half, |y| < cy / fy * z + 2 half (cam = (p - t) @ R)."""
    fx, fy, cx, cy = intrinsic
    sx, sy = np.float32(cx / fx), np.float32(cy / fy)
    grid = grid.astype(np.float32).copy()
    poses = np.asarray(poses, np.float32)
    coords = cell_coords(G)
    edge = np.zeros(grid.shape, bool)     # cells within rounding of a frustum plane (the bmm's summation order is cuBLAS')
    for cas in range(grid.shape[0]):
        p, half = cell_centers(cas, G, bound, coords)
        h2 = np.float32(half * np.float32(2.0))
        seen = np.zeros(G ** 3, bool)
        near = np.zeros(G ** 3, bool)
        for P in poses:
            d = p - P[:3, 3]
            cam = d @ P[:3, :3]
            z = cam[:, 2]
            mx = np.abs(cam[:, 0]) - (sx * z + h2)
            my = np.abs(cam[:, 1]) - (sy * z + h2)
            seen |= (z > 0) & (mx < 0) & (my < 0)
            tol = 1e-5 * (1.0 + np.abs(cam).max(-1))
            near |= (np.abs(z) < tol) | (np.abs(mx) < tol) | (np.abs(my) < tol)
        grid[cas, ~seen] = -1.0
        edge[cas] = near
    return grid, edge


def linspace(lo, hi, n):
    """torch.linspace in float32: start + i step for the first half, end - (n - 1 - i) step for the second, each a fused
    multiply-add on the GPU (one rounding: evaluated here in float64, where the product and sum are exact, then rounded)."""
    lo, hi = np.float32(lo), np.float32(hi)
    step = np.float64(np.float32((hi - lo) / np.float32(n - 1)))
    i = np.arange(n).astype(np.float64)
    return np.where(i < n // 2, np.float64(lo) + step * i, np.float64(hi) - step * (n - 1 - i)).astype(np.float32)


def astar_lattice(lo, hi, side):
    """quad_plot.py:75-77: coods = stack(meshgrid(linspace, linspace, linspace), -1) ('ij') -> [side^3, 3]."""
    ax = [linspace(l, h, side) for l, h in zip(lo, hi)]
    return np.stack(np.meshgrid(*ax, indexing="ij"), -1).reshape(-1, 3).astype(np.float32)


def astar_occupied(sigma, side, k, thresh):
    """quad_plot.py:80-85: MaxPool3d(kernel_size = k)(sigma [side, side, side]) > thresh -> bool [(side // k)^3]."""
    g = side // k
    v = sigma.reshape(side, side, side)[:g * k, :g * k, :g * k].reshape(g, k, g, k, g, k)
    return v.max(axis=(1, 3, 5)) > thresh
