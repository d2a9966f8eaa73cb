"""Synthetic scenes of the BASELINE.json configurations: analytic "stonehenge" occupancy — a ring of 12
upright boxes (0.12 x 0.12 x 0.5) at radius 0.7 plus a floor slab, density 20 inside / 0 outside, on the [cascade, 128^3]
Morton-ordered grid the renderer marches (renderer.py:464-483 cell centres).  Same arithmetic as tests/util.py."""
import math

import torch

from . import raymarching


def stonehenge_density_grid(cascade=2, H=128, bound=2.0, device="cuda"):
    ar = torch.arange(H, dtype=torch.int32, device=device)
    xx, yy, zz = torch.meshgrid(ar, ar, ar, indexing="ij")
    coords = torch.stack([xx.reshape(-1), yy.reshape(-1), zz.reshape(-1)], -1).contiguous()
    idx = raymarching.morton3D(coords).long()
    unit = 2.0 * coords.float() / (H - 1) - 1.0
    grid = torch.zeros(cascade, H ** 3, device=device)
    ang = torch.arange(12, device=device) * (2 * math.pi / 12)
    centres = torch.stack([0.7 * torch.cos(ang), 0.7 * torch.sin(ang)], -1)
    for cas in range(cascade):
        b = min(2 ** cas, bound)
        p = unit * (b - b / H)
        occ = (p[:, 2] < 0.02) & (p[:, 2] > -0.1) & (p[:, 0].abs() < 1.0) & (p[:, 1].abs() < 1.0)
        for c in centres:
            occ |= ((p[:, 0] - c[0]).abs() < 0.06) & ((p[:, 1] - c[1]).abs() < 0.06) & (p[:, 2] >= 0.0) & (p[:, 2] < 0.5)
        grid[cas, idx] = torch.where(occ, 20.0, 0.0)
    return grid


def install_stonehenge(model, thresh=10.0):
    """Fill model.density_grid / density_bitfield (cuda_ray=True) with the analytic scene."""
    g = stonehenge_density_grid(model.cascade, model.grid_size, float(model.bound), model.density_grid.device)
    model.density_grid.copy_(g)
    model.density_bitfield = raymarching.packbits(model.density_grid, thresh, model.density_bitfield)
    model.mean_density = float(g.clamp(min=0).mean())
    return model
