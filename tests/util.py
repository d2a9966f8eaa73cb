"""Shared helpers for the parity tests (seeded synthetic inputs of the BASELINE.json shapes)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def hashgrid_config(bound=2.0, table_std=0.5, seed=0, level_dim=2, num_levels=16, log2_hashmap_size=19):
    """Config-2 encoder: 16 levels x 2 feats, 2^19 rows cap, base 16 -> 2048*bound (network.py:31, encoding.py:45-49)."""
    from oracle import oracle as O
    offsets, pls = O.level_table(3, num_levels, level_dim, 2.0, 16, log2_hashmap_size, 2048 * bound)
    rng = np.random.default_rng(seed)
    table = rng.uniform(-table_std, table_std, (int(offsets[-1]), level_dim)).astype(np.float32)
    return dict(offsets=offsets, per_level_scale=pls, S=float(np.log2(pls)), H=16, table=table, bound=bound,
                L=num_levels, C=level_dim)


def sigma_net_weights(seed=1, in_dim=32, hidden=64, out=16):
    """torch nn.Linear default init: U(-1/sqrt(in), 1/sqrt(in)), bias-free (network.py:33-47)."""
    rng = np.random.default_rng(seed)
    W0 = rng.uniform(-1, 1, (hidden, in_dim)).astype(np.float32) / np.sqrt(in_dim).astype(np.float32)
    W1 = rng.uniform(-1, 1, (out, hidden)).astype(np.float32) / np.sqrt(hidden).astype(np.float32)
    return W0, W1


def rel_err(a, b, floor=1e-30):
    """max |a-b| / max|b| — the relative error used for the fp32 (1e-5) and fp16 (2e-3) bars."""
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b)) / max(float(np.max(np.abs(b))), floor))


def rel_err_cols(a, b, axis=-1, floor=1e-30):
    """Worst per-column relative error: max over columns c of max|a[..., c] - b[..., c]| / max|b[..., c]|.  The global
    rel_err lets a channel 100x smaller than the largest one hide a wrong value; this normalises every column (geo_feat
    channel, dy_dx component, table-gradient level ...) by its own scale."""
    a = np.moveaxis(np.asarray(a, np.float64), axis, -1).reshape(-1, np.shape(a)[axis])
    b = np.moveaxis(np.asarray(b, np.float64), axis, -1).reshape(-1, np.shape(b)[axis])
    num = np.max(np.abs(a - b), axis=0)
    den = np.maximum(np.max(np.abs(b), axis=0), floor)
    return float(np.max(num / den))


def level_rel_err(a, b, offsets, floor=1e-30):
    """Worst per-level relative error of two [rows, C] tables (levels delimited by `offsets`)."""
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    worst = 0.0
    for l in range(len(offsets) - 1):
        lo, hi = int(offsets[l]), int(offsets[l + 1])
        den = max(float(np.max(np.abs(b[lo:hi]))), floor)
        worst = max(worst, float(np.max(np.abs(a[lo:hi] - b[lo:hi]))) / den)
    return worst


def load_ref(name):
    """Reference CUDA extension built by oracle/build_ref.py, or None."""
    from oracle import build_ref
    return build_ref.load_ref(name)


def stonehenge_grid(cascade=2, H=128, bound=2.0):
    """ring of 12 upright boxes (0.12 x 0.12 x 0.5) at radius 0.7 + floor slab z < 0.02,
    density 20 inside / 0 outside, on the [cascade, H^3] grid in Morton order (renderer.py:464-483 cell centres)."""
    from oracle import oracle as O
    ar = np.arange(H, dtype=np.int32)
    xx, yy, zz = np.meshgrid(ar, ar, ar, indexing="ij")
    coords = np.stack([xx.ravel(), yy.ravel(), zz.ravel()], -1).astype(np.int32)
    idx = O.morton3D(coords).astype(np.int64)
    unit = 2.0 * coords.astype(np.float32) / (H - 1) - 1.0
    grid = np.zeros((cascade, H ** 3), np.float32)
    ang = np.arange(12) * (2 * np.pi / 12)
    centres = np.stack([0.7 * np.cos(ang), 0.7 * np.sin(ang)], -1)
    for cas in range(cascade):
        b = min(2 ** cas, bound)
        p = unit * (b - b / H)
        occ = (p[:, 2] < 0.02) & (p[:, 2] > -0.1) & (np.abs(p[:, 0]) < 1.0) & (np.abs(p[:, 1]) < 1.0)
        for c in centres:
            occ |= (np.abs(p[:, 0] - c[0]) < 0.06) & (np.abs(p[:, 1] - c[1]) < 0.06) & (p[:, 2] >= 0.0) & (p[:, 2] < 0.5)
        grid[cas, idx] = np.where(occ, 20.0, 0.0)
    return grid


def orbit_rays(n_rays, seed=0, radius=1.2, height=0.4, H=800, W=800, focal=1111.1, jitter=0.0):
    """Pinhole rays (nerf/utils.py:52-116 conventions: x right, y down, z forward) from a camera on an orbit looking
    at the origin; `n_rays` random pixels."""
    rng = np.random.default_rng(seed)
    ang = rng.uniform(0, 2 * np.pi)
    eye = np.array([radius * np.cos(ang), radius * np.sin(ang), height], np.float64) + rng.normal(0, jitter, 3)
    fwd = -eye / np.linalg.norm(eye)
    right = np.cross(fwd, np.array([0.0, 0.0, 1.0])); right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    R = np.stack([right, down, fwd], -1)
    pix = rng.integers(0, H * W, n_rays)
    i = (pix % W).astype(np.float64) + 0.5
    j = (pix // W).astype(np.float64) + 0.5
    d_cam = np.stack([(i - W / 2) / focal, (j - H / 2) / focal, np.ones_like(i)], -1)
    d = d_cam @ R.T
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    o = np.broadcast_to(eye, d.shape)
    return np.ascontiguousarray(o, np.float32), np.ascontiguousarray(d, np.float32)
