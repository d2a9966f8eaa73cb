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


# ------------------------------------------------------------------ float64 replica of NeRFRenderer.run (torch, autograd)
def sh4_torch(d):
    """Real spherical harmonics, degree 4 (the polynomials of shencoder.cu:51-69 / orc_sh_encode), any dtype, differentiable."""
    import torch
    x, y, z = d[..., 0], d[..., 1], d[..., 2]
    xy, xz, yz, x2, y2, z2 = x * y, x * z, y * z, x * x, y * y, z * z
    c1, c2 = 0.48860251190291987, 1.0925484305920792
    v = [torch.full_like(x, 0.28209479177387814), -c1 * y, c1 * z, -c1 * x, c2 * xy, -c2 * yz,
         0.94617469575755997 * z2 - 0.31539156525251999, -c2 * xz, 0.54627421529603959 * x2 - 0.54627421529603959 * y2,
         0.59004358992664352 * y * (-3.0 * x2 + y2), 2.8906114426405538 * xy * z, 0.45704579946446572 * y * (1.0 - 5.0 * z2),
         0.3731763325901154 * z * (5.0 * z2 - 3.0), 0.45704579946446572 * x * (1.0 - 5.0 * z2), 1.4453057213202769 * z * (x2 - y2),
         0.59004358992664352 * x * (-x2 + 3.0 * y2)]
    return torch.stack(v, -1)


def hashgrid_encode_f64(x01, table, offsets, scales, x01_f32=None):
    """Hash-grid forward (gridencoder.cu:75-178, D = 3, hash type) in float64 torch: the CELL of every point is the one the
    fp32 kernels pick — floor(fl32(fma(x01_f32, scale, 0.5))), gridencoder.cu:134-137, with x01_f32 the kernels' own fp32
    normalised position — because the features are continuous across a cell face but their gradient is not: which side of
    a face a point within rounding of it falls on is an fp32 tie like a ReLU at zero.  The fractional position is the fp32
    one as well (see below); everything after it is exact.  Differentiable w.r.t. x01."""
    import torch
    L = len(offsets) - 1
    primes = (1, 2654435761, 805459861)
    outs = []
    for l in range(L):
        rows = int(offsets[l + 1] - offsets[l])
        scale = float(scales[l])
        step = int(np.ceil(scale)) + 2
        pos = x01 * scale + 0.5
        x32 = x01.detach().float() if x01_f32 is None else x01_f32
        pos32 = (x32.double() * scale + 0.5).float().double()                 # one rounding, like the kernels' fma
        pg = torch.floor(pos32)
        # the interpolation weights are taken at the FP32 position (ulp 5e-4 cells at the finest level: part of what the
        # reference computes, identical in every fp32 implementation), their derivative d(pos)/d(x01) = scale is exact
        f = (pos32 - pg) + (pos - pos.detach())
        pg = pg.long()
        dense = step ** 3 <= rows
        acc = 0
        for k in range(8):
            w, idx, stride = 1.0, 0, 1
            for dd in range(3):
                bit = (k >> dd) & 1
                c = pg[:, dd] + bit
                w = w * (f[:, dd] if bit else 1 - f[:, dd])
                if dense:
                    idx = idx + c * stride
                    stride *= step
                else:
                    idx = (idx ^ ((c * primes[dd]) & 0xFFFFFFFF)) if dd else (c & 0xFFFFFFFF)
            acc = acc + w[:, None] * table[(idx % rows) + int(offsets[l])]
        outs.append(acc)
    return torch.cat(outs, -1)


def run_f64(model, rays_o, rays_d, near, far, scales, num_steps=512, bg=1.0):
    """NeRFRenderer.run (renderer.py:125-254, upsample_steps = 0, perturb = False) of `model` (hash-grid NeRFNetwork) in
    float64: same sample depths (from the fp32 near / far), exact arithmetic everywhere else.  rays_* double leaf tensors."""
    import torch
    N = rays_o.shape[0]
    # sample depths: the fp32 values the renderer itself uses (its own op order), then exact arithmetic on them
    z32 = near[:, None] + (far - near)[:, None] * torch.linspace(0.0, 1.0, num_steps, device=rays_o.device)[None, :].expand(N, num_steps)
    near, far = near.double()[:, None], far.double()[:, None]
    z = z32.double()
    gap = ((far.float() - near.float()) / num_steps).double()
    box = model.aabb_infer.double()
    p = rays_o[:, None, :] + rays_d[:, None, :] * z[..., None]
    # The clip to the box (renderer.py:160) is decided by the FP32 evaluation (the last sample of a ray sits on the box face:
    # whether it is clipped — zero gradient on that axis — is a rounding matter): same decision here, exact arithmetic after
    # it.  torch.max / torch.min split the gradient equally at exact ties.
    p32 = rays_o.detach().float()[:, None, :] + rays_d.detach().float()[:, None, :] * z32[..., None]
    b32 = model.aabb_infer
    g = ((p32 > b32[:3]) & (p32 < b32[3:])).double() + 0.5 * ((p32 == b32[:3]) | (p32 == b32[3:])).double()
    p = torch.min(torch.max(p32, b32[:3]), b32[3:]).double() + g * (p - p.detach())
    p = p.reshape(-1, 3)
    enc = model.encoder
    x01_f32 = (p.detach().float() + float(model.bound)) / (2 * float(model.bound))      # grid.py:144 in fp32, as the kernels do
    feat = hashgrid_encode_f64((p + model.bound) / (2 * model.bound), enc.embeddings.detach().double(),
                               enc.offsets.cpu().numpy(), scales, x01_f32)
    h = torch.relu(feat @ model.sigma_net[0].weight.detach().double().t()) @ model.sigma_net[1].weight.detach().double().t()
    sigma = torch.exp(h[:, 0]).view(N, num_steps)
    gaps = torch.cat([z[:, 1:] - z[:, :-1], gap.expand(N, 1)], -1)
    alpha = 1 - torch.exp(-gaps * model.density_scale * sigma)
    w = alpha * torch.cat([torch.ones_like(alpha[:, :1]), 1 - alpha + 1e-15], -1).cumprod(-1)[:, :-1]
    d = rays_d[:, None, :].expand(N, num_steps, 3).reshape(-1, 3)
    c = torch.cat([sh4_torch(d), h[:, 1:]], -1)
    for i, layer in enumerate(model.color_net):
        c = c @ layer.weight.detach().double().t()
        if i < 2:
            c = torch.relu(c)
    rgb = torch.sigmoid(c).view(N, num_steps, 3) * (w > 1e-4)[..., None]
    return (w[..., None] * rgb).sum(-2) + (1 - w.sum(-1))[:, None] * bg
