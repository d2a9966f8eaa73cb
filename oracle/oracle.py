"""TEST INFRASTRUCTURE ONLY — numpy/ctypes front end of the CPU oracle (oracle/ngp_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
Every function names the reference file:line it restates (details in the C source).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libngp_oracle.so")


def build(force=False):
    src = os.path.join(_HERE, "ngp_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "libngp_oracle.so"], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.orc_pcg32_jitter.restype = C.c_float
        _lib.orc_pcg32_jitter.argtypes = [C.c_uint64, C.c_uint64]
        _lib.orc_num_threads.restype = C.c_int
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def num_threads():
    return lib().orc_num_threads()


def set_num_threads(n):
    lib().orc_set_num_threads(C.c_int(n))


# ---------------------------------------------------------------- level table (host logic)
def level_table(input_dim=3, num_levels=16, level_dim=2, per_level_scale=2.0, base_resolution=16,
                log2_hashmap_size=19, desired_resolution=None, align_corners=False):
    """gridencoder/grid.py:98-124: offsets[L+1] (int32), per_level_scale (float64)."""
    if desired_resolution is not None:
        per_level_scale = np.exp2(np.log2(desired_resolution / base_resolution) / (num_levels - 1))
    offsets, off = [], 0
    cap = 2 ** log2_hashmap_size
    for i in range(num_levels):
        res = int(np.ceil(base_resolution * per_level_scale ** i))
        n = min(cap, (res if align_corners else res + 1) ** input_dim)
        n = int(np.ceil(n / 8) * 8)
        offsets.append(off)
        off += n
    offsets.append(off)
    return np.array(offsets, dtype=np.int32), float(per_level_scale)


def level_scales(L, per_level_scale, H):
    """gridencoder.cu:126 with S = np.log2(per_level_scale) passed as C float (grid.py:33)."""
    s = np.empty(L, np.float32)
    lib().orc_grid_level_scales(C.c_int(L), C.c_float(np.log2(per_level_scale)), C.c_uint32(H), _p(s))
    return s


# ---------------------------------------------------------------- grid encoder
def grid_encode_forward(x01, table, offsets, scales, calc_grad_inputs=False, gridtype=0, align_corners=False,
                        want_indices=False):
    """gridencoder.cu:75-224.  Returns out [L,B,C], dy_dx [B,L,D,C] | None, indices [B,L,2^D] | None."""
    x01 = _f32(x01); table = _f32(table)
    B, D = x01.shape; Cc = table.shape[1]; L = len(offsets) - 1
    out = np.empty((L, B, Cc), np.float32)
    dy = np.empty((B, L, D, Cc), np.float32) if calc_grad_inputs else None
    ix = np.empty((B, L, 1 << D), np.uint32) if want_indices else None
    lib().orc_grid_encode_forward(_p(x01), _p(table), _p(offsets), _p(scales), C.c_int64(B), C.c_int(D), C.c_int(Cc),
                                  C.c_int(L), C.c_uint32(gridtype), C.c_int(int(align_corners)), _p(out), _p(dy), _p(ix))
    return out, dy, ix


def grid_backward_table(grad_lbc, x01, offsets, scales, C_, gridtype=0, align_corners=False):
    """gridencoder.cu:227-314 (float64 accumulation).  grad [L,B,C] -> grad_table [rows,C]."""
    grad_lbc = _f32(grad_lbc); x01 = _f32(x01)
    L, B, Cc = grad_lbc.shape; D = x01.shape[1]
    gt = np.empty((int(offsets[-1]), Cc), np.float32)
    lib().orc_grid_backward_table(_p(grad_lbc), _p(x01), _p(offsets), _p(scales), C.c_int64(B), C.c_int(D),
                                  C.c_int(Cc), C.c_int(L), C.c_uint32(gridtype), C.c_int(int(align_corners)), _p(gt))
    return gt


def grid_backward_input(grad_lbc, dy_dx):
    """gridencoder.cu:317-343."""
    grad_lbc = _f32(grad_lbc); dy_dx = _f32(dy_dx)
    L, B, Cc = grad_lbc.shape; D = dy_dx.shape[2]
    gx = np.empty((B, D), np.float32)
    lib().orc_grid_backward_input(_p(grad_lbc), _p(dy_dx), C.c_int64(B), C.c_int(D), C.c_int(Cc), C.c_int(L), _p(gx))
    return gx


def field_query(x, bound, table, offsets, scales, W0, W1, grad_sigma=None, grad_geo=None, gridtype=0,
                align_corners=False):
    """network.py:126-143 density() (+ backward to x).  Returns sigma, geo, grad_x|None."""
    x = _f32(x); table = _f32(table); W0 = _f32(W0); W1 = _f32(W1)
    B = x.shape[0]; Cc = table.shape[1]; L = len(offsets) - 1
    hidden, n_out = W0.shape[0], W1.shape[0]
    sigma = np.empty(B, np.float32); geo = np.empty((B, n_out - 1), np.float32)
    gs = None if grad_sigma is None else _f32(grad_sigma)
    gg = None if grad_geo is None else _f32(grad_geo)
    gx = None if gs is None else np.empty((B, 3), np.float32)
    lib().orc_field_query(_p(x), C.c_float(bound), _p(table), _p(offsets), _p(scales), C.c_int64(B), C.c_int(Cc),
                          C.c_int(L), C.c_uint32(gridtype), C.c_int(int(align_corners)), _p(W0), _p(W1),
                          C.c_int(hidden), C.c_int(n_out), _p(sigma), _p(geo), _p(gs), _p(gg), _p(gx))
    return sigma, geo, gx


def mlp_forward(x, weights, dims, out_act=0):
    """nn.Linear stack, ReLU hidden, bias-free (network.py:104-122).  weights: list of [out,in]."""
    x = _f32(x)
    flat = np.concatenate([_f32(w).ravel() for w in weights])
    dims_a = np.array(dims, np.int32)
    y = np.empty((x.shape[0], dims[-1]), np.float32)
    lib().orc_mlp_forward(_p(x), _p(flat), _p(dims_a), C.c_int(len(dims) - 1), C.c_int(out_act), C.c_int64(x.shape[0]), _p(y))
    return y


def sh_encode(dirs, degree=4, calc_grad_inputs=False):
    """shencoder.cu:44-70 (+ Jacobian).  Returns out [B,deg^2], dy_dx [B,3,deg^2] | None."""
    dirs = _f32(dirs); B = dirs.shape[0]
    out = np.empty((B, degree * degree), np.float32)
    dy = np.empty((B, 3, degree * degree), np.float32) if calc_grad_inputs else None
    lib().orc_sh_encode(_p(dirs), C.c_int64(B), C.c_int(degree), _p(out), _p(dy))
    return out, dy


def field_color(x, d, bound, table, offsets, scales, Ws, Wc, grad_sigma=None, grad_rgb=None):
    """NeRFNetwork.forward(x, d) (network.py:95-124) and its backward to (x, d): density() restated by orc_field_query,
    SH(4) by orc_sh_encode (shencoder.cu:44-70 and its Jacobian :131-298), then the colour net 31-64-64-3 + sigmoid
    (network.py:113-122) and the chain rule through it in float64 numpy.  Ws = [W0, W1] sigma net, Wc = [W0, W1, W2]
    colour net, nn.Linear layouts.  Returns dict(sigma, rgb, z=[pre-activations of the two hidden colour layers],
    grad_x, grad_d) (gradients only when grad_sigma / grad_rgb are given)."""
    x = _f32(x); d = _f32(d)
    B = x.shape[0]
    want_grad = grad_sigma is not None or grad_rgb is not None
    sigma, geo, _ = field_query(x, bound, table, offsets, scales, Ws[0], Ws[1], None)
    sh, dsh = sh_encode(d, 4, calc_grad_inputs=want_grad)
    h0 = np.concatenate([sh, geo], -1).astype(np.float64)
    W0, W1, W2 = (np.asarray(w, np.float64) for w in Wc)
    z1 = h0 @ W0.T; h1 = np.maximum(z1, 0)
    z2 = h1 @ W1.T; h2 = np.maximum(z2, 0)
    z3 = h2 @ W2.T
    rgb = 1.0 / (1.0 + np.exp(-z3))
    out = dict(sigma=sigma, rgb=rgb, z=[z1, z2])
    if want_grad:
        gs = np.zeros(B, np.float32) if grad_sigma is None else _f32(grad_sigma)
        gr = np.zeros((B, 3), np.float64) if grad_rgb is None else np.asarray(grad_rgb, np.float64)
        dz3 = gr * rgb * (1.0 - rgb)
        dz2 = (dz3 @ W2) * (z2 > 0)
        dz1 = (dz2 @ W1) * (z1 > 0)
        dh0 = dz1 @ W0
        out["grad_d"] = np.einsum("bk,bak->ba", dh0[:, :16], dsh.astype(np.float64))
        _, _, gx = field_query(x, bound, table, offsets, scales, Ws[0], Ws[1], gs, dh0[:, 16:].astype(np.float32))
        out["grad_x"] = gx
    return out


# ---------------------------------------------------------------- ray ops
def near_far_from_aabb(rays_o, rays_d, aabb, min_near=0.2):
    """raymarching.cu:93-147."""
    rays_o = _f32(rays_o); rays_d = _f32(rays_d); aabb = _f32(aabb)
    N = rays_o.shape[0]
    nears = np.empty(N, np.float32); fars = np.empty(N, np.float32)
    lib().orc_near_far_from_aabb(_p(rays_o), _p(rays_d), _p(aabb), C.c_int64(N), C.c_float(min_near), _p(nears), _p(fars))
    return nears, fars


def morton3D(coords):
    coords = np.ascontiguousarray(coords, np.int32)
    out = np.empty(coords.shape[0], np.int32)
    lib().orc_morton3D(_p(coords), C.c_int64(coords.shape[0]), _p(out))
    return out


def morton3D_invert(indices):
    indices = np.ascontiguousarray(indices, np.int32)
    out = np.empty((indices.shape[0], 3), np.int32)
    lib().orc_morton3D_invert(_p(indices), C.c_int64(indices.shape[0]), _p(out))
    return out


def packbits(grid, thresh):
    """raymarching.cu:269-291.  grid [C,H^3] f32 -> u8 [C*H^3/8]."""
    grid = _f32(grid)
    N = grid.size // 8
    bits = np.empty(N, np.uint8)
    lib().orc_packbits(_p(grid), C.c_int64(N), C.c_float(thresh), _p(bits))
    return bits


def march_rays_train(rays_o, rays_d, bound, bitfield, cascade, H, nears, fars, M, dt_gamma=0.0, max_steps=1024,
                     perturb=False, counter=None):
    """raymarching.cu:313-484, canonical order.  Returns xyzs[M,3], dirs[M,3], deltas[M,2], rays[N,3], counter[2]."""
    rays_o = _f32(rays_o); rays_d = _f32(rays_d); nears = _f32(nears); fars = _f32(fars)
    bitfield = np.ascontiguousarray(bitfield, np.uint8)
    N = rays_o.shape[0]
    xyzs = np.zeros((M, 3), np.float32); dirs = np.zeros((M, 3), np.float32); deltas = np.zeros((M, 2), np.float32)
    rays = np.empty((N, 3), np.int32)
    counter = np.zeros(2, np.int32) if counter is None else counter
    lib().orc_march_rays_train(_p(rays_o), _p(rays_d), _p(bitfield), C.c_float(bound), C.c_float(dt_gamma),
                               C.c_uint32(max_steps), C.c_int64(N), C.c_uint32(cascade), C.c_uint32(H), C.c_int64(M),
                               _p(nears), _p(fars), _p(xyzs), _p(dirs), _p(deltas), _p(rays), _p(counter),
                               C.c_uint32(int(perturb)))
    return xyzs, dirs, deltas, rays, counter


def composite_rays_train_forward(sigmas, rgbs, deltas, rays):
    """raymarching.cu:505-582."""
    sigmas = _f32(sigmas); rgbs = _f32(rgbs); deltas = _f32(deltas); rays = np.ascontiguousarray(rays, np.int32)
    M, N = sigmas.shape[0], rays.shape[0]
    ws = np.empty(N, np.float32); depth = np.empty(N, np.float32); image = np.empty((N, 3), np.float32)
    lib().orc_composite_rays_train_forward(_p(sigmas), _p(rgbs), _p(deltas), _p(rays), C.c_int64(M), C.c_int64(N),
                                           _p(ws), _p(depth), _p(image))
    return ws, depth, image


def composite_rays_train_backward(grad_ws, grad_image, sigmas, rgbs, deltas, rays, weights_sum, image):
    """raymarching.cu:606-688."""
    a = [_f32(v) for v in (grad_ws, grad_image, sigmas, rgbs, deltas)]
    rays = np.ascontiguousarray(rays, np.int32); weights_sum = _f32(weights_sum); image = _f32(image)
    M, N = a[2].shape[0], rays.shape[0]
    gs = np.zeros(M, np.float32); gc = np.zeros((M, 3), np.float32)
    lib().orc_composite_rays_train_backward(_p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), _p(a[4]), _p(rays), _p(weights_sum),
                                            _p(image), C.c_int64(M), C.c_int64(N), _p(gs), _p(gc))
    return gs, gc


def march_rays(n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, bound, bitfield, cascade, H, nears, fars,
               dt_gamma=0.0, max_steps=1024, perturb=0):
    """raymarching.cu:706-814."""
    rays_alive = np.ascontiguousarray(rays_alive, np.int32)
    rays_t, rays_o, rays_d, nears, fars = (_f32(v) for v in (rays_t, rays_o, rays_d, nears, fars))
    bitfield = np.ascontiguousarray(bitfield, np.uint8)
    Mi = n_alive * n_step
    xyzs = np.zeros((Mi, 3), np.float32); dirs = np.zeros((Mi, 3), np.float32); deltas = np.zeros((Mi, 2), np.float32)
    lib().orc_march_rays(C.c_int64(n_alive), C.c_uint32(n_step), _p(rays_alive), _p(rays_t), _p(rays_o), _p(rays_d),
                         C.c_float(bound), C.c_float(dt_gamma), C.c_uint32(max_steps), C.c_uint32(cascade), C.c_uint32(H),
                         _p(bitfield), _p(nears), _p(fars), _p(xyzs), _p(dirs), _p(deltas), C.c_uint32(int(perturb)))
    return xyzs, dirs, deltas


def composite_rays(n_alive, n_step, rays_alive, rays_t, sigmas, rgbs, deltas, weights_sum, depth, image):
    """raymarching.cu:828-913.  In place on rays_alive, rays_t, weights_sum, depth, image (numpy, contiguous)."""
    for v, dt in ((rays_alive, np.int32), (rays_t, np.float32), (weights_sum, np.float32), (depth, np.float32),
                  (image, np.float32)):
        assert v.dtype == dt and v.flags.c_contiguous
    sigmas = _f32(sigmas); rgbs = _f32(rgbs); deltas = _f32(deltas)
    lib().orc_composite_rays(C.c_int64(n_alive), C.c_uint32(n_step), _p(rays_alive), _p(rays_t), _p(sigmas), _p(rgbs),
                             _p(deltas), _p(weights_sum), _p(depth), _p(image))


def pcg32_jitter(seed, n):
    return float(lib().orc_pcg32_jitter(seed, n))


# --------------------------------------------------------------------------------------------- FFMLP (numpy restatement)
def ffmlp_split_weights(weights, input_dim, hidden_dim, num_layers, padded_out=16):
    """Flat fp16 weight vector -> [W_in, W_h..., W_out], each [out, in] (ffmlp.cu:632 layout)."""
    w = np.asarray(weights)
    mats, o = [], 0
    for (r, c) in [(hidden_dim, input_dim)] + [(hidden_dim, hidden_dim)] * (num_layers - 1) + [(padded_out, hidden_dim)]:
        mats.append(w[o:o + r * c].reshape(r, c)); o += r * c
    assert o == w.size
    return mats


def ffmlp_forward(inputs, weights, input_dim, hidden_dim, num_layers, relu=True):
    """kernel_mlp_fused (ffmlp.cu:331-407): h_l = act(h_{l-1} W_l^T) rounded to fp16 after every layer; the products
    are accumulated here in fp32 (the reference accumulates in fp16 wmma fragments; the 2e-3 bar covers the difference).
    Returns (outputs [B,16] f16, forward_buffer [n,B,hidden] f16)."""
    mats = ffmlp_split_weights(weights, input_dim, hidden_dim, num_layers)
    h = np.asarray(inputs, np.float16)
    fwd = []
    for l in range(num_layers):
        z = h.astype(np.float32) @ mats[l].astype(np.float32).T
        if relu:
            z = np.maximum(z, 0)
        h = z.astype(np.float16)
        fwd.append(h)
    out = (h.astype(np.float32) @ mats[-1].astype(np.float32).T).astype(np.float16)
    return out, np.stack(fwd)


def ffmlp_backward(grad, inputs, weights, forward_buffer, input_dim, hidden_dim, num_layers, relu=True):
    """kernel_mlp_fused_backward (ffmlp.cu:410-518) + the weight-gradient products (ffmlp.cu:795-887).
    Returns (grad_inputs [B,in] f16, grad_weights flat f16, backward_buffer [n,B,hidden] f16); backward_buffer[k] is the
    gradient at hidden layer n-1-k (post-ReLU-transfer), as the reference lays it out."""
    mats = ffmlp_split_weights(weights, input_dim, hidden_dim, num_layers)
    g = np.asarray(grad, np.float16)
    x = np.asarray(inputs, np.float16)
    fwd = np.asarray(forward_buffer, np.float16)
    n = num_layers
    gw = [None] * (n + 1)
    bwd = []
    gw[n] = g.astype(np.float32).T @ fwd[n - 1].astype(np.float32)
    d = g.astype(np.float32) @ mats[n].astype(np.float32)
    for k in range(n):
        layer = n - 1 - k                      # hidden layer this gradient belongs to
        if relu:
            d = d * (fwd[layer] > 0)
        dh = d.astype(np.float16)
        bwd.append(dh)
        src = fwd[layer - 1] if layer > 0 else x
        gw[layer] = dh.astype(np.float32).T @ src.astype(np.float32)
        d = dh.astype(np.float32) @ mats[layer].astype(np.float32)
    grad_inputs = d.astype(np.float16)
    grad_weights = np.concatenate([m.reshape(-1) for m in gw]).astype(np.float16)
    return grad_inputs, grad_weights, np.stack(bwd)
