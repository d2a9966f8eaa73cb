"""Fused field query (hash-grid encode -> sigma MLP -> trunc_exp) used by NeRFNetwork.density().

Reference path: nerf/network.py:126-143 (density), gridencoder/grid.py:140-156, activation.py:5-17.
One kernel forward, one kernel backward-to-x; the per-point features / hidden activations never reach HBM.
"""
import numpy as np
import torch
from torch.autograd import Function

from . import _lib


def fused_supported(encoder, W0, W1, x):
    """The fused kernels cover the configuration nerf-navigation uses: 3-D, 16 levels x 2 features, fp32 table,
    sigma net 32 -> 64 -> 16 (network.py:33-47).  Inside an autocast region they still run (in fp32: the kernels take fp32
    pointers and autocast does not touch custom Functions); the callers cast geo_feat / rgb to the autocast dtype, which is
    what the reference's half Linear layers return there (sigma stays fp32: activation.py:7 casts trunc_exp's input)."""
    return (encoder.input_dim == 3 and encoder.num_levels == 16 and encoder.level_dim == 2
            and encoder.embeddings.dtype == torch.float32 and tuple(W0.shape) == (64, 32)
            and tuple(W1.shape) == (16, 64) and x.dtype == torch.float32)


class _field_query(Function):
    """(x [B,3] world coords, embeddings, W0, W1) -> sigma [B], geo_feat [B,15].  Backward: dL/dx only
    (parameter gradients go through the unfused path, see NeRFNetwork.density).

    When x requires a gradient the forward is the single-table-pass kernel that also returns the Jacobian rows
    d sigma_i / d x_i (ngp_field_query_forward_jacobian): sigma_i depends on x_i alone, so the backward of sigma is
    grad_sigma_i * jac_i, a streaming kernel.  Only if a gradient also arrives through geo_feat does the backward run the
    second table pass (ngp_field_query_backward, from the ReLU bits the forward saved)."""

    @staticmethod
    def forward(ctx, x, embeddings, W0, W1, offsets, bound, S, H, gridtype, align_corners):
        _lib.require_cuda(x, embeddings, W0, W1, offsets)
        x = x.contiguous()
        embeddings = embeddings.contiguous(); W0 = W0.contiguous(); W1 = W1.contiguous()
        B = x.shape[0]
        sigma = torch.empty(B, device=x.device, dtype=torch.float32)
        geo = torch.empty(B, 15, device=x.device, dtype=torch.float32)
        need_bwd = ctx.needs_input_grad[0]
        ctx.set_materialize_grads(False)
        with torch.cuda.device(x.device):
            if need_bwd:
                mask = torch.empty(B, device=x.device, dtype=torch.int64)
                jac = torch.empty(B, 3, device=x.device, dtype=torch.float32)
                _lib.check(_lib.lib().ngp_field_query_forward_jacobian(
                    _lib.ptr(x), float(bound), _lib.ptr(embeddings), _lib.ptr(offsets), S, H, gridtype, int(align_corners),
                    _lib.ptr(W0), _lib.ptr(W1), _lib.ptr(sigma), _lib.ptr(geo), _lib.ptr(mask), _lib.ptr(jac), B, _lib.stream()),
                    "field_query_forward_jacobian")
                ctx.save_for_backward(x, embeddings, W0, W1, offsets, sigma, mask, jac)
            else:
                _lib.check(_lib.lib().ngp_field_query_forward(
                    _lib.ptr(x), float(bound), _lib.ptr(embeddings), _lib.ptr(offsets), S, H, gridtype, int(align_corners),
                    _lib.ptr(W0), _lib.ptr(W1), _lib.ptr(sigma), _lib.ptr(geo), None, B, _lib.stream()),
                    "field_query_forward")
        ctx.cfg = (float(bound), S, H, gridtype, int(align_corners))
        return sigma, geo

    @staticmethod
    def backward(ctx, grad_sigma, grad_geo):
        x, embeddings, W0, W1, offsets, sigma, mask, jac = ctx.saved_tensors
        bound, S, H, gridtype, align = ctx.cfg
        B = x.shape[0]
        grad_x = torch.empty(B, 3, device=x.device, dtype=torch.float32)
        if grad_sigma is None and grad_geo is None:
            return (grad_x.zero_(),) + (None,) * 9
        with torch.cuda.device(x.device):
            if grad_geo is None:
                _lib.check(_lib.lib().ngp_field_query_backward_jacobian(
                    _lib.ptr(jac), _lib.ptr(grad_sigma.contiguous().float()), _lib.ptr(grad_x), B, _lib.stream()),
                    "field_query_backward_jacobian")
            else:
                grad_sigma = torch.zeros_like(sigma) if grad_sigma is None else grad_sigma.contiguous().float()
                gg = grad_geo.contiguous().float()
                _lib.check(_lib.lib().ngp_field_query_backward(
                    _lib.ptr(x), bound, _lib.ptr(embeddings), _lib.ptr(offsets), S, H, gridtype, align, _lib.ptr(W0),
                    _lib.ptr(W1), _lib.ptr(sigma), _lib.ptr(mask), _lib.ptr(grad_sigma), _lib.ptr(gg), _lib.ptr(grad_x), B,
                    _lib.stream()), "field_query_backward")
        return grad_x, None, None, None, None, None, None, None, None, None


def _host_chunks(B, chunk=1 << 20):
    """Chunk sizes covering B points: a short first chunk (so the first kernel starts early) and a short last one (so little
    of the last download is exposed), full chunks in between."""
    if B <= chunk:
        return [B] if B else []
    head = [chunk // 4, chunk // 2]
    if B < 2 * sum(head) + chunk:
        return [min(chunk, B - a) for a in range(0, B, chunk)]
    mid = B - 2 * sum(head)
    return head + [chunk] * (mid // chunk) + ([mid % chunk] if mid % chunk else []) + head[::-1]


_host_streams = {}


def _streams_for(dev):
    if dev not in _host_streams:
        _host_streams[dev] = (torch.cuda.Stream(dev), torch.cuda.Stream(dev))
    return _host_streams[dev]


class _field_query_host(Function):
    """density() for points that live in PINNED HOST memory (the planner keeps its states on the host until the query,
    nav/quad_plot.py:224-250): (x_host [B,3]) -> sigma [B], geo_feat [B,15] on the device; backward: dL/dx back in pinned
    host memory.  Forward: the upload of chunk k+1 overlaps the kernel of chunk k (the Jacobian-returning forward of
    _field_query when x requires a gradient); backward: the download of chunk k overlaps the kernel of chunk k+1 (PCIe is
    full duplex, copies on their own streams), so the drop-in call `model.density(x_pinned)` costs about
    max(copies, kernels) instead of their sum."""

    @staticmethod
    def forward(ctx, x, embeddings, W0, W1, offsets, bound, S, H, gridtype, align_corners):
        dev = embeddings.device
        _lib.require_cuda(embeddings, W0, W1, offsets)
        embeddings = embeddings.contiguous(); W0 = W0.contiguous(); W1 = W1.contiguous()
        B = x.shape[0]
        need_bwd = ctx.needs_input_grad[0]
        ctx.set_materialize_grads(False)
        xd = torch.empty(B, 3, device=dev)
        sigma = torch.empty(B, device=dev)
        geo = torch.empty(B, 15, device=dev)
        mask = torch.empty(B, device=dev, dtype=torch.int64) if need_bwd else None
        jac = torch.empty(B, 3, device=dev) if need_bwd else None
        s_in, _ = _streams_for(dev)
        L = _lib.lib()
        with torch.cuda.device(dev):
            cur = torch.cuda.current_stream()
            s_in.wait_stream(cur)                   # the buffers above were allocated on the current stream
            a = 0
            for n in _host_chunks(B):
                with torch.cuda.stream(s_in):
                    xd[a:a + n].copy_(x[a:a + n], non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record(s_in)
                cur.wait_event(ev)
                if need_bwd:
                    _lib.check(L.ngp_field_query_forward_jacobian(
                        xd[a:a + n].data_ptr(), float(bound), _lib.ptr(embeddings), _lib.ptr(offsets), S, H, gridtype,
                        int(align_corners), _lib.ptr(W0), _lib.ptr(W1), sigma[a:a + n].data_ptr(), geo[a:a + n].data_ptr(),
                        mask[a:a + n].data_ptr(), jac[a:a + n].data_ptr(), n, _lib.stream()), "field_query_forward_jacobian")
                else:
                    _lib.check(L.ngp_field_query_forward(
                        xd[a:a + n].data_ptr(), float(bound), _lib.ptr(embeddings), _lib.ptr(offsets), S, H, gridtype,
                        int(align_corners), _lib.ptr(W0), _lib.ptr(W1), sigma[a:a + n].data_ptr(), geo[a:a + n].data_ptr(),
                        None, n, _lib.stream()), "field_query_forward")
                a += n
            xd.record_stream(s_in)
        if need_bwd:
            ctx.save_for_backward(xd, embeddings, W0, W1, offsets, sigma, mask, jac)
        ctx.cfg = (float(bound), S, H, gridtype, int(align_corners))
        return sigma, geo

    @staticmethod
    def backward(ctx, grad_sigma, grad_geo):
        xd, embeddings, W0, W1, offsets, sigma, mask, jac = ctx.saved_tensors
        bound, S, H, gridtype, align = ctx.cfg
        dev = xd.device
        B = xd.shape[0]
        gxh = torch.empty(B, 3, pin_memory=True)
        if grad_sigma is None and grad_geo is None:
            return (gxh.zero_(),) + (None,) * 9
        grad_sigma = torch.zeros_like(sigma) if grad_sigma is None else grad_sigma.contiguous().float()
        gg = None if grad_geo is None else grad_geo.contiguous().float()
        gxd = torch.empty(B, 3, device=dev)
        _, s_out = _streams_for(dev)
        L = _lib.lib()
        with torch.cuda.device(dev):
            cur = torch.cuda.current_stream()
            a = 0
            for n in _host_chunks(B):
                if gg is None:
                    _lib.check(L.ngp_field_query_backward_jacobian(
                        jac[a:a + n].data_ptr(), grad_sigma[a:a + n].data_ptr(), gxd[a:a + n].data_ptr(), n, _lib.stream()),
                        "field_query_backward_jacobian")
                else:
                    _lib.check(L.ngp_field_query_backward(
                        xd[a:a + n].data_ptr(), bound, _lib.ptr(embeddings), _lib.ptr(offsets), S, H, gridtype, align,
                        _lib.ptr(W0), _lib.ptr(W1), sigma[a:a + n].data_ptr(), mask[a:a + n].data_ptr(),
                        grad_sigma[a:a + n].data_ptr(), gg[a:a + n].data_ptr(), gxd[a:a + n].data_ptr(), n, _lib.stream()),
                        "field_query_backward")
                ev = torch.cuda.Event()
                ev.record(cur)
                with torch.cuda.stream(s_out):
                    s_out.wait_event(ev)
                    gxh[a:a + n].copy_(gxd[a:a + n], non_blocking=True)
                a += n
            gxd.record_stream(s_out)
            s_out.synchronize()                     # the caller reads x.grad on the host right away
        return gxh, None, None, None, None, None, None, None, None, None


def field_query(x, encoder, W0, W1, bound):
    """x [..., 3] -> (sigma [...], geo_feat [..., 15]).  x on the device, or in pinned host memory (pipelined copies)."""
    prefix = list(x.shape[:-1])
    S = float(np.log2(encoder.per_level_scale))
    if not x.is_cuda:
        if not x.is_pinned():
            raise RuntimeError("expected a CUDA tensor or a PINNED host tensor (nerf_navigation_b200 has no CPU path)")
        sigma, geo = _field_query_host.apply(x.reshape(-1, 3), encoder.embeddings, W0, W1, encoder.offsets, bound, S,
                                             int(encoder.base_resolution), encoder.gridtype_id, encoder.align_corners)
        return sigma.view(prefix), geo.view(prefix + [15])
    sigma, geo = _field_query.apply(x.reshape(-1, 3), encoder.embeddings, W0, W1, encoder.offsets, bound, S,
                                    int(encoder.base_resolution), encoder.gridtype_id, encoder.align_corners)
    return sigma.view(prefix), geo.view(prefix + [15])


def fused_color_supported(model, x, d):
    """The fused sample kernel covers nav's network: fused_supported() + SH degree 4 + colour net 31-64-64-3."""
    cn = model.color_net
    return (len(cn) == 3 and tuple(cn[0].weight.shape) == (64, 31) and tuple(cn[1].weight.shape) == (64, 64)
            and tuple(cn[2].weight.shape) == (3, 64) and getattr(model.encoder_dir, "degree", None) == 4
            and d.dtype == torch.float32
            and fused_supported(model.encoder, model.sigma_net[0].weight, model.sigma_net[1].weight, x))


class _field_color(Function):
    """(xyzs [M,3], dirs [M,3]) -> sigma [M], rgb [M,3] with a frozen field; backward: dL/dxyzs, dL/ddirs.
    One kernel each way (csrc/renderfield.cu); replaces GridEncoder + 2 Linear + trunc_exp + SHEncoder + cat + 3 Linear +
    sigmoid of NeRFNetwork.forward (network.py:95-124)."""

    @staticmethod
    def forward(ctx, x, d, embeddings, offsets, Ws0, Ws1, Wc0, Wc1, Wc2, bound, S, H, gridtype, align_corners):
        _lib.require_cuda(x, d, embeddings)
        x = x.contiguous(); d = d.contiguous()
        ws = [w.contiguous() for w in (Ws0, Ws1, Wc0, Wc1, Wc2)]
        M = x.shape[0]
        sigma = torch.empty(M, device=x.device, dtype=torch.float32)
        rgb = torch.empty(M, 3, device=x.device, dtype=torch.float32)
        need_bwd = ctx.needs_input_grad[0] or ctx.needs_input_grad[1]
        masks = torch.empty(M, 12, device=x.device, dtype=torch.int16) if need_bwd else None
        with torch.cuda.device(x.device):
            _lib.check(_lib.lib().ngp_field_color_forward(
                _lib.ptr(x), _lib.ptr(d), float(bound), _lib.ptr(embeddings), _lib.ptr(offsets), S, H, gridtype,
                int(align_corners), *[_lib.ptr(w) for w in ws], _lib.ptr(sigma), _lib.ptr(rgb), _lib.ptr(masks), M,
                _lib.stream()), "field_color_forward")
        if need_bwd:
            ctx.save_for_backward(x, d, embeddings, offsets, *ws, sigma, rgb, masks)
        ctx.cfg = (float(bound), S, H, gridtype, int(align_corners))
        return sigma, rgb

    @staticmethod
    def backward(ctx, grad_sigma, grad_rgb):
        x, d, embeddings, offsets, Ws0, Ws1, Wc0, Wc1, Wc2, sigma, rgb, masks = ctx.saved_tensors
        bound, S, H, gridtype, align = ctx.cfg
        M = x.shape[0]
        grad_sigma = torch.zeros_like(sigma) if grad_sigma is None else grad_sigma.contiguous().float()
        grad_rgb = torch.zeros_like(rgb) if grad_rgb is None else grad_rgb.contiguous().float()
        gx = torch.empty(M, 3, device=x.device, dtype=torch.float32)
        gd = torch.empty(M, 3, device=x.device, dtype=torch.float32) if ctx.needs_input_grad[1] else None
        with torch.cuda.device(x.device):
            _lib.check(_lib.lib().ngp_field_color_backward(
                _lib.ptr(x), _lib.ptr(d), bound, _lib.ptr(embeddings), _lib.ptr(offsets), S, H, gridtype, align,
                _lib.ptr(Ws0), _lib.ptr(Ws1), _lib.ptr(Wc0), _lib.ptr(Wc1), _lib.ptr(Wc2), _lib.ptr(sigma), _lib.ptr(rgb),
                _lib.ptr(masks), _lib.ptr(grad_sigma), _lib.ptr(grad_rgb), _lib.ptr(gx), _lib.ptr(gd), M, _lib.stream()),
                "field_color_backward")
        return (gx, gd) + (None,) * 12


def field_color(model, x, d):
    """Fused NeRFNetwork.forward(x, d) for a frozen field: x, d [..., 3] -> (sigma [...], rgb [..., 3])."""
    prefix = list(x.shape[:-1])
    enc = model.encoder
    S = float(np.log2(enc.per_level_scale))
    sigma, rgb = _field_color.apply(
        x.reshape(-1, 3), d.reshape(-1, 3), enc.embeddings.detach(), enc.offsets, model.sigma_net[0].weight.detach(),
        model.sigma_net[1].weight.detach(), model.color_net[0].weight.detach(), model.color_net[1].weight.detach(),
        model.color_net[2].weight.detach(), model.bound, S, int(enc.base_resolution), enc.gridtype_id, enc.align_corners)
    return sigma.view(prefix), rgb.view(prefix + [3])


# render_rays_infer: where the occupancy marching happens (tcgen05 implementation):
#   "segments" (or True)  full-occupancy pre-pass that records every ray's runs of samples; one C call, no host read
#   "lists"               pre-pass in two passes that leaves explicit (t, dt) lists; one host read sizes them
#   False                 inside the persistent kernel
PREMARCH = "segments"
SEGMENT_CAPACITY = 32   # runs per ray kept by the "segments" pre-pass; rays with more go through the marcher-inside kernel


@torch.no_grad()
def render_rays_infer(model, rays_o, rays_d, aabb, bg_color=None, dt_gamma=0.0, max_steps=1024, return_stats=False,
                      return_counts=False, premarch=None):
    """Fused eval branch of NeRFRenderer.run_cuda (renderer.py:325-374): near/far, the alive-ray loop of march_rays ->
    forward(x, d) -> composite_rays, and the background / depth post-processing in ONE kernel (csrc/render.cu).
    rays_o, rays_d [N,3] -> image [N,3], depth [N], weights_sum [N] (already post-processed like run_cuda's results);
    return_stats appends the workspace (its second i64 = samples evaluated), return_counts the per-ray int32 sample counts.
    premarch (default: module flag PREMARCH, tcgen05 implementation only): the occupancy marching runs as a thread-per-ray
    pre-pass at full occupancy and the persistent kernel only evaluates and composites — same samples, same results.
    "segments": one pass records each ray's runs of samples (SEGMENT_CAPACITY per ray), no host read; "lists": count + write
    passes leave explicit (t, dt) lists sized by one host read; False keeps the marcher inside the kernel (no scratch)."""
    _lib.require_cuda(rays_o, rays_d, model.density_bitfield)
    rays_o = rays_o.contiguous().float()
    rays_d = rays_d.contiguous().float()
    N = rays_o.shape[0]
    dev = rays_o.device
    enc = model.encoder
    image = torch.empty(N, 3, device=dev)
    depth = torch.empty(N, device=dev)
    wsum = torch.empty(N, device=dev)
    bg, stride = None, 0
    if bg_color is not None:
        bg = torch.as_tensor(bg_color, dtype=torch.float32, device=dev)
        if bg.numel() == 1:
            bg = bg.reshape(1).expand(3)
        if bg.numel() == 3:
            bg = bg.reshape(3).contiguous()
        else:
            bg, stride = bg.expand(N, 3).contiguous(), 3
    work = torch.empty(_lib.lib().ngp_render_rays_infer_workspace_bytes() // 8, device=dev, dtype=torch.int64)
    counts = torch.empty(N, device=dev, dtype=torch.int32) if return_counts else None
    ws = [w.detach().contiguous() for w in (model.sigma_net[0].weight, model.sigma_net[1].weight, model.color_net[0].weight,
                                            model.color_net[1].weight, model.color_net[2].weight)]
    L = _lib.lib()
    if premarch is None:
        premarch = PREMARCH
    if premarch is True:
        premarch = "segments"
    if not (N > 0 and L.ngp_set_field_impl(-1) == 1):
        premarch = False
    common = lambda: (_lib.ptr(rays_o), _lib.ptr(rays_d), N, _lib.ptr(aabb), float(model.min_near),
                      _lib.ptr(model.density_bitfield), float(model.bound), float(dt_gamma), int(max_steps), int(model.cascade),
                      int(model.grid_size))
    field = lambda: (_lib.ptr(enc.embeddings.detach()), _lib.ptr(enc.offsets), float(np.log2(enc.per_level_scale)),
                     int(enc.base_resolution), enc.gridtype_id, int(enc.align_corners), *[_lib.ptr(w) for w in ws],
                     float(model.density_scale), _lib.ptr(bg), stride, _lib.ptr(image), _lib.ptr(depth), _lib.ptr(wsum),
                     _lib.ptr(counts))
    aabb = aabb.contiguous()
    with torch.cuda.device(dev):
        if premarch == "segments":
            sw_bytes = L.ngp_render_rays_infer_segments_workspace_bytes(N, SEGMENT_CAPACITY)
            work = torch.empty(sw_bytes // 8, device=dev, dtype=torch.int64)
            _lib.check(L.ngp_render_rays_infer_segments(*common(), *field(), SEGMENT_CAPACITY, _lib.ptr(work), sw_bytes,
                                                        _lib.stream()), "render_rays_infer_segments")
            work = work[:64]      # the header: [1] = samples evaluated, [26] = rays that took the overflow path
        elif premarch:
            plan_counts = torch.empty(N, device=dev, dtype=torch.int32)
            plan_ends = torch.empty(N, device=dev, dtype=torch.int64)
            pw_bytes = L.ngp_render_plan_workspace_bytes(N)
            pw = torch.empty(pw_bytes, device=dev, dtype=torch.uint8)
            _lib.check(L.ngp_render_plan(*common(), _lib.ptr(plan_counts), _lib.ptr(plan_ends), _lib.ptr(pw), pw_bytes,
                                         _lib.stream()), "render_plan")
            total = int(plan_ends[-1].item())             # the one host read: sizes the sample lists
            samples = torch.empty(max(total, 1), 2, device=dev)
            _lib.check(L.ngp_render_rays_infer_planned(*common(), *field(), _lib.ptr(plan_counts), _lib.ptr(plan_ends),
                                                       _lib.ptr(samples), total, _lib.ptr(work), _lib.stream()),
                       "render_rays_infer_planned")
        else:
            _lib.check(L.ngp_render_rays_infer(*common(), *field(), _lib.ptr(work), _lib.stream()), "render_rays_infer")
    out = (image, depth, wsum)
    if return_stats:
        out = out + (work,)
    if return_counts:
        out = out + (counts,)
    return out


class HostFieldQuery:
    """Field query for HOST-resident points (pinned memory): sigma and d(sum_i f(sigma_i))/dx for points that live on the
    host, e.g. a planner running on the CPU (the reference's nav/ code keeps its states in host tensors until the query,
    nav/quad_plot.py:224-250).  The batch is cut into chunks and H2D copy / forward+backward kernels / D2H copy of
    consecutive chunks overlap on three streams (PCIe is full duplex), so the step costs max(copies, kernels) instead of
    their sum.  `dloss_dsigma(sigma_chunk) -> grad_chunk` supplies the upstream gradient (elementwise in sigma); a pair
    `(a0, b0)` instead of a callable means dL/dsigma_i = a0 + b0 * sigma_i (b0 = 2/B: mean(sigma^2), the planner's density
    cost) and runs each chunk as ONE kernel (ngp_field_query_value_grad); a callable costs one more streaming kernel
    (Jacobian-returning forward, then dL/dsigma * jac): the table is gathered once either way.  `sigma_hook(sigma_chunk)`, if given, is called
    on the compute stream for every chunk (e.g. to accumulate the loss value on the device).  With `ramp=True` the first
    and last chunks are shorter (chunk/4, chunk/2) so less of the first upload and last download is exposed."""

    def __init__(self, model, max_points, chunk=1 << 19, ramp=True):
        enc = model.encoder
        if not fused_supported(enc, model.sigma_net[0].weight, model.sigma_net[1].weight,
                               torch.empty(0, 3, dtype=torch.float32)):
            raise RuntimeError("HostFieldQuery needs the fused field configuration (16x2 hash grid, 32-64-16 sigma net, fp32)")
        self.model, self.chunk, self.ramp = model, int(chunk), ramp
        dev = enc.embeddings.device
        self.dev = dev
        self.s_in, self.s_cmp, self.s_out = (torch.cuda.Stream(dev) for _ in range(3))
        self.nbuf = 3
        c = self.chunk
        self.bufs = [dict(x=torch.empty(c, 3, device=dev), sigma=torch.empty(c, device=dev), gx=torch.empty(c, 3, device=dev),
                          e_in=torch.cuda.Event(), e_cmp=torch.cuda.Event(), e_out=torch.cuda.Event())
                     for _ in range(self.nbuf)]
        self.S = float(np.log2(enc.per_level_scale))

    def _schedule(self, B):
        """Chunk sizes covering B points, in order."""
        c = self.chunk
        levels = int(self.ramp)          # True -> 1 halving step ... here 2 = (c/4, c/2)
        levels = 2 if self.ramp is True else levels
        if levels <= 0 or (c >> levels) < 1 or B < 4 * c:
            return [min(c, B - a) for a in range(0, B, c)]
        head = [c >> k for k in range(levels, 0, -1)]
        mid = B - 2 * sum(head)
        sizes = head + [c] * (mid // c) + ([mid % c] if mid % c else []) + head[::-1]
        assert sum(sizes) == B
        return sizes

    @torch.no_grad()
    def __call__(self, x_host, dloss_dsigma, sigma_host=None, grad_host=None, sigma_hook=None):
        """x_host [B,3] pinned fp32 -> writes sigma_host [B] / grad_host [B,3] (pinned, optional).  Returns when the
        results are on the host."""
        m, enc, L = self.model, self.model.encoder, _lib.lib()
        if not (x_host.is_pinned() and x_host.dtype == torch.float32 and x_host.is_contiguous()):
            raise RuntimeError("HostFieldQuery: x_host must be a contiguous pinned fp32 tensor")
        B = x_host.shape[0]
        W0 = m.sigma_net[0].weight.detach(); W1 = m.sigma_net[1].weight.detach()
        tab = enc.embeddings.detach(); off = enc.offsets
        H, gt, al, bound = int(enc.base_resolution), enc.gridtype_id, int(enc.align_corners), float(m.bound)
        cur = torch.cuda.current_stream(self.dev)
        for s in (self.s_in, self.s_cmp, self.s_out):
            s.wait_stream(cur)
        with torch.cuda.device(self.dev):
            affine = None if callable(dloss_dsigma) else tuple(float(v) for v in dloss_dsigma)
            a = 0
            for k, n in enumerate(self._schedule(B)):
                if k:
                    a += prev_n
                prev_n = n
                b = self.bufs[k % self.nbuf]
                with torch.cuda.stream(self.s_in):
                    self.s_in.wait_event(b["e_out"])          # the slot's previous results have left the device
                    b["x"][:n].copy_(x_host[a:a + n], non_blocking=True)
                    b["e_in"].record(self.s_in)
                with torch.cuda.stream(self.s_cmp):
                    self.s_cmp.wait_event(b["e_in"])
                    st = self.s_cmp.cuda_stream
                    if affine is not None:
                        _lib.check(L.ngp_field_query_value_grad(b["x"].data_ptr(), bound, tab.data_ptr(), off.data_ptr(), self.S,
                                                                H, gt, al, W0.data_ptr(), W1.data_ptr(), None, None, affine[0],
                                                                affine[1], b["sigma"].data_ptr(), None, b["gx"].data_ptr(), n,
                                                                st), "field_query_value_grad")
                        if sigma_hook is not None:
                            sigma_hook(b["sigma"][:n])
                    else:
                        # sigma and d sigma / dx from ONE table pass (b["gx"] holds the Jacobian rows), then dL/dx = dL/dsigma * jac
                        _lib.check(L.ngp_field_query_forward_jacobian(b["x"].data_ptr(), bound, tab.data_ptr(), off.data_ptr(),
                                                                      self.S, H, gt, al, W0.data_ptr(), W1.data_ptr(),
                                                                      b["sigma"].data_ptr(), None, None, b["gx"].data_ptr(), n, st),
                                   "field_query_forward_jacobian")
                        if sigma_hook is not None:
                            sigma_hook(b["sigma"][:n])
                        gs = dloss_dsigma(b["sigma"][:n]).contiguous().float()
                        _lib.check(L.ngp_field_query_backward_jacobian(b["gx"].data_ptr(), gs.data_ptr(), b["gx"].data_ptr(), n, st),
                                   "field_query_backward_jacobian")
                        gs.record_stream(self.s_cmp)
                    b["e_cmp"].record(self.s_cmp)
                with torch.cuda.stream(self.s_out):
                    self.s_out.wait_event(b["e_cmp"])
                    if sigma_host is not None:
                        sigma_host[a:a + n].copy_(b["sigma"][:n], non_blocking=True)
                    if grad_host is not None:
                        grad_host[a:a + n].copy_(b["gx"][:n], non_blocking=True)
                    b["e_out"].record(self.s_out)
        self.s_out.synchronize()


def _linear_small(x, W, transpose_w):
    """x [..., k] fp32 -> x W^T (transpose_w = 0) or x W (1) through csrc/linear.cu; the result is a fresh (non-view)
    tensor of shape [..., n] so in-place activations may follow."""
    x2 = x.contiguous()
    M = x2.numel() // x2.shape[-1] if x2.shape[-1] else 0
    y = torch.empty(*x.shape[:-1], W.shape[1] if transpose_w else W.shape[0], device=x2.device, dtype=torch.float32)
    with torch.cuda.device(x2.device):
        _lib.check(_lib.lib().ngp_linear_small(_lib.ptr(x2), _lib.ptr(W), _lib.ptr(y), M, W.shape[0], W.shape[1],
                                               int(transpose_w), _lib.stream()), "linear_small")
    return y


class _linear_nobias(Function):
    """y = x W^T for the bias-free Linear layers of the field MLPs when they train (nerf/network.py:33-68): forward and
    dL/dx are csrc/linear.cu (one streaming pass, 3xTF32 tensor-core products), the weight gradient dW = dY^T X (K = all
    samples of the step, output <= 64x64) is csrc/wgrad.cu."""

    @staticmethod
    def forward(ctx, x, W):
        ctx.save_for_backward(x, W)
        return _linear_small(x, W.contiguous(), 0)

    @staticmethod
    def backward(ctx, gy):
        x, W = ctx.saved_tensors
        gx = gW = None
        gy = gy.contiguous()
        if ctx.needs_input_grad[0]:
            gx = _linear_small(gy, W.contiguous(), 1)
        if ctx.needs_input_grad[1]:
            x2 = x.reshape(-1, x.shape[-1]).contiguous()
            g2 = gy.reshape(-1, gy.shape[-1])
            M = x2.shape[0]
            L = _lib.lib()
            nb = L.ngp_linear_wgrad_workspace_bytes(M)
            ws = torch.empty(nb, dtype=torch.uint8, device=x.device)
            gW = torch.empty_like(W)
            with torch.cuda.device(x.device):
                _lib.check(L.ngp_linear_wgrad(_lib.ptr(g2), _lib.ptr(x2), _lib.ptr(gW), M, W.shape[0], W.shape[1], _lib.ptr(ws),
                                              nb, _lib.stream()), "linear_wgrad")
        return gx, gW


def linear_nobias(layer, x):
    """nn.Linear(bias=False) call that routes the weight gradient through csrc/wgrad.cu when it applies."""
    W = layer.weight
    if (layer.bias is None and x.is_cuda and x.dtype == torch.float32 and W.dtype == torch.float32 and W.requires_grad
            and torch.is_grad_enabled() and W.shape[0] <= 64 and W.shape[1] <= 64 and not torch.is_autocast_enabled()):
        return _linear_nobias.apply(x, W)
    return layer(x)


@torch.no_grad()
def field_value_grad(model, x, coef_a=None, coef_b=None, a0=1.0, b0=0.0, want_geo=False):
    """sigma(x) and u * d sigma / d x in ONE kernel (csrc/fieldvg.cu), u_i = a_i + b_i sigma_i per point (scalars a0 / b0
    when the arrays are None).  a0 = 1, b0 = 0: the Jacobian rows d sigma_i / d x_i, from which ANY loss of the densities
    back-propagates with one elementwise product (sigma_i depends on x_i only); coef_b = 2 c_i: the gradient of the
    planner's collision cost sum_i c_i sigma_i^2 (nav/quad_plot.py:224-250) directly.  x [..., 3] -> sigma [...],
    grad [..., 3] (, geo_feat [..., 15])."""
    enc = model.encoder
    W0, W1 = model.sigma_net[0].weight.detach().contiguous(), model.sigma_net[1].weight.detach().contiguous()
    if not fused_supported(enc, W0, W1, x):
        raise RuntimeError("field_value_grad needs the fused field configuration (16x2 hash grid, 32-64-16 sigma net, fp32)")
    _lib.require_cuda(x, enc.embeddings)
    prefix = list(x.shape[:-1])
    xf = x.detach().reshape(-1, 3).contiguous()
    B = xf.shape[0]
    sigma = torch.empty(B, device=x.device); gx = torch.empty(B, 3, device=x.device)
    geo = torch.empty(B, 15, device=x.device) if want_geo else None
    ca = None if coef_a is None else coef_a.reshape(-1).contiguous().float()
    cb = None if coef_b is None else coef_b.reshape(-1).contiguous().float()
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().ngp_field_query_value_grad(
            _lib.ptr(xf), float(model.bound), _lib.ptr(enc.embeddings.detach()), _lib.ptr(enc.offsets),
            float(np.log2(enc.per_level_scale)), int(enc.base_resolution), enc.gridtype_id, int(enc.align_corners),
            _lib.ptr(W0), _lib.ptr(W1), _lib.ptr(ca), _lib.ptr(cb), float(a0), float(b0), _lib.ptr(sigma), _lib.ptr(geo),
            _lib.ptr(gx), B, _lib.stream()), "field_query_value_grad")
    out = (sigma.view(prefix), gx.view(prefix + [3]))
    return out + (geo.view(prefix + [15]),) if want_geo else out


class _sigma_with_jacobian(Function):
    """sigma(x) whose backward is one elementwise product with the Jacobian rows computed in the forward's own kernel."""

    @staticmethod
    def forward(ctx, x, model):
        sigma, jac = field_value_grad(model, x)
        ctx.save_for_backward(jac)
        return sigma

    @staticmethod
    def backward(ctx, grad_sigma):
        (jac,) = ctx.saved_tensors
        return grad_sigma.unsqueeze(-1) * jac, None


def sigma_single_pass(model, x):
    """Differentiable sigma(x) for a frozen field when only the density is needed (the planner's closure): forward AND the
    later backward cost one table pass in total."""
    return _sigma_with_jacobian.apply(x, model)
