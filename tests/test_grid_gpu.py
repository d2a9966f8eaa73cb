"""GPU parity: hash-grid encoder (C ABI) vs the CPU oracle and vs the reference's own CUDA build."""
import numpy as np
import pytest
import torch

from util import hashgrid_config, rel_err, rel_err_cols, level_rel_err, load_ref

pytestmark = pytest.mark.gpu


def _dev_scales(ngp, cfg):
    s = torch.empty(cfg["L"], device="cuda")
    ngp.check(ngp.lib().ngp_grid_level_scales(s.data_ptr(), cfg["L"], cfg["S"], cfg["H"], ngp.stream()))
    return s.cpu().numpy()


def _forward(ngp, x01, table, offsets, cfg, layout=1, calc=False, want_idx=False, D=3):
    B = x01.shape[0]; L, C = cfg["L"], table.shape[1]
    out = torch.empty((B, L * C) if layout == 1 else (L, B, C), device="cuda", dtype=table.dtype)
    dy = torch.empty(B, L, D, C, device="cuda", dtype=table.dtype) if calc else None
    ix = torch.empty(B, L, 1 << D, device="cuda", dtype=torch.int32) if want_idx else None
    ngp.check(ngp.lib().ngp_grid_encode_forward(
        x01.data_ptr(), table.data_ptr(), 0 if table.dtype == torch.float32 else 1, offsets.data_ptr(), out.data_ptr(),
        layout, B, D, C, L, cfg["S"], cfg["H"], int(calc), ngp.ptr(dy), 0, 0, ngp.ptr(ix), ngp.stream()), "fwd")
    return out, dy, ix


@pytest.mark.parametrize("B", [1, 1000, 65537])
def test_forward_vs_oracle(ngp, oracle, B):
    cfg = hashgrid_config()
    rng = np.random.default_rng(B)
    x = rng.uniform(0, 1, (B, 3)).astype(np.float32)
    x[0] = [0.0, 1.0, 0.5]
    if B > 10:
        x[3] = [1.5, 0.2, 0.2]   # out of range -> zeros (gridencoder.cu:99-123)
        x[4] = [0.2, -0.1, 0.2]
    scales = _dev_scales(ngp, cfg)
    ref_out, ref_dy, ref_ix = oracle.grid_encode_forward(x, cfg["table"], cfg["offsets"], scales, True, want_indices=True)
    xt = torch.from_numpy(x).cuda(); tt = torch.from_numpy(cfg["table"]).cuda(); ot = torch.from_numpy(cfg["offsets"]).cuda()
    out, dy, ix = _forward(ngp, xt, tt, ot, cfg, layout=0, calc=True, want_idx=True)
    assert np.array_equal(ix.cpu().numpy().view(np.uint32), ref_ix)          # bit-exact indices
    assert rel_err(out.cpu().numpy(), ref_out) < 1e-5
    assert rel_err(dy.cpu().numpy(), ref_dy) < 1e-5
    if B > 10:
        # per (level, channel) feature column and per (level, dim, channel) Jacobian column, each against its own scale
        assert rel_err_cols(out.cpu().numpy().transpose(1, 0, 2).reshape(B, -1), ref_out.transpose(1, 0, 2).reshape(B, -1)) < 1e-5
        assert rel_err_cols(dy.cpu().numpy().reshape(B, -1), np.asarray(ref_dy).reshape(B, -1)) < 1e-5
    out1, _, _ = _forward(ngp, xt, tt, ot, cfg, layout=1)
    assert torch.equal(out1.view(B, cfg["L"], 2).permute(1, 0, 2), out)     # layouts agree (grid.py:52)


def test_device_scales_match_libm(ngp, oracle):
    cfg = hashgrid_config()
    dev = _dev_scales(ngp, cfg)
    host = oracle.level_scales(cfg["L"], cfg["per_level_scale"], cfg["H"])
    # device exp2f is within 2 ulp of libm; cell indices are compared with the DEVICE scales on both sides
    assert np.allclose(dev, host, rtol=3e-7)


def test_backward_input_vs_oracle(ngp, oracle):
    cfg = hashgrid_config()
    B = 20000
    rng = np.random.default_rng(7)
    x = rng.uniform(0, 1, (B, 3)).astype(np.float32)
    x[5] = [1.2, 0.5, 0.5]
    g = rng.standard_normal((B, 32)).astype(np.float32)
    scales = _dev_scales(ngp, cfg)
    _, dy, _ = oracle.grid_encode_forward(x, cfg["table"], cfg["offsets"], scales, True)
    ref = oracle.grid_backward_input(np.ascontiguousarray(g.reshape(B, 16, 2).transpose(1, 0, 2)), dy)
    xt = torch.from_numpy(x).cuda(); tt = torch.from_numpy(cfg["table"]).cuda(); ot = torch.from_numpy(cfg["offsets"]).cuda()
    gt = torch.from_numpy(g).cuda(); gx = torch.empty(B, 3, device="cuda")
    ngp.check(ngp.lib().ngp_grid_encode_backward_input(gt.data_ptr(), 1, xt.data_ptr(), tt.data_ptr(), 0, ot.data_ptr(),
                                                      gx.data_ptr(), B, 3, 2, 16, cfg["S"], cfg["H"], 0, 0, ngp.stream()))
    assert rel_err(gx.cpu().numpy(), ref) < 1e-5


@pytest.mark.parametrize("D,C", [(2, 2), (3, 1), (3, 4), (3, 8), (2, 8)])
def test_other_shapes_vs_oracle(ngp, oracle, D, C):
    offsets, pls = oracle.level_table(D, 8, C, 2.0, 16, 15, 512)
    cfg = dict(L=8, S=float(np.log2(pls)), H=16)
    rng = np.random.default_rng(D * 10 + C)
    table = rng.uniform(-1, 1, (int(offsets[-1]), C)).astype(np.float32)
    x = rng.uniform(0, 1, (3000, D)).astype(np.float32)
    scales = _dev_scales(ngp, cfg)
    ref_out, ref_dy, ref_ix = oracle.grid_encode_forward(x, table, offsets, scales, True, want_indices=True)
    out, dy, ix = _forward(ngp, torch.from_numpy(x).cuda(), torch.from_numpy(table).cuda(),
                           torch.from_numpy(offsets).cuda(), cfg, layout=0, calc=True, want_idx=True, D=D)
    assert np.array_equal(ix.cpu().numpy().view(np.uint32), ref_ix)
    assert rel_err(out.cpu().numpy(), ref_out) < 1e-5
    assert rel_err(dy.cpu().numpy(), ref_dy) < 1e-5


def test_forward_vs_reference_cuda(ngp):
    """Strongest pin: the reference's own kernel (oracle/_ref, built from /root/reference) on identical tensors."""
    ref = load_ref("gridencoder")
    if ref is None:
        pytest.skip("oracle/_ref/gridencoder not built")
    cfg = hashgrid_config()
    B = 200000
    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.rand(B, 3, device="cuda", generator=g)
    tt = torch.from_numpy(cfg["table"]).cuda(); ot = torch.from_numpy(cfg["offsets"]).cuda()
    out_ref = torch.empty(16, B, 2, device="cuda"); dy_ref = torch.empty(B, 96, device="cuda")
    ref.grid_encode_forward(x, tt, ot, out_ref, B, 3, 2, 16, cfg["S"], cfg["H"], True, dy_ref, 0, False)
    out, dy, _ = _forward(ngp, x, tt, ot, cfg, layout=0, calc=True)
    torch.cuda.synchronize()
    # same device exp2f + same fma contraction => identical cells; allow only fp32 rounding differences
    assert rel_err(out.cpu().numpy(), out_ref.cpu().numpy()) < 1e-6
    assert rel_err(dy.view(B, 96).cpu().numpy(), dy_ref.cpu().numpy()) < 1e-6
    frac_equal = (out == out_ref).float().mean().item()
    assert frac_equal > 0.99, frac_equal
    # input backward vs reference kernel_input_backward
    gr = torch.randn(B, 32, device="cuda", generator=g)
    gref_in = torch.zeros(B, 3, device="cuda"); gemb = torch.zeros_like(tt)
    ref.grid_encode_backward(gr.view(B, 16, 2).permute(1, 0, 2).contiguous(), x, tt, ot, gemb, B, 3, 2, 16, cfg["S"],
                             cfg["H"], True, dy_ref, gref_in, 0, False)
    gx = torch.empty(B, 3, device="cuda")
    ngp.check(ngp.lib().ngp_grid_encode_backward_input(gr.data_ptr(), 1, x.data_ptr(), tt.data_ptr(), 0, ot.data_ptr(),
                                                      gx.data_ptr(), B, 3, 2, 16, cfg["S"], cfg["H"], 0, 0, ngp.stream()))
    assert rel_err(gx.cpu().numpy(), gref_in.cpu().numpy()) < 1e-5


def _table_grad(ngp, g, x, cfg, layout=1):
    from nerf_navigation_b200.gridencoder.table_grad import grid_backward_table
    tt = torch.from_numpy(cfg["table"]).cuda(); ot = torch.from_numpy(cfg["offsets"]).cuda()
    B = x.shape[0]
    return grid_backward_table(g, x, tt, ot, (B, 3, 2, 16, cfg["S"], cfg["H"], 0), False, layout)


@pytest.mark.parametrize("method", ["fixed", "sorted"])
def test_table_gradient_vs_oracle_and_deterministic(ngp, oracle, method):
    from nerf_navigation_b200.gridencoder import table_grad
    cfg = hashgrid_config()
    B = 50000
    rng = np.random.default_rng(11)
    x = rng.uniform(0, 1, (B, 3)).astype(np.float32)
    x[7] = [1.1, 0.5, 0.5]                                  # out of range: contributes nothing
    g = rng.standard_normal((B, 32)).astype(np.float32)
    g[100:200] *= 1e-3                                      # small contributions next to large ones
    scales = _dev_scales(ngp, cfg)
    ref = oracle.grid_backward_table(np.ascontiguousarray(g.reshape(B, 16, 2).transpose(1, 0, 2)), x, cfg["offsets"], scales, 2)
    xt, gt = torch.from_numpy(x).cuda(), torch.from_numpy(g).cuda()
    prev = table_grad.METHOD
    try:
        table_grad.METHOD = method
        a = _table_grad(ngp, gt, xt, cfg)
        b = _table_grad(ngp, gt, xt, cfg)
        assert torch.equal(a, b)                            # bit-identical run to run (no float atomics)
        if method == "fixed":
            # integer sums do not depend on the order of the contributions at all: shuffle the batch
            perm = torch.randperm(B, device="cuda")
            c = _table_grad(ngp, gt[perm].contiguous(), xt[perm].contiguous(), cfg)
            assert torch.equal(a, c)
            # fp16 upstream gradient (autocast training) and the [L, B, C] layout take the same path
            gl0 = gt.view(B, 16, 2).permute(1, 0, 2).contiguous()
            assert torch.equal(_table_grad(ngp, gl0, xt, cfg, layout=0), a)
        if method == "fixed":
            # integers cannot carry inf / NaN: an overflowed upstream gradient (fp16 training) gives an all-NaN table
            # gradient, which is what GradScaler's inf check needs to skip the step
            gbad = gt.clone(); gbad[17, 3] = float("inf")
            assert not torch.isfinite(_table_grad(ngp, gbad, xt, cfg)).any()
            gbad[17, 3] = float("nan")
            assert not torch.isfinite(_table_grad(ngp, gbad, xt, cfg)).any()
            assert torch.equal(_table_grad(ngp, gt, xt, cfg), a)      # and the next call is clean again
    finally:
        table_grad.METHOD = prev
    assert rel_err(a.cpu().numpy(), ref) < 1e-5
    assert level_rel_err(a.cpu().numpy(), ref, cfg["offsets"]) < 1e-5      # every level against its own largest entry
    assert (a.cpu().numpy() != 0).sum() == (ref != 0).sum()


def test_table_gradient_atomic_variant(ngp, oracle):
    """table_grad.DETERMINISTIC = False: the reference's own method (float atomics).  Same sums up to fp32 summation order,
    same set of touched rows; rows with a single contribution and zero-gradient samples (the marcher's padding) are exact."""
    from nerf_navigation_b200.gridencoder import table_grad
    cfg = hashgrid_config()
    B = 50000
    rng = np.random.default_rng(12)
    x = rng.uniform(0, 1, (B, 3)).astype(np.float32)
    x[7] = [1.1, 0.5, 0.5]
    x[B // 2:] = 0.0                                        # zero-padded samples all sit on one corner ...
    g = rng.standard_normal((B, 32)).astype(np.float32)
    g[B // 2:] = 0.0                                        # ... and carry no gradient
    xt, gt = torch.from_numpy(x).cuda(), torch.from_numpy(g).cuda()
    det = _table_grad(ngp, gt, xt, cfg)
    try:
        table_grad.DETERMINISTIC = False
        atom = _table_grad(ngp, gt, xt, cfg)
        atom_l0 = _table_grad(ngp, gt.view(B, 16, 2).permute(1, 0, 2).contiguous(), xt, cfg, layout=0)
    finally:
        table_grad.DETERMINISTIC = True
    assert rel_err(atom.cpu().numpy(), det.cpu().numpy()) < 1e-5 and rel_err(atom_l0.cpu().numpy(), det.cpu().numpy()) < 1e-5
    assert torch.equal(atom != 0, det != 0)
    # the finest level is sparse at this batch size: most touched rows have one contribution -> exact there
    fine = slice(int(cfg["offsets"][15]), int(cfg["offsets"][16]))
    assert (atom[fine] == det[fine]).float().mean() > 0.9


def test_table_gradient_vs_reference_cuda(ngp):
    ref = load_ref("gridencoder")
    if ref is None:
        pytest.skip("oracle/_ref/gridencoder not built")
    cfg = hashgrid_config()
    B = 100000
    gen = torch.Generator(device="cuda").manual_seed(5)
    x = torch.rand(B, 3, device="cuda", generator=gen)
    gr = torch.randn(B, 32, device="cuda", generator=gen)
    tt = torch.from_numpy(cfg["table"]).cuda(); ot = torch.from_numpy(cfg["offsets"]).cuda()
    gemb = torch.zeros_like(tt); dummy = torch.zeros(1, device="cuda")
    ref.grid_encode_backward(gr.view(B, 16, 2).permute(1, 0, 2).contiguous(), x, tt, ot, gemb, B, 3, 2, 16, cfg["S"],
                             cfg["H"], False, dummy, dummy, 0, False)
    ours = _table_grad(ngp, gr, x, cfg)
    assert rel_err(ours.cpu().numpy(), gemb.cpu().numpy()) < 1e-5   # atomics order noise only
    assert level_rel_err(ours.cpu().numpy(), gemb.cpu().numpy(), cfg["offsets"]) < 1e-5


def test_grid_encoder_module_autograd(ngp, oracle):
    """GridEncoder module (grid.py:93-156 contract): forward, dL/dx and dL/dtable through autograd."""
    from nerf_navigation_b200.gridencoder import GridEncoder
    torch.manual_seed(0)
    enc = GridEncoder(desired_resolution=4096).cuda()
    enc.embeddings.data.uniform_(-0.5, 0.5)
    assert enc.embeddings.shape == (6328848, 2) and enc.offsets.dtype == torch.int32 and enc.output_dim == 32
    x = (torch.rand(5000, 3, device="cuda") * 4 - 2).requires_grad_()
    out = enc(x, bound=2)
    w = torch.randn_like(out)
    (out * w).sum().backward()
    cfg = dict(L=16, S=float(np.log2(enc.per_level_scale)), H=16)
    scales = _dev_scales(ngp, cfg)
    x01 = ((x.detach() + 2) / 4).cpu().numpy()
    ro, rdy, _ = oracle.grid_encode_forward(x01, enc.embeddings.detach().cpu().numpy(), enc.offsets.cpu().numpy(), scales, True)
    assert rel_err(out.detach().cpu().numpy(), ro.transpose(1, 0, 2).reshape(5000, 32)) < 1e-5
    gl = np.ascontiguousarray(w.cpu().numpy().reshape(5000, 16, 2).transpose(1, 0, 2))
    assert rel_err(x.grad.cpu().numpy(), oracle.grid_backward_input(gl, rdy) / 4.0) < 1e-5
    rt = oracle.grid_backward_table(gl, x01, enc.offsets.cpu().numpy(), scales, 2)
    assert rel_err(enc.embeddings.grad.cpu().numpy(), rt) < 1e-5
    # fp16 table under autocast (grid.py:36-39): output dtype half, 2e-3 bar
    with torch.autocast("cuda", dtype=torch.float16):
        oh = enc(x.detach(), bound=2)
    assert oh.dtype == torch.float16
    assert rel_err(oh.detach().float().cpu().numpy(), ro.transpose(1, 0, 2).reshape(5000, 32)) < 2e-3
