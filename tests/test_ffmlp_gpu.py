"""GPU parity: FFMLP (tcgen05 kernels) vs the numpy oracle and vs the reference's own CUDA build (fp16 bar: 2e-3)."""
import numpy as np
import pytest
import torch

from util import rel_err, load_ref

pytestmark = pytest.mark.gpu


def _weights(input_dim, num_layers, seed=0):
    n = 64 * (input_dim + 64 * (num_layers - 1) + 16)
    rng = np.random.default_rng(seed)
    return (rng.uniform(-1, 1, n) * np.sqrt(3 / 64)).astype(np.float16)


def _run_forward(ngp, x, w, input_dim, num_layers, training, act=0):
    B = x.shape[0]
    xt, wt = torch.from_numpy(x).cuda(), torch.from_numpy(w).cuda()
    out = torch.empty(B, 16, device="cuda", dtype=torch.half)
    fwd = torch.empty(num_layers, B, 64, device="cuda", dtype=torch.half) if training else None
    ngp.check(ngp.lib().ngp_ffmlp_forward(xt.data_ptr(), wt.data_ptr(), B, input_dim, 16, 64, num_layers, act, 6,
                                          ngp.ptr(fwd), out.data_ptr(), ngp.stream()), "ffmlp_forward")
    torch.cuda.synchronize()
    return out.cpu().numpy(), None if fwd is None else fwd.cpu().numpy()


@pytest.mark.parametrize("B,input_dim,num_layers,training", [(128, 32, 2, True), (1, 32, 2, False), (1000, 32, 3, True),
                                                              (4099, 64, 2, True), (300, 16, 4, False), (70000, 32, 3, False)])
def test_forward_vs_oracle(ngp, oracle, B, input_dim, num_layers, training):
    rng = np.random.default_rng(B)
    x = rng.uniform(-1, 1, (B, input_dim)).astype(np.float16)
    w = _weights(input_dim, num_layers, B)
    ro, rf = oracle.ffmlp_forward(x, w, input_dim, 64, num_layers)
    out, fwd = _run_forward(ngp, x, w, input_dim, num_layers, training)
    assert rel_err(out, ro) < 2e-3
    if training:
        assert rel_err(fwd, rf) < 2e-3


def test_forward_no_activation(ngp, oracle):
    rng = np.random.default_rng(5)
    x = rng.uniform(-1, 1, (513, 32)).astype(np.float16)
    w = _weights(32, 2, 5)
    ro, rf = oracle.ffmlp_forward(x, w, 32, 64, 2, relu=False)
    out, fwd = _run_forward(ngp, x, w, 32, 2, True, act=6)
    assert rel_err(out, ro) < 2e-3 and rel_err(fwd, rf) < 2e-3


def test_forward_vs_reference_cuda(ngp):
    ref = load_ref("ffmlp")
    if ref is None:
        pytest.skip("oracle/_ref/ffmlp not built")
    ref.allocate_splitk(4)
    B, input_dim, n = 4096, 32, 3
    rng = np.random.default_rng(9)
    x = rng.uniform(-1, 1, (B, input_dim)).astype(np.float16)
    w = _weights(input_dim, n, 9)
    xt, wt = torch.from_numpy(x).cuda(), torch.from_numpy(w).cuda()
    r_out = torch.empty(B, 16, device="cuda", dtype=torch.half)
    r_fwd = torch.empty(n, B, 64, device="cuda", dtype=torch.half)
    ref.ffmlp_forward(xt, wt, B, input_dim, 16, 64, n, 0, 6, r_fwd, r_out)
    torch.cuda.synchronize()
    out, fwd = _run_forward(ngp, x, w, input_dim, n, True)
    assert rel_err(out, r_out.cpu().numpy()) < 2e-3
    assert rel_err(fwd, r_fwd.cpu().numpy()) < 2e-3


def test_unsupported_raises(ngp):
    x = torch.zeros(128, 32, device="cuda", dtype=torch.half)
    w = torch.zeros(64 * (32 + 64 + 16), device="cuda", dtype=torch.half)
    out = torch.empty(128, 16, device="cuda", dtype=torch.half)
    rc = ngp.lib().ngp_ffmlp_forward(x.data_ptr(), w.data_ptr(), 128, 32, 16, 128, 2, 0, 6, None, out.data_ptr(), ngp.stream())
    assert rc != 0 and b"hidden_dim" in ngp.lib().ngp_last_error()
