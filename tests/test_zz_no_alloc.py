"""ngp_linear_wgrad_static — the allocation-free variant of ngp_linear_wgrad (csrc/wgrad.cu): per-CTA partials in a file-scope
static __device__ array of the library instead of a caller workspace.  It has no counterpart in the reference.
Called through the C ABI, twice in a row, against the existing entry point: the two
run the same kernels with the same launch shape, so the results are bit-identical (and within the 1e-5 of the fp64
product that tests/test_training_gpu.py::test_linear_wgrad_kernel asks of the existing one).  Collected last on
purpose."""
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_M = 65536    # NGP_LINEAR_WGRAD_STATIC_MAX_M (include/ngp_b200.h)


def _existing(ngp, dY, X, n_out, n_in):
    L = ngp.lib()
    M = dY.shape[0]
    nb = L.ngp_linear_wgrad_workspace_bytes(M)
    ws = torch.empty(max(nb, 1), dtype=torch.uint8, device="cuda")
    dW = torch.empty(n_out, n_in, device="cuda")
    ngp.check(L.ngp_linear_wgrad(dY.data_ptr(), X.data_ptr(), dW.data_ptr(), M, n_out, n_in, ws.data_ptr(), nb, ngp.stream()),
              "linear_wgrad")
    return dW


def _static(ngp, dY, X, n_out, n_in, dW):
    ngp.check(ngp.lib().ngp_linear_wgrad_static(dY.data_ptr(), X.data_ptr(), dW.data_ptr(), dY.shape[0], n_out, n_in,
                                                ngp.stream()), "linear_wgrad_static")


def test_static_scratch_variant_matches_existing_entry_point_on_two_calls(ngp):
    g = torch.Generator(device="cuda").manual_seed(7)
    # call 1 fills more of the shared array (more CTAs, 64 output rows) than call 2 reads: stale contents must not leak
    shapes = [(MAX_M, 64, 64), (5001, 16, 31)]
    outs = [torch.empty(n_out, n_in, device="cuda") for _, n_out, n_in in shapes]     # allocated before the calls
    data = []
    for M, n_out, n_in in shapes:
        data.append((torch.randn(M, n_out, device="cuda", generator=g), torch.randn(M, n_in, device="cuda", generator=g)))
    for (M, n_out, n_in), (dY, X), dW in zip(shapes, data, outs):                       # two calls in a row
        _static(ngp, dY, X, n_out, n_in, dW)
    torch.cuda.synchronize()
    for (M, n_out, n_in), (dY, X), dW in zip(shapes, data, outs):
        want = _existing(ngp, dY, X, n_out, n_in)
        assert torch.equal(dW, want), (M, n_out, n_in)
        ref64 = dY.double().t() @ X.double()
        err = float((dW.double() - ref64).abs().max() / ref64.abs().max())
        assert err < 1e-5, err
    # the same call again gives the same bits (the array's old contents play no part)
    again = torch.empty_like(outs[0])
    _static(ngp, data[0][0], data[0][1], shapes[0][1], shapes[0][2], again)
    assert torch.equal(again, outs[0])


def test_static_scratch_variant_limits_and_edges(ngp):
    L = ngp.lib()
    dY = torch.zeros(8, 16, device="cuda"); X = torch.zeros(8, 32, device="cuda"); dW = torch.full((16, 32), 3.0, device="cuda")
    rc = L.ngp_linear_wgrad_static(dY.data_ptr(), X.data_ptr(), dW.data_ptr(), MAX_M + 1, 16, 32, ngp.stream())
    assert rc != 0 and b"exceeds the static scratch limit" in L.ngp_last_error()
    with pytest.raises(RuntimeError, match="static scratch limit"):
        ngp.check(L.ngp_linear_wgrad_static(dY.data_ptr(), X.data_ptr(), dW.data_ptr(), 1 << 20, 16, 32, ngp.stream()), "static")
    assert L.ngp_linear_wgrad_static(dY.data_ptr(), X.data_ptr(), dW.data_ptr(), 8, 65, 32, ngp.stream()) != 0     # n_out > 64
    assert L.ngp_linear_wgrad_static(None, X.data_ptr(), dW.data_ptr(), 8, 16, 32, ngp.stream()) != 0
    ngp.check(L.ngp_linear_wgrad_static(dY.data_ptr(), X.data_ptr(), dW.data_ptr(), 0, 16, 32, ngp.stream()), "M = 0")
    assert float(dW.abs().max()) == 0.0                                                # empty batch: dW = 0
