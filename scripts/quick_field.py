"""Quick device timing of the fused field query at config-2 size (dev aid, not the bench)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from util import hashgrid_config, sigma_net_weights
from nerf_navigation_b200 import _lib as ngp

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 22
cfg = hashgrid_config()
W0, W1 = sigma_net_weights()
dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
tt, ot, w0, w1 = dev(cfg["table"]), dev(cfg["offsets"]), dev(W0), dev(W1)
x = torch.rand(B, 3, device="cuda") * 2 - 1
gs = torch.randn(B, device="cuda")
sigma = torch.empty(B, device="cuda"); geo = torch.empty(B, 15, device="cuda")
mask = torch.empty(B, device="cuda", dtype=torch.int64); gx = torch.empty(B, 3, device="cuda")
L = ngp.lib()
def fwd():
    ngp.check(L.ngp_field_query_forward(x.data_ptr(), 2.0, tt.data_ptr(), ot.data_ptr(), cfg["S"], 16, 0, 0, w0.data_ptr(), w1.data_ptr(), sigma.data_ptr(), geo.data_ptr(), mask.data_ptr(), B, ngp.stream()))
def bwd():
    ngp.check(L.ngp_field_query_backward(x.data_ptr(), 2.0, tt.data_ptr(), ot.data_ptr(), cfg["S"], 16, 0, 0, w0.data_ptr(), w1.data_ptr(), sigma.data_ptr(), mask.data_ptr(), gs.data_ptr(), None, gx.data_ptr(), B, ngp.stream()))
def enc():
    out = torch.empty(B, 32, device="cuda")
    x01 = (x + 2) / 4
    def f():
        ngp.check(L.ngp_grid_encode_forward(x01.data_ptr(), tt.data_ptr(), 0, ot.data_ptr(), out.data_ptr(), 1, B, 3, 2, 16, cfg["S"], 16, 0, None, 0, 0, None, ngp.stream()))
    return f
def timeit(f, n=10):
    for _ in range(3): f()
    torch.cuda.synchronize()
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): f()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
for impl in ((0, 1) if len(sys.argv) < 3 else (int(sys.argv[2]),)):
  L.ngp_set_field_impl(impl)
  tf, tb, te = timeit(fwd), timeit(bwd), timeit(enc())
  print(f"impl={'tc5' if impl else 'mma'} B={B} fwd {tf:.3f} ms  bwd {tb:.3f} ms  enc-only {te:.3f} ms  -> {B/(tf+tb)/1e3:.1f} Msamples/s fwd+bwd; gather GB/s fwd {B*1024/tf/1e6:.0f} bwd {B*1024/tb/1e6:.0f}")
