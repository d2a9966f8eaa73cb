"""dev aid: time ngp_field_query_value_grad against forward + backward at 4M uniform points."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_navigation_b200 import _lib as ngp
from nerf_navigation_b200.nerf.network import NeRFNetwork
torch.manual_seed(0)
dev = torch.device("cuda", 0)
m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=False).to(dev).eval()
enc = m.encoder
B = 1 << 22
x = torch.rand(B, 3, device=dev) * 2 - 1
sigma = torch.empty(B, device=dev); geo = torch.empty(B, 15, device=dev); gx = torch.empty(B, 3, device=dev)
mask = torch.empty(B, device=dev, dtype=torch.int64); gs = torch.empty(B, device=dev)
W0 = m.sigma_net[0].weight.detach().contiguous(); W1 = m.sigma_net[1].weight.detach().contiguous()
tab = enc.embeddings.detach(); off = enc.offsets
S = float(np.log2(enc.per_level_scale)); H = 16; L = ngp.lib(); st = torch.cuda.current_stream().cuda_stream
def vg(g=True):
    ngp.check(L.ngp_field_query_value_grad(x.data_ptr(), 2.0, tab.data_ptr(), off.data_ptr(), S, H, 0, 0, W0.data_ptr(), W1.data_ptr(),
                                           None, None, 0.0, 2.0 / B, sigma.data_ptr(), geo.data_ptr() if g else None, gx.data_ptr(), B, st))
def two():
    ngp.check(L.ngp_field_query_forward(x.data_ptr(), 2.0, tab.data_ptr(), off.data_ptr(), S, H, 0, 0, W0.data_ptr(), W1.data_ptr(),
                                        sigma.data_ptr(), geo.data_ptr(), mask.data_ptr(), B, st))
    torch.mul(sigma, 2.0 / B, out=gs)
    ngp.check(L.ngp_field_query_backward(x.data_ptr(), 2.0, tab.data_ptr(), off.data_ptr(), S, H, 0, 0, W0.data_ptr(), W1.data_ptr(),
                                         sigma.data_ptr(), mask.data_ptr(), gs.data_ptr(), None, gx.data_ptr(), B, st))
def t(fn, n=30):
    for _ in range(5): fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
print("value_grad (geo)  ms", t(vg), " Msamples/s", B / t(vg) / 1e3)
print("value_grad (-geo) ms", t(lambda: vg(False)))
print("fwd + bwd         ms", t(two), " Msamples/s", B / t(two) / 1e3)
