"""dev aid: network_ff.density at 4M points, fused encode->FFMLP kernel vs encoder kernel + FFMLP kernel."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_navigation_b200.nerf.network_ff import NeRFNetwork as NetFF
m = NetFF(encoding="hashgrid", bound=2, cuda_ray=False).cuda().eval()
m.encoder.embeddings.data.uniform_(-0.5, 0.5)
B = 1 << 22
x = torch.rand(B, 3, device="cuda") * 2 - 1
def t(fn, n=20):
    for _ in range(3): fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
with torch.no_grad():
    for fused in (True, False):
        m.fused_eval = fused
        ms = t(lambda: m.density(x))
        print(f"network_ff.density fused={fused}: {ms:.3f} ms per {B} points = {B / ms / 1e3:.0f} Msamples/s")
