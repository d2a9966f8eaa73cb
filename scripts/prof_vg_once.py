"""dev aid: ONE launch of ngp_field_query_value_grad at 4M uniform points (for -DD5_STATS=1 builds / ncu)."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_navigation_b200 import _lib as ngp
from nerf_navigation_b200.nerf.network import NeRFNetwork
torch.manual_seed(0)
m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=False).cuda().eval()
enc = m.encoder
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 22
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1
x = torch.rand(B, 3, device="cuda") * 2 - 1
sigma = torch.empty(B, device="cuda"); gx = torch.empty(B, 3, device="cuda")
W0 = m.sigma_net[0].weight.detach().contiguous(); W1 = m.sigma_net[1].weight.detach().contiguous()
L = ngp.lib(); st = torch.cuda.current_stream().cuda_stream
for _ in range(n):
    ngp.check(L.ngp_field_query_value_grad(x.data_ptr(), 2.0, enc.embeddings.data_ptr(), enc.offsets.data_ptr(), float(np.log2(enc.per_level_scale)),
                                           16, 0, 0, W0.data_ptr(), W1.data_ptr(), None, None, 0.0, 2.0 / B, sigma.data_ptr(), None, gx.data_ptr(), B, st))
torch.cuda.synchronize()
print("done", float(sigma.mean()), float(gx.abs().mean()))
