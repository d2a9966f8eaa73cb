"""Dev aid for ncu: the density kernels at the bench size (4,194,304 uniform points), two launches each — forward, backward,
forward_jacobian (single table pass) and its streaming backward."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_navigation_b200 import _lib as ngp
from nerf_navigation_b200.nerf.network import NeRFNetwork
torch.manual_seed(0)
m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=False).cuda().eval()
enc = m.encoder
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 22
x = torch.rand(B, 3, device="cuda") * 2 - 1
sigma = torch.empty(B, device="cuda"); geo = torch.empty(B, 15, device="cuda"); gx = torch.empty(B, 3, device="cuda")
jac = torch.empty(B, 3, device="cuda"); mask = torch.empty(B, device="cuda", dtype=torch.int64); gs = torch.randn(B, device="cuda")
W0 = m.sigma_net[0].weight.detach().contiguous(); W1 = m.sigma_net[1].weight.detach().contiguous()
L = ngp.lib(); st = torch.cuda.current_stream().cuda_stream
a = (x.data_ptr(), 2.0, enc.embeddings.data_ptr(), enc.offsets.data_ptr(), float(np.log2(enc.per_level_scale)), 16, 0, 0,
     W0.data_ptr(), W1.data_ptr())
for _ in range(2):
    ngp.check(L.ngp_field_query_forward(*a, sigma.data_ptr(), geo.data_ptr(), mask.data_ptr(), B, st))
    ngp.check(L.ngp_field_query_backward(*a, sigma.data_ptr(), mask.data_ptr(), gs.data_ptr(), None, gx.data_ptr(), B, st))
    ngp.check(L.ngp_field_query_forward_jacobian(*a, sigma.data_ptr(), geo.data_ptr(), mask.data_ptr(), jac.data_ptr(), B, st))
    ngp.check(L.ngp_field_query_backward_jacobian(jac.data_ptr(), gs.data_ptr(), gx.data_ptr(), B, st))
torch.cuda.synchronize()
print("ok", float(sigma.mean()), float(gx.abs().mean()))
