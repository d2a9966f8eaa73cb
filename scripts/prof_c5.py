"""Dev aid for ncu: one forward + one backward of the fused colour field at M samples (ray-like input)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_navigation_b200 import _lib
from nerf_navigation_b200.field import field_color
from nerf_navigation_b200.nerf.network import NeRFNetwork

M = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 21
impl = int(sys.argv[2]) if len(sys.argv) > 2 else 1
torch.manual_seed(0)
m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True).cuda().eval()
m.encoder.embeddings.data.uniform_(-0.5, 0.5)
_lib.lib().ngp_set_field_impl(impl)
R = M // 64
x = (torch.rand(R, 1, 3, device="cuda") * 2 - 1 + torch.linspace(0, 0.22, 64, device="cuda").view(1, 64, 1) *
     torch.nn.functional.normalize(torch.randn(R, 1, 3, device="cuda"), dim=-1)).reshape(-1, 3).contiguous()
d = torch.nn.functional.normalize(torch.randn(M, 3, device="cuda"), dim=-1)
for _ in range(2):
    xa = x.clone().requires_grad_(); da = d.clone().requires_grad_()
    s, c = field_color(m, xa, da)
    torch.autograd.backward([s, c], [torch.ones_like(s), torch.ones_like(c)])
torch.cuda.synchronize()
print("ok", float(xa.grad.abs().sum()))
