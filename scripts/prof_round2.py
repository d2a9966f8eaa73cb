"""Dev aid for ONE `ncu --set full` capture of the round-2 tensor-core kernels at bench-like sizes, two launches each:
k_cfield5_forward / backward (4,194,304 ray samples), k_render5_rays_infer (640,000 rays = one 800x800 frame),
k_ffmlp5_forward (inference and training, 4,194,304 rows, 32-64-64-16) and k_ffmlp_backward."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from nerf_navigation_b200 import _lib as ngp
from nerf_navigation_b200 import raymarching as rm
from nerf_navigation_b200.field import field_color, render_rays_infer
from nerf_navigation_b200.nerf.network import NeRFNetwork
from util import orbit_rays, stonehenge_grid

torch.manual_seed(0)
dev = torch.device("cuda", 0)
m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10).cuda().eval()
m.encoder.embeddings.data.uniform_(-0.5, 0.5)
for p in m.parameters():
    p.requires_grad_(False)
m.density_grid.copy_(torch.from_numpy(stonehenge_grid()))
m.density_bitfield = rm.packbits(m.density_grid, 10.0, m.density_bitfield)
M = 1 << 22
R = M // 64
x = (torch.rand(R, 1, 3, device=dev) * 2 - 1 + torch.linspace(0, 0.22, 64, device=dev).view(1, 64, 1) *
     torch.nn.functional.normalize(torch.randn(R, 1, 3, device=dev), dim=-1)).reshape(-1, 3).contiguous()
d = torch.nn.functional.normalize(torch.randn(M, 3, device=dev), dim=-1)
for _ in range(2):
    xa = x.clone().requires_grad_(); da = d.clone().requires_grad_()
    s, c = field_color(m, xa, da)
    torch.autograd.backward([s, c], [torch.ones_like(s), torch.ones_like(c)])
o, dd = orbit_rays(640000, 3)
ro, rd = torch.from_numpy(o).cuda(), torch.from_numpy(dd).cuda()
for _ in range(2):
    out = render_rays_infer(m, ro, rd, m.aabb_infer, 1.0, return_stats=True)
L = ngp.lib(); st = torch.cuda.current_stream().cuda_stream
B, n, ind = 1 << 22, 2, 32
nw = 64 * (ind + 64 * (n - 1) + 16)
xh = (torch.rand(B, ind, device=dev) * 2 - 1).half()
w = ((torch.rand(nw, device=dev) * 2 - 1) * float(np.sqrt(3 / 64))).half()
grad = torch.zeros(B, 16, device=dev, dtype=torch.half); grad[:, :3] = 0.05
outp = torch.empty(B, 16, device=dev, dtype=torch.half); fwd = torch.empty(n, B, 64, device=dev, dtype=torch.half)
gx = torch.empty(B, ind, device=dev, dtype=torch.half); gw = torch.empty(nw, device=dev, dtype=torch.half)
wsb = L.ngp_ffmlp_backward_workspace_bytes(B, ind, n); wsp = torch.empty(wsb, device=dev, dtype=torch.uint8)
for _ in range(2):
    ngp.check(L.ngp_ffmlp_forward(xh.data_ptr(), w.data_ptr(), B, ind, 16, 64, n, 0, 6, None, outp.data_ptr(), st))
    ngp.check(L.ngp_ffmlp_forward(xh.data_ptr(), w.data_ptr(), B, ind, 16, 64, n, 0, 6, fwd.data_ptr(), outp.data_ptr(), st))
    ngp.check(L.ngp_ffmlp_backward(grad.data_ptr(), xh.data_ptr(), w.data_ptr(), fwd.data_ptr(), B, ind, 16, 64, n, 0, 6, 1, None,
                                   gx.data_ptr(), gw.data_ptr(), wsp.data_ptr(), wsb, st))
# network_ff.density through the fused encode -> FFMLP kernel, and two training steps (fixed-point table gradient, warp-per-ray
# compositing backward, two-level weight gradient)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
from nerf_navigation_b200.nerf.network_ff import NeRFNetwork as NetFF
mff = NetFF(encoding="hashgrid", bound=2, cuda_ray=False).cuda().eval()
pts = torch.rand(1 << 22, 3, device=dev) * 2 - 1
with torch.no_grad():
    for _ in range(2):
        mff.density(pts)
import bench_train as BT
tm, opt, tro, trd, tgt = BT.setup(dev, 4096)
for _ in range(2):
    BT.step(tm, opt, tro, trd, tgt, 1)
torch.cuda.synchronize()
print("ok", int(out[3][1]), float(xa.grad.abs().sum()))
