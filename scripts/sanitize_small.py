"""dev aid: the round-2 kernels once each at small, ragged sizes (written for compute-sanitizer memcheck, which is closed on
this pool; run plain it still exercises every new entry point with sizes that are not multiples of a tile)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "scripts"))
from nerf_navigation_b200 import _lib as ngp, field, raymarching as rm
from nerf_navigation_b200.nerf.network import NeRFNetwork
from nerf_navigation_b200.nerf.network_ff import NeRFNetwork as NetFF
from util import orbit_rays, stonehenge_grid
torch.manual_seed(0)
dev = torch.device("cuda", 0)
m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=25, min_near=0.2, density_thresh=10).cuda().eval()
m.encoder.embeddings.data.uniform_(-0.5, 0.5)
for p in m.parameters():
    p.requires_grad_(False)
m.density_grid.copy_(torch.from_numpy(stonehenge_grid()))
m.density_bitfield = rm.packbits(m.density_grid, 10.0, m.density_bitfield)
# density: forward, Jacobian forward + both backwards, single-pass, pinned host
x = (torch.rand(5003, 3, device=dev) * 4.2 - 2.1)
with torch.no_grad():
    m.density(x)
xi = x.clone().requires_grad_(True)
out = m.density(xi); (out["sigma"] ** 2).sum().backward()
xi = x.clone().requires_grad_(True)
out = m.density(xi); ((out["sigma"] ** 2).sum() + out["geo_feat"].sum()).backward()
field.field_value_grad(m, x, a0=0.0, b0=2.0 / 5003, want_geo=True)
xh = x.cpu().pin_memory().requires_grad_(True)
(m.density(xh)["sigma"] ** 2).mean().backward()
# colour field fwd / bwd
d = torch.nn.functional.normalize(torch.randn(5003, 3, device=dev), dim=-1)
xa = x.clone().requires_grad_(True); da = d.clone().requires_grad_(True)
s, c = field.field_color(m, xa, da); (s.sum() + c.sum()).backward()
# renderer: runs (default), runs with overflow, lists, marcher inside
o, dd = orbit_rays(3001, 3)
ro, rd = torch.from_numpy(o).cuda(), torch.from_numpy(dd).cuda()
for mode, cap in (("segments", 32), ("segments", 1), ("lists", 32), (False, 32)):
    field.SEGMENT_CAPACITY = cap
    img = field.render_rays_infer(m, ro, rd, m.aabb_infer, 1.0, premarch=mode, return_counts=True)
field.SEGMENT_CAPACITY = 32
# network_ff fused density, FFMLP training fwd/bwd
mf = NetFF(encoding="hashgrid", bound=2, cuda_ray=False).cuda().eval()
with torch.no_grad():
    mf.density(x)
mf.train()
h = mf.density(x); (h["sigma"].float().sum() + h["geo_feat"].float().sum()).backward()
# one training step (fixed-point table gradient, warp compositing backward, wgrad) + the static-scratch wgrad
import bench_train as BT
tm, opt, tro, trd, tgt = BT.setup(dev, 1024)
BT.step(tm, opt, tro, trd, tgt, 1)
L = ngp.lib()
dY = torch.randn(3000, 16, device=dev); X = torch.randn(3000, 31, device=dev); dW = torch.empty(16, 31, device=dev)
ngp.check(L.ngp_linear_wgrad_static(dY.data_ptr(), X.data_ptr(), dW.data_ptr(), 3000, 16, 31, ngp.stream()), "static")
torch.cuda.synchronize()
print("sanitize_small ok")
