"""dev aid: time the fused inference renderer on one 800x800 frame of the bench scene (orbit rays, stonehenge occupancy)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from nerf_navigation_b200 import raymarching as rm
from nerf_navigation_b200.field import render_rays_infer
from nerf_navigation_b200.nerf.network import NeRFNetwork
from util import orbit_rays, stonehenge_grid
n = int(sys.argv[1]) if len(sys.argv) > 1 else 640000
torch.manual_seed(0)
m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10).cuda().eval()
m.density_grid.copy_(torch.from_numpy(stonehenge_grid()))
m.density_bitfield = rm.packbits(m.density_grid, 10.0, m.density_bitfield)
o, d = orbit_rays(n, 3)
ro, rd = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
for _ in range(2):
    out = render_rays_infer(m, ro, rd, m.aabb_infer, 1.0, return_stats=True)
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(3):
    out = render_rays_infer(m, ro, rd, m.aabb_infer, 1.0, return_stats=True)
b.record(); torch.cuda.synchronize()
ms = a.elapsed_time(b) / 3
ns = int(out[3][1])
print(f"{ms:.2f} ms per {n} rays, {ns} samples -> {ns / ms / 1e6:.3f} Gsamples/s  checksum {float(out[0].sum()):.4f}")
