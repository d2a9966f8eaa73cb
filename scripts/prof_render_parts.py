"""dev aid: where one 800x800 frame of the pre-marched renderer spends its time (plan = count + scan, write, persistent kernel)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from torch.profiler import profile, ProfilerActivity
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
    render_rays_infer(m, ro, rd, m.aabb_infer, 1.0)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        render_rays_infer(m, ro, rd, m.aabb_infer, 1.0)
    torch.cuda.synchronize()
for e in sorted(prof.key_averages(), key=lambda e: -e.device_time_total)[:8]:
    print(f"{e.device_time_total / 3e3:9.3f} ms/frame  x{e.count // 3}  {e.key[:90]}")
