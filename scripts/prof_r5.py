"""Dev aid for ncu: two launches of the fused inference renderer at n rays (default 600,000; impl 1 = tcgen05)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from nerf_navigation_b200 import _lib
from nerf_navigation_b200 import raymarching as rm
from nerf_navigation_b200.field import render_rays_infer
from nerf_navigation_b200.nerf.network import NeRFNetwork
from util import orbit_rays, stonehenge_grid

n = int(sys.argv[1]) if len(sys.argv) > 1 else 600000
impl = int(sys.argv[2]) if len(sys.argv) > 2 else 1
torch.manual_seed(0)
m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10).cuda().eval()
m.encoder.embeddings.data.uniform_(-0.5, 0.5)
m.density_grid.copy_(torch.from_numpy(stonehenge_grid()))
m.density_bitfield = rm.packbits(m.density_grid, 10.0, m.density_bitfield)
_lib.lib().ngp_set_field_impl(impl)
o, d = orbit_rays(n, 3)
ro, rd = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
for _ in range(2):
    out = render_rays_infer(m, ro, rd, m.aabb_infer, 1.0, return_stats=True)
torch.cuda.synchronize()
print("ok", int(out[3][1]))
