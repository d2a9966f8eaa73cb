"""Dev check: tcgen05 inference renderer (csrc/render5.cu) against the mma.sync kernel (csrc/render.cu) on the same rays —
images, depths, weight sums and per-ray sample counts — plus timings of both.  Usage: python scripts/check_r5.py [n_rays] [density_scale]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from nerf_navigation_b200 import _lib  # noqa: E402
from nerf_navigation_b200 import raymarching as rm  # noqa: E402
from nerf_navigation_b200.field import render_rays_infer  # noqa: E402
from nerf_navigation_b200.nerf.network import NeRFNetwork  # noqa: E402
from util import orbit_rays, stonehenge_grid  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
    scale = float(sys.argv[2]) if len(sys.argv) > 2 else 25.0
    torch.manual_seed(0)
    m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=scale, min_near=0.2, density_thresh=10).cuda()
    m.encoder.embeddings.data.uniform_(-0.5, 0.5)
    m.sigma_net[1].weight.data[0].abs_().mul_(6.0)
    m.density_grid.copy_(torch.from_numpy(stonehenge_grid()))
    m.density_bitfield = rm.packbits(m.density_grid, 10.0, m.density_bitfield)
    m.eval()
    L = _lib.lib()
    for nn in sorted({1, 37, 4096, n}):
        o, d = orbit_rays(nn, 3)
        ro, rd = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
        res = {}
        for impl in (0, 1):
            L.ngp_set_field_impl(impl)
            out = render_rays_infer(m, ro, rd, m.aabb_infer, 1.0, return_stats=True, return_counts=True)
            torch.cuda.synchronize()
            res[impl] = out
            print(f"n={nn} impl={'tc5' if impl else 'mma'} ok: samples {int(out[3][1])} rays handed out {int(out[3][0])}", flush=True)
        a, b = res[1], res[0]
        dimg = float((a[0] - b[0]).abs().max())
        fin = torch.isfinite(b[1])
        ddep = float((a[1][fin] - b[1][fin]).abs().max()) if fin.any() else 0.0
        print(f"n={nn}: max |image diff| {dimg:.3e}  max |depth diff| {ddep:.3e}  count mismatches {int((a[4] != b[4]).sum())}"
              f"  mean count {float(a[4].float().mean()):.1f}  terminated early (ws > 0.9999): {int((a[2] > 0.9999).sum())}", flush=True)
    # timings at the largest size
    for impl in (0, 1):
        L.ngp_set_field_impl(impl)
        for _ in range(2):
            render_rays_infer(m, ro, rd, m.aabb_infer, 1.0)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            out = render_rays_infer(m, ro, rd, m.aabb_infer, 1.0, return_stats=True)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        ns = int(out[3][1])
        print(f"impl {'tc5' if impl else 'mma'}: {ms:.2f} ms for {nn} rays, {ns} samples -> {ns / ms / 1e6:.3f} Gsamples/s, "
              f"{nn / ms / 1e3:.2f} Mrays/s", flush=True)
        st = out[3][2:].cpu().numpy().astype(float)
        if impl == 1 and st.any():   # -DR5_STATS=1 build: cycle accounting per role (lane 0 of every warp, summed over the grid)
            names = {0: "gather wait rows_full", 1: "gather wait f_empty", 2: "gather total", 3: "mma wait rows_full",
                     4: "mma wait f_full", 5: "mma wait t_free", 6: "mma wait a_ready", 7: "mma total", 8: "chain wait rows_full",
                     9: "chain wait d_full", 10: "chain total", 11: "chain groups of 2", 12: "chain groups", 13: "marcher passes",
                     14: "marcher builds", 15: "marcher passes E < part", 16: "marcher passes buffer busy", 17: "marcher total",
                     18: "marcher end wait", 19: "marcher list entries", 20: "marcher sum E", 21: "marcher rows built"}
            tot = {0: 2, 1: 2, 3: 7, 4: 7, 5: 7, 6: 7, 8: 10, 9: 10, 18: 17}
            for k, nm in names.items():
                extra = f"  ({100 * st[k] / st[tot[k]]:.1f} % of role time)" if k in tot and st[tot[k]] else ""
                print(f"   stat {nm:28s} {st[k]:.4g}{extra}")
            print(f"   per build: passes {st[13] / max(st[14], 1):.2f}  list entries {st[19] / max(st[14], 1):.1f}  rows {st[21] / max(st[14], 1):.1f}"
                  f"   cycles per pass {st[17] / max(st[13], 1):.0f}")
    L.ngp_set_field_impl(1)


if __name__ == "__main__":
    t0 = time.time()
    main()
    print(f"done in {time.time() - t0:.1f} s")
