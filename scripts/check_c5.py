"""Dev check: tcgen05 colour-field kernels (csrc/cfield5.cu) against the mma.sync kernels (csrc/renderfield.cu) on the same
inputs, forward and backward, plus timings of both.  Usage: python scripts/check_c5.py [M]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_navigation_b200 import _lib  # noqa: E402
from nerf_navigation_b200.field import field_color  # noqa: E402
from nerf_navigation_b200.nerf.network import NeRFNetwork  # noqa: E402


def run(model, x, d, gs, gr):
    xa = x.clone().requires_grad_()
    da = d.clone().requires_grad_()
    s, c = field_color(model, xa, da)
    torch.autograd.backward([s, c], [gs, gr])
    return s.detach(), c.detach(), xa.grad, da.grad


def timeit(fn, n=5):
    fn(); fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    M = int(sys.argv[1]) if len(sys.argv) > 1 else 100003
    torch.manual_seed(0)
    m = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True).cuda().eval()
    m.encoder.embeddings.data.uniform_(-0.5, 0.5)
    L = _lib.lib()
    for name, x in (("uniform", torch.rand(M, 3, device="cuda") * 4.2 - 2.1),      # a few points outside the box
                    ("ray-like", (torch.rand(M // 64 + 1, 1, 3, device="cuda") * 2 - 1 +
                                  torch.linspace(0, 0.2, 64, device="cuda").view(1, 64, 1) *
                                  torch.randn(M // 64 + 1, 1, 3, device="cuda")).reshape(-1, 3)[:M].contiguous())):
        d = torch.nn.functional.normalize(torch.randn(M, 3, device="cuda"), dim=-1)
        gs = torch.randn(M, device="cuda") * 0.1
        gr = torch.randn(M, 3, device="cuda")
        L.ngp_set_field_impl(0)
        ref = run(m, x, d, gs, gr)
        L.ngp_set_field_impl(1)
        new = run(m, x, d, gs, gr)
        torch.cuda.synchronize()
        for nm, a, b in zip(("sigma", "rgb", "grad_x", "grad_d"), new, ref):
            e = (a - b).abs() / b.abs().max()
            err = float(e.max())
            bad = int((~torch.isfinite(a)).sum())
            print(f"[{name}] {nm}: rel err tc5 vs mma {err:.3e}  nonfinite {bad}  elements > 1e-5: {int((e > 1e-5).sum())} of {e.numel()}"
                  f"  rms {float((e.double() ** 2).mean().sqrt()):.2e}")
        for impl in (0, 1):
            L.ngp_set_field_impl(impl)
            xa = x.clone().requires_grad_(); da = d.clone().requires_grad_()
            with torch.no_grad():
                tf = timeit(lambda: field_color(m, x, d))
            s, c = field_color(m, xa, da)
            tb = timeit(lambda: torch.autograd.backward([s, c], [gs, gr], retain_graph=True))
            print(f"[{name}] impl {'tc5' if impl else 'mma'}: fwd {tf:.3f} ms ({M / tf / 1e6:.2f} Gsamples/s)  "
                  f"bwd {tb:.3f} ms ({M / tb / 1e6:.2f} Gsamples/s)")
    L.ngp_set_field_impl(1)


if __name__ == "__main__":
    t0 = time.time()
    main()
    print(f"done in {time.time() - t0:.1f} s")
