"""Dev aid: host-fed field query (field.HostFieldQuery) — chunk size / ramp / single-pass sweep at config-2 size."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from nerf_navigation_b200.nerf.network import NeRFNetwork
from nerf_navigation_b200.field import HostFieldQuery

B = 1 << 22
dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=False).to(dev).eval()
xh = (torch.rand(B, 3) * 2 - 1).pin_memory()
gxh = torch.empty(B, 3).pin_memory()
loss_acc = torch.zeros((), device=dev)

def dloss(sig):
    loss_acc.add_((sig * sig).sum())
    return sig * (2.0 / B)

def hook(sig):
    loss_acc.add_((sig * sig).sum())

ref = None
for chunk in (1 << 19, 3 << 18, 1 << 20):
    for ramp in (1, 2, 3, 4):
        for mode in ("callable", "affine"):
            hq = HostFieldQuery(model, B, chunk=chunk, ramp=ramp)
            def step():
                loss_acc.zero_()
                if mode == "callable":
                    hq(xh, dloss, None, gxh)
                else:
                    hq(xh, (0.0, 2.0 / B), None, gxh, sigma_hook=hook)
                return loss_acc.item() / B
            for _ in range(3):
                l = step()
            if ref is None:
                ref = (gxh.clone(), l)
            err = float((gxh - ref[0]).abs().max() / ref[0].abs().max())
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(20):
                step()
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) / 20
            print(f"chunk {chunk:8d} ramp {ramp!s:5} {mode:8}: {dt*1e3:.3f} ms  {B/dt/1e6:.0f} Msamples/s  grad rel diff {err:.1e} loss {l:.6f}", flush=True)
