"""Dev aid: where the drop-in host-fed call spends its time — NeRFNetwork.density(x_pinned)['sigma'] -> mean(sigma^2)
-> backward() -> x.grad (pinned) + loss.item(), config-2 size.  Host wall per phase, device time per phase, and the
torch-profiler kernel / memcpy table of one step."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
from nerf_navigation_b200.nerf.network import NeRFNetwork

B = 1 << 22
dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=False).to(dev).eval()
for p in model.parameters():
    p.requires_grad_(False)
xh = (torch.rand(B, 3) * 2 - 1).pin_memory()


def step(marks=None):
    t = [time.perf_counter()]
    xin = xh.detach().requires_grad_(True)
    s = model.density(xin)["sigma"]
    t.append(time.perf_counter())
    loss = (s ** 2).mean()
    t.append(time.perf_counter())
    loss.backward()
    t.append(time.perf_counter())
    v = loss.item()
    t.append(time.perf_counter())
    if marks is not None:
        marks.append([b - a for a, b in zip(t, t[1:])])
    return v, xin.grad


for _ in range(5):
    step()
torch.cuda.synchronize()
marks = []
t0 = time.perf_counter()
for _ in range(20):
    step(marks)
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 20
m = torch.tensor(marks).mean(0) * 1e3
print(f"step {dt*1e3:.3f} ms = {B/dt/1e6:.0f} Msamples/s; host wall per phase [ms]: density {m[0]:.3f}  loss {m[1]:.3f}  backward {m[2]:.3f}  item {m[3]:.3f}")
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    step(); torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=14, max_name_column_width=60))
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
t0 = ev[0].time_range.start
for e in ev:
    print(f"{(e.time_range.start - t0)/1e3:8.3f} +{(e.time_range.end - e.time_range.start)/1e3:7.3f} ms  {e.name[:70]}")
