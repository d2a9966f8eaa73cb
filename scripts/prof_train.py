"""Dev aid: torch.profiler kernel table of the training step (config 5), deterministic and atomic table gradient."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import bench_train as BT
from nerf_navigation_b200.gridencoder import table_grad
from torch.profiler import profile, ProfilerActivity

dev = torch.device("cuda", 0)
tm, opt, ro, rd, tgt = BT.setup(dev, 4096)
for mode in (True, False):
    table_grad.DETERMINISTIC = mode
    for _ in range(5):
        BT.step(tm, opt, ro, rd, tgt, 1)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(4):
            BT.step(tm, opt, ro, rd, tgt, 1)
        torch.cuda.synchronize()
    print("DETERMINISTIC =", mode)
    rows = [(e.key, e.device_time_total / 4, e.count / 4) for e in prof.key_averages() if e.device_time_total > 0 and e.key.startswith(("k_", "void", "cutlass", "ampere", "sm", "Memset", "Memcpy", "nv"))]
    rows.sort(key=lambda r: -r[1])
    tot = sum(r[1] for r in rows)
    print(f"  kernel time per step {tot/1e3:.3f} ms")
    for k, t, c in rows[:22]:
        print(f"  {t:9.1f} us  x{c:4.1f}  {k[:110]}")
