import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import bench_estimator as be
from torch.profiler import profile, ProfilerActivity
dev = torch.device('cuda', 0)
model, poses, pix, target, kw = be.setup(dev, 64, 4096)
for ch in (8, 16, 32, 64):
    be.step(model, poses, pix, target, kw, ch); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(3): be.step(model, poses, pix, target, kw, ch)
    b.record(); torch.cuda.synchronize()
    print('chunk', ch, 'ms/step', a.elapsed_time(b) / 3, 'rays/s', 64 * 4096 / (a.elapsed_time(b) / 3e3))
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    be.step(model, poses, pix, target, kw, 32); torch.cuda.synchronize()
print(prof.key_averages().table(sort_by='cuda_time_total', row_limit=25, max_name_column_width=60))
