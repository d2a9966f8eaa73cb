"""Dev aid: torch-profiler kernel table of one estimator step on 12-states (all hypotheses in one launch chain)."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import bench_estimator as be
from torch.profiler import profile, ProfilerActivity
dev = torch.device('cuda', 0)
hyp = int(sys.argv[1]) if len(sys.argv) > 1 else 256
model, poses, pix, target, kw = be.setup(dev, hyp, 4096)
st = model.estimator_states
for _ in range(2):
    be.step_states(model, st, pix, target, kw)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(3): be.step_states(model, st, pix, target, kw)
b.record(); torch.cuda.synchronize()
print('ms/step', a.elapsed_time(b) / 3, 'rays/s', hyp * 4096 / (a.elapsed_time(b) / 3e3))
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    be.step_states(model, st, pix, target, kw); torch.cuda.synchronize()
print(prof.key_averages().table(sort_by='cuda_time_total', row_limit=22, max_name_column_width=70))
