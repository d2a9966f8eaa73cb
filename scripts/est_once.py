"""Dev aid: two estimator steps on 12-states (all hypotheses in one launch chain), nothing else — the program to put under
`ncu -k regex:... --set full` for the ray-op kernels around the two field kernels (run it without ncu first)."""
import sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import bench_estimator as be
dev = torch.device('cuda', 0)
hyp = int(sys.argv[1]) if len(sys.argv) > 1 else 256
model, poses, pix, target, kw = be.setup(dev, hyp, 4096)
st = model.estimator_states
for _ in range(2):
    losses, grads = be.step_states(model, st, pix, target, kw)
torch.cuda.synchronize()
print('ok', float(losses.min()), float(grads.abs().mean()))
