"""One small invocation of each ray-path / training kernel for `ncu --set full` captures (dev aid).
  ncu --set full --clock-control none --import-source on -k regex:'k_(render_field|march|composite|tg_|adam)' -c 24 \
      -o gpurun_out/prof_paths python scripts/prof_kernels.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import bench_estimator as BE
import bench_train as BT

dev = torch.device("cuda", 0)
model, poses, pix, target, kw = BE.setup(dev, 32, 4096)
BE.step(model, poses, pix, target, kw, 32)          # 131,072 rays, ~9.8 M samples: march, field fwd/bwd, composite fwd/bwd
torch.cuda.synchronize()
tm, opt, ro, rd, tgt = BT.setup(dev, 4096)
BT.step(tm, opt, ro, rd, tgt, 1)                    # table gradient (keys, sort, reduce, stitch), Adam
torch.cuda.synchronize()
print("ok")
