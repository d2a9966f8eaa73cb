"""Two training steps (config 5) with the atomic table gradient, for `ncu --set full` captures of the training-path kernels
(k_tg_atomic, k_linear_small, k_wgrad_partial, k_march_count_warp / k_march_write_warp, k_adam_step).  Dev aid."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import bench_train as BT
from nerf_navigation_b200.gridencoder import table_grad

dev = torch.device("cuda", 0)
tm, opt, ro, rd, tgt = BT.setup(dev, 4096)
table_grad.DETERMINISTIC = os.environ.get("DET", "0") == "1"
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    BT.step(tm, opt, ro, rd, tgt, 1)
torch.cuda.synchronize()
print("ok")
