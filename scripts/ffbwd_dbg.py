import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_navigation_b200 import _lib as ngp
dev = torch.device("cuda", 0)
L = ngp.lib(); st = torch.cuda.current_stream().cuda_stream
B, n, ind = 1 << 22, 2, 32
nw = 64 * (ind + 64 * (n - 1) + 16)
xh = (torch.rand(B, ind, device=dev) * 2 - 1).half()
w = ((torch.rand(nw, device=dev) * 2 - 1) * float(np.sqrt(3 / 64))).half()
grad = torch.zeros(B, 16, device=dev, dtype=torch.half); grad[:, :3] = 0.05
outp = torch.empty(B, 16, device=dev, dtype=torch.half); fwd = torch.empty(n, B, 64, device=dev, dtype=torch.half)
gx = torch.empty(B, ind, device=dev, dtype=torch.half); gw = torch.empty(nw, device=dev, dtype=torch.half)
wsb = L.ngp_ffmlp_backward_workspace_bytes(B, ind, n); wsp = torch.empty(wsb, device=dev, dtype=torch.uint8)
ngp.check(L.ngp_ffmlp_forward(xh.data_ptr(), w.data_ptr(), B, ind, 16, 64, n, 0, 6, fwd.data_ptr(), outp.data_ptr(), st))
for i in range(6):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    ngp.check(L.ngp_ffmlp_backward(grad.data_ptr(), xh.data_ptr(), w.data_ptr(), fwd.data_ptr(), B, ind, 16, 64, n, 0, 6, 1, None,
                                   gx.data_ptr(), gw.data_ptr(), wsp.data_ptr(), wsb, st))
    b.record(); torch.cuda.synchronize()
    print(i, a.elapsed_time(b))
