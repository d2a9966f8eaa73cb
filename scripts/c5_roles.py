"""Dev aid: where the roles of k_cfield5_forward wait, at the estimator's size and sample ordering.  Needs a library built
with -DC5_STATS:  VFILES=cfield5.cu scripts/build_variant.sh c5stats -DC5_STATS
                  NGP_B200_LIB=nerf_navigation_b200/build/variants/libngp_c5stats.so python scripts/c5_roles.py"""
import ctypes, sys, os, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import bench_estimator as be
from nerf_navigation_b200 import _lib
dev = torch.device('cuda', 0)
hyp = int(sys.argv[1]) if len(sys.argv) > 1 else 256
model, poses, pix, target, kw = be.setup(dev, hyp, 4096)
st = model.estimator_states
for _ in range(2):
    be.step_states(model, st, pix, target, kw)
torch.cuda.synchronize()
L = _lib.lib()
out = (ctypes.c_ulonglong * 8)()
L.ngp_c5_stats.argtypes = [ctypes.c_void_p, ctypes.c_int]
assert L.ngp_c5_stats(out, 1) == 0
be.step_states(model, st, pix, target, kw)
torch.cuda.synchronize()
assert L.ngp_c5_stats(out, 0) == 0
v = [int(x) for x in out]
g = lambda a, b: f"{100.0 * a / max(b, 1):5.1f} %"
print("cycles summed over reporting threads:", v)
print("gather warps : waiting for the slot's feature columns (f_empty)", g(v[0], v[1]), " of their time")
print("chain warps  : waiting for a layer's accumulator (d_full)      ", g(v[2], v[3]))
print("MMA threads  : waiting for features (f_full)", g(v[4], v[7]), " for the slot's last epilogue (t_free)", g(v[5], v[7]),
      " for the next A operand (a_ready)", g(v[6], v[7]), " issuing", g(v[7] - v[4] - v[5] - v[6], v[7]))
