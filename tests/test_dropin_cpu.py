"""Drop-in proof on the CPU box (north_star: "drops in unchanged for nav/ and nerf/renderer.py"): with
`compat.install_as_reference_modules(models=False)` the reference's OWN, unmodified nerf/renderer.py, nerf/network.py and
nerf/network_ff.py are imported from /root/reference and run on top of the B200 op packages:

  * every call site of `raymarching.*`, `get_encoder`, `trunc_exp` and `FFMLP` in those three files binds against the
    signature of the object this repo registers under that name (positional count and keyword names), found by walking the
    reference's AST — no hand-kept argument list;
  * the reference's NeRFNetwork / NeRFNetwork-ff construct over our encoders, expose the attributes callers read, and their
    state dicts have the keys and shapes a reference checkpoint has.

Skipped where /root/reference does not exist (the GPU box); tests/test_dropin_gpu.py covers the GPU side there through
the host mirrors (same ops, same call order)."""
import ast
import inspect
import os
import sys
from unittest.mock import MagicMock

import pytest
import torch

REF = os.environ.get("NERFNAV_REFERENCE", "/root/reference")
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "nerf")), reason="the reference tree is not on this box")


@pytest.fixture(scope="module")
def ref_nerf():
    """The reference's `nerf` package imported over the B200 op packages."""
    from nerf_navigation_b200 import compat
    saved = dict(sys.modules)
    for name in ["imageio", "tensorboardX", "matplotlib", "matplotlib.pyplot", "cv2", "trimesh", "mcubes", "torch_ema", "lpips",
                 "dearpygui", "dearpygui.dearpygui", "tqdm", "rich", "rich.console", "scipy.spatial.transform"]:
        if name not in sys.modules:
            sys.modules[name] = MagicMock()
    names = compat.install_as_reference_modules(models=False)
    assert set(names) == {"gridencoder", "raymarching", "shencoder", "ffmlp", "encoding", "activation"}
    sys.path.insert(0, REF)
    try:
        import nerf.network as net
        import nerf.network_ff as net_ff
        import nerf.renderer as ren
        assert os.path.realpath(ren.__file__).startswith(os.path.realpath(REF))      # the reference's own file, unmodified
        assert os.path.realpath(net.__file__).startswith(os.path.realpath(REF))
        yield net, net_ff, ren
    finally:
        sys.path.remove(REF)
        for k in list(sys.modules):
            if k not in saved:
                del sys.modules[k]
        compat.uninstall()


def _calls(path, owner=None, func=None):
    """(lineno, n_positional, keyword names) of every call `owner.func(...)` / `func(...)` in the file."""
    out = []
    for node in ast.walk(ast.parse(open(path).read())):
        if not isinstance(node, ast.Call):
            continue
        f = node.func
        if owner is not None and isinstance(f, ast.Attribute) and isinstance(f.value, ast.Name) and f.value.id == owner:
            out.append((f.attr, node.lineno, len(node.args), [k.arg for k in node.keywords]))
        elif func is not None and isinstance(f, ast.Name) and f.id == func:
            out.append((func, node.lineno, len(node.args), [k.arg for k in node.keywords]))
    return out


def _bind(fn, npos, kw, skip_first):
    sig = inspect.signature(fn)
    args = [None] * (npos + (1 if skip_first else 0))
    sig.bind(*args, **{k: None for k in kw})      # raises TypeError if the call would not bind


def test_every_reference_call_site_binds(ref_nerf):
    import raymarching as rm                       # resolves to nerf_navigation_b200.raymarching
    import encoding
    import activation
    import ffmlp
    assert rm.__name__ == "nerf_navigation_b200.raymarching"
    checked = 0
    for fname in ("renderer.py", "network.py", "network_ff.py"):
        path = os.path.join(REF, "nerf", fname)
        for name, line, npos, kw in _calls(path, owner="raymarching"):
            if name == "sph_from_ray":             # background sphere (bg_radius > 0): out of scope, nav uses -1
                continue
            target = getattr(rm, name)             # Function.apply alias: the signature is forward(ctx, ...)
            cls = target.__self__                  # the autograd.Function class behind the alias
            assert cls.__name__ == "_" + name
            try:
                _bind(cls.forward, npos, kw, skip_first=True)
            except TypeError as e:
                pytest.fail(f"{fname}:{line} raymarching.{name}: {e}")
            checked += 1
        for name, line, npos, kw in _calls(path, func="get_encoder"):
            _bind(encoding.get_encoder, npos, kw, skip_first=False)
            checked += 1
        for name, line, npos, kw in _calls(path, func="trunc_exp"):
            _bind(activation._trunc_exp.forward, npos, kw, skip_first=True)
            checked += 1
        for name, line, npos, kw in _calls(path, func="FFMLP"):
            _bind(ffmlp.FFMLP.__init__, npos, kw, skip_first=True)
            checked += 1
    assert checked >= 20


def test_reference_models_construct_over_b200_ops(ref_nerf):
    net, net_ff, ren = ref_nerf
    from nerf_navigation_b200.gridencoder import GridEncoder
    from nerf_navigation_b200.shencoder import SHEncoder
    from nerf_navigation_b200.ffmlp import FFMLP
    m = net.NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10)
    assert isinstance(m, ren.NeRFRenderer) and type(m).__module__ == "nerf.network"
    assert isinstance(m.encoder, GridEncoder) and isinstance(m.encoder_dir, SHEncoder)
    assert m.in_dim == 32 and m.in_dim_dir == 16 and m.cascade == 2 and m.grid_size == 128
    sd = m.state_dict()
    want = {"encoder.embeddings": (6328848, 2), "encoder.offsets": (17,), "sigma_net.0.weight": (64, 32),
            "sigma_net.1.weight": (16, 64), "color_net.0.weight": (64, 31), "color_net.1.weight": (64, 64),
            "color_net.2.weight": (3, 64), "density_grid": (2, 128 ** 3), "density_bitfield": (2 * 128 ** 3 // 8,),
            "step_counter": (16, 2), "aabb_train": (6,), "aabb_infer": (6,)}
    for k, shape in want.items():
        assert tuple(sd[k].shape) == shape, k
    # the same checkpoint loads into this repo's mirror and back (strict)
    from nerf_navigation_b200.nerf.network import NeRFNetwork as Mirror
    mirror = Mirror(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10)
    mirror.load_state_dict(sd, strict=True)
    m.load_state_dict(mirror.state_dict(), strict=True)
    # every attribute the reference's renderer / trainer reads on the model exists on the mirror too
    for attr in ("bound", "cascade", "grid_size", "density_scale", "min_near", "density_thresh", "bg_radius", "cuda_ray",
                 "mean_density", "iter_density", "mean_count", "local_step", "density_grid", "density_bitfield", "step_counter",
                 "aabb_train", "aabb_infer", "render", "run", "run_cuda", "update_extra_state", "mark_untrained_grid",
                 "reset_extra_state", "density", "color", "get_params"):
        assert hasattr(mirror, attr) and hasattr(m, attr), attr
    # the fully fused variant
    f = net_ff.NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=False)
    assert isinstance(f.sigma_net, FFMLP) and isinstance(f.color_net, FFMLP)
    assert f.sigma_net.weights.numel() == 64 * (32 + 64 + 16) and f.color_net.weights.numel() == 64 * (32 + 64 * 2 + 16)
    from nerf_navigation_b200.nerf.network_ff import NeRFNetwork as MirrorFF
    MirrorFF(encoding="hashgrid", bound=2, cuda_ray=False).load_state_dict(f.state_dict(), strict=True)


def test_reference_methods_and_mirror_methods_take_the_same_arguments(ref_nerf):
    """The host mirrors keep the reference's method signatures (names, order, defaults) for everything nav/ and Trainer call."""
    net, net_ff, ren = ref_nerf
    from nerf_navigation_b200.nerf import renderer as mren, network as mnet
    for cls_r, cls_m, names in ((ren.NeRFRenderer, mren.NeRFRenderer,
                                 ["__init__", "run", "run_cuda", "render", "update_extra_state", "mark_untrained_grid",
                                  "reset_extra_state", "density", "color", "forward"]),
                                (net.NeRFNetwork, mnet.NeRFNetwork, ["__init__", "forward", "density", "color", "get_params"])):
        for n in names:
            pr = list(inspect.signature(getattr(cls_r, n)).parameters.values())
            pm = list(inspect.signature(getattr(cls_m, n)).parameters.values())
            assert [p.name for p in pr] == [p.name for p in pm], (cls_r.__name__, n)
            assert [p.default for p in pr] == [p.default for p in pm], (cls_r.__name__, n)
    assert inspect.signature(ren.sample_pdf) == inspect.signature(mren.sample_pdf)
