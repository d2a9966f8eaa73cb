"""Register the B200 modules under the reference's top-level import names.

Two levels of drop-in:
  * ops only (`install_as_reference_modules(models=False)`): the extension packages the reference's own, unmodified
    `nerf/network.py`, `nerf/network_ff.py` and `nerf/renderer.py` import — `gridencoder`, `raymarching`, `shencoder`,
    `ffmlp`, `encoding`, `activation` — resolve to the B200 implementations; the reference's model / renderer code then
    runs on top of them (tests/test_dropin_cpu.py binds every call site, tests/test_dropin_gpu.py runs it);
  * ops + models (default): additionally `nerf.network`, `nerf.network_ff`, `nerf.renderer` resolve to the host mirrors
    here, which add the fused paths (one-kernel field queries and inference render).
"""
import importlib
import sys

_OPS = {
    "gridencoder": "nerf_navigation_b200.gridencoder",
    "raymarching": "nerf_navigation_b200.raymarching",
    "shencoder": "nerf_navigation_b200.shencoder",
    "ffmlp": "nerf_navigation_b200.ffmlp",
    "encoding": "nerf_navigation_b200.encoding",
    "activation": "nerf_navigation_b200.activation",
}
_MODELS = {
    "nerf.network": "nerf_navigation_b200.nerf.network",
    "nerf.network_ff": "nerf_navigation_b200.nerf.network_ff",
    "nerf.renderer": "nerf_navigation_b200.nerf.renderer",
}
_MAP = {**_OPS, **_MODELS}


def install_as_reference_modules(overwrite=True, models=True):
    """After this call `import gridencoder`, `import raymarching`, `from ffmlp import FFMLP` ... (and, with models=True,
    `from nerf.network import NeRFNetwork`) resolve to the B200 implementations.  `nerf` itself is left to the reference
    package (its utils / provider / gui stay the reference's).  Returns the names that were registered."""
    table = _MAP if models else _OPS
    done = []
    for ref_name, ours in table.items():
        if overwrite or ref_name not in sys.modules:
            sys.modules[ref_name] = importlib.import_module(ours)
            done.append(ref_name)
    return sorted(done)


def uninstall():
    """Remove the aliases again (tests)."""
    for ref_name, ours in _MAP.items():
        mod = sys.modules.get(ref_name)
        if mod is not None and mod.__name__ == ours:
            del sys.modules[ref_name]
