"""Register the B200 modules under the reference's top-level import names."""
import importlib
import sys

_MAP = {
    "gridencoder": "nerf_navigation_b200.gridencoder",
    "raymarching": "nerf_navigation_b200.raymarching",
    "shencoder": "nerf_navigation_b200.shencoder",
    "encoding": "nerf_navigation_b200.encoding",
    "activation": "nerf_navigation_b200.activation",
    "nerf.network": "nerf_navigation_b200.nerf.network",
    "nerf.renderer": "nerf_navigation_b200.nerf.renderer",
}


def install_as_reference_modules(overwrite=True):
    """After this call `import gridencoder`, `import raymarching`, `from nerf.network import NeRFNetwork` ... resolve to
    the B200 implementations.  `nerf` itself is left to the reference package (utils/provider stay the reference's)."""
    for ref_name, ours in _MAP.items():
        if overwrite or ref_name not in sys.modules:
            sys.modules[ref_name] = importlib.import_module(ours)
    return sorted(_MAP)
