"""GPU side of the drop-in: after compat.install_as_reference_modules() the reference's import statements
(`from nerf.network import NeRFNetwork`, `import raymarching`, `from gridencoder import GridEncoder`, `from ffmlp import FFMLP`,
`from encoding import get_encoder`, `from activation import trunc_exp`: main_nerf.py:6-12, simulate.py:24-31, nerf/network.py:5-7)
resolve to the B200 modules, and the three closures nav/ uses (simulate.py:343-347) run through them."""
import sys

import pytest
import torch

from util import stonehenge_grid

pytestmark = pytest.mark.gpu


def test_reference_import_names_run_on_the_gpu():
    from nerf_navigation_b200 import compat
    saved = dict(sys.modules)
    try:
        names = compat.install_as_reference_modules()
        assert {"nerf.network", "nerf.network_ff", "nerf.renderer", "ffmlp", "raymarching"} <= set(names)
        from nerf.network import NeRFNetwork            # noqa: the reference's import line
        from nerf.network_ff import NeRFNetwork as NeRFNetworkFF
        import raymarching
        from gridencoder import GridEncoder
        from ffmlp import FFMLP
        from encoding import get_encoder
        from activation import trunc_exp
        assert NeRFNetwork.__module__ == "nerf_navigation_b200.nerf.network"
        torch.manual_seed(0)
        model = NeRFNetwork(encoding="hashgrid", bound=2, cuda_ray=True, density_scale=1, min_near=0.2, density_thresh=10).cuda()
        model.encoder.embeddings.data.uniform_(-0.5, 0.5)
        model.density_grid.copy_(torch.from_numpy(stonehenge_grid()))
        model.density_bitfield = raymarching.packbits(model.density_grid, 10.0, model.density_bitfield)
        model.eval()
        assert isinstance(model.encoder, GridEncoder)
        # simulate.py:343 — the planner's density closure, with the gradient to the points
        x = (torch.rand(20, 100, 3, device="cuda") * 2 - 1).requires_grad_(True)
        sigma = model.density(x.reshape(-1, 3))["sigma"]
        (sigma ** 2).mean().backward()
        assert sigma.shape == (2000,) and torch.isfinite(x.grad).all() and x.grad.abs().sum() > 0
        # simulate.py:345 — the estimator's render closure (eval: the fused kernel; with grad: the train branch)
        o = torch.tensor([[1.2, 0.0, 0.4]], device="cuda").repeat(512, 1)
        d = torch.nn.functional.normalize(-o + 0.3 * (torch.rand(512, 3, device="cuda") - 0.5), dim=-1)
        with torch.no_grad():
            out = model.render(o[None], d[None], staged=True, bg_color=1.0, perturb=False)
        assert out["image"].shape == (1, 512, 3) and out["depth"].shape == (1, 512)
        # the FFMLP variant and the free functions
        ff = NeRFNetworkFF(encoding="hashgrid", bound=2, cuda_ray=False).cuda().eval()
        assert isinstance(ff.sigma_net, FFMLP)
        with torch.no_grad():
            s2 = ff.density(torch.rand(300, 3, device="cuda") * 2 - 1)["sigma"]
        assert s2.shape == (300,) and torch.isfinite(s2).all()
        enc, dim = get_encoder("sphere_harmonics")
        assert dim == 16 and torch.allclose(trunc_exp(torch.zeros(3, device="cuda")), torch.ones(3, device="cuda"))
    finally:
        compat.uninstall()
        for k in list(sys.modules):
            if k not in saved and (k == "nerf" or k.startswith("nerf.")):
                del sys.modules[k]
