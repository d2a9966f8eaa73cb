"""NeRFNetwork (FFMLP variant) — host-side mirror of the reference's nerf/network_ff.py:10-200: hash-grid encoder ->
FFMLP(32 -> 64 -> 16) -> trunc_exp for density, [SH(16) | geo_feat(15) | 0] -> FFMLP(32 -> 64 -> 64 -> 3) -> sigmoid for
colour, with the same constructor kwargs, parameter names (`encoder.embeddings`, `sigma_net.weights`,
`color_net.weights`) and density()/forward()/color() contracts.  The MLPs run on the tcgen05 kernels (csrc/ffmlp.cu),
the encoders on the grid / SH kernels.  Background model (bg_radius > 0) is out of scope as in nerf/network.py.
"""
import numpy as np
import torch

from .. import _lib
from ..activation import trunc_exp
from ..encoding import get_encoder
from ..ffmlp import FFMLP
from .renderer import NeRFRenderer


class NeRFNetwork(NeRFRenderer):
    def __init__(self, encoding="hashgrid", encoding_dir="sphere_harmonics", num_layers=2, hidden_dim=64,
                 geo_feat_dim=15, num_layers_color=3, hidden_dim_color=64, bound=1, **kwargs):
        super().__init__(bound, **kwargs)
        if self.bg_radius > 0:
            raise NotImplementedError("background model (bg_radius > 0) is out of scope; nav uses bg_radius=-1")
        self.num_layers = num_layers
        self.hidden_dim = hidden_dim
        self.geo_feat_dim = geo_feat_dim
        self.encoder, self.in_dim = get_encoder(encoding, desired_resolution=2048 * bound)
        self.sigma_net = FFMLP(input_dim=self.in_dim, output_dim=1 + self.geo_feat_dim, hidden_dim=self.hidden_dim,
                               num_layers=self.num_layers)
        self.num_layers_color = num_layers_color
        self.hidden_dim_color = hidden_dim_color
        self.encoder_dir, self.in_dim_color = get_encoder(encoding_dir)
        self.in_dim_color += self.geo_feat_dim + 1  # padded to 32 as the reference does (network_ff.py:42)
        self.color_net = FFMLP(input_dim=self.in_dim_color, output_dim=3, hidden_dim=self.hidden_dim_color,
                               num_layers=self.num_layers_color)

    def _color_in(self, d, geo_feat):
        d = self.encoder_dir(d)
        p = torch.zeros_like(geo_feat[..., :1])  # manual input padding (network_ff.py:67-68)
        return torch.cat([d.to(geo_feat.dtype), geo_feat, p], dim=-1)

    def forward(self, x, d):
        """network_ff.py:52-75 -> (sigma [N], rgb [N, 3])."""
        h = self.sigma_net(self.encoder(x, bound=self.bound))
        sigma = trunc_exp(h[..., 0])
        rgb = torch.sigmoid(self.color_net(self._color_in(d, h[..., 1:])))
        return sigma, rgb

    fused_eval = True  # inference density() through the fused encode -> FFMLP kernel (ngp_field_query_ff_forward)

    def _sigma_net_fused(self, x):
        """encoder + sigma_net in one kernel when nothing needs a gradient (the features never reach HBM); None otherwise."""
        enc, net = self.encoder, self.sigma_net
        if not (self.fused_eval and x.is_cuda and x.dtype == torch.float32 and hasattr(enc, "embeddings")
                and not (torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in self.parameters())))
                and enc.input_dim == 3 and enc.num_levels == 16 and enc.level_dim == 2 and enc.embeddings.dtype == torch.float32
                and net.input_dim == 32 and net.hidden_dim == 64 and net.padded_output_dim == 16 and net.activation in (0, 6)):
            return None
        x = x.contiguous()
        B = x.shape[0]
        w = net.weights.detach()
        w = w if w.dtype == torch.half else w.half()
        out = torch.empty(B, 16, device=x.device, dtype=torch.half)
        with torch.cuda.device(x.device):
            _lib.check(_lib.lib().ngp_field_query_ff_forward(
                _lib.ptr(x), float(self.bound), _lib.ptr(enc.embeddings.detach()), _lib.ptr(enc.offsets),
                float(np.log2(enc.per_level_scale)), int(enc.base_resolution), enc.gridtype_id, int(enc.align_corners), _lib.ptr(w),
                net.num_layers, net.activation, _lib.ptr(out), B, _lib.stream()), "field_query_ff_forward")
        return out[:, :net.output_dim] if net.output_dim != 16 else out

    def density(self, x):
        """network_ff.py:76-89 -> {'sigma', 'geo_feat'}; leading dimensions are kept."""
        prefix = x.shape[:-1]
        h = self._sigma_net_fused(x.reshape(-1, 3))
        if h is None:
            h = self.sigma_net(self.encoder(x.reshape(-1, 3), bound=self.bound))
        return {"sigma": trunc_exp(h[..., 0]).view(*prefix), "geo_feat": h[..., 1:].reshape(*prefix, self.geo_feat_dim)}

    def color(self, x, d, mask=None, geo_feat=None, **kwargs):
        """network_ff.py:92-135 (masked colour query)."""
        if mask is not None:
            rgbs = torch.zeros(mask.shape[0], 3, dtype=x.dtype, device=x.device)
            if not mask.any():
                return rgbs
            x, d, geo_feat = x[mask], d[mask], geo_feat[mask]
        h = torch.sigmoid(self.color_net(self._color_in(d, geo_feat)))
        if mask is not None:
            rgbs[mask] = h.to(rgbs.dtype)
        else:
            rgbs = h
        return rgbs

    def get_params(self, lr):
        """network_ff.py:138-151."""
        return [
            {"params": self.encoder.parameters(), "lr": lr},
            {"params": self.sigma_net.parameters(), "lr": lr},
            {"params": self.encoder_dir.parameters(), "lr": lr},
            {"params": self.color_net.parameters(), "lr": lr},
        ]
