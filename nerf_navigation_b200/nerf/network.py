"""NeRFNetwork — host-side mirror of the reference's nerf/network.py:10-206 (same ctor kwargs, module/parameter names
`encoder.embeddings`, `encoder.offsets`, `sigma_net.{i}.weight`, `color_net.{i}.weight`, same density()/forward()/color()
contracts) with the hot path routed to the B200 kernels:

* density(x): one fused kernel (hash-grid encode + sigma MLP + trunc_exp), one fused backward-to-x kernel, in eval
  mode (the planner / estimator case: simulate.py:191 model.eval(), :343-347) or whenever the field parameters do
  not require grad, i.e. the field is frozen and gradients flow to x only.  In training mode the unfused GridEncoder -> nn.Linear path runs so autograd produces table/MLP gradients.
* forward(x, d) / color(): GridEncoder + SHEncoder kernels + nn.Linear (torch) — see renderer.py for the fused ray path.

The background model (`bg_radius > 0`, network.py:70-92, 145-160) is out of scope: nav uses bg_radius=-1.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from ..activation import trunc_exp
from ..encoding import get_encoder
from ..field import field_color, field_query, fused_color_supported, fused_supported, linear_nobias
from .renderer import NeRFRenderer


class NeRFNetwork(NeRFRenderer):
    def __init__(self, encoding="hashgrid", encoding_dir="sphere_harmonics", encoding_bg="hashgrid", num_layers=2,
                 hidden_dim=64, geo_feat_dim=15, num_layers_color=3, hidden_dim_color=64, num_layers_bg=2,
                 hidden_dim_bg=64, bound=1, **kwargs):
        super().__init__(bound, **kwargs)
        if self.bg_radius > 0:
            raise NotImplementedError("background model (bg_radius > 0) is out of scope; nav uses bg_radius=-1")
        self.num_layers = num_layers
        self.hidden_dim = hidden_dim
        self.geo_feat_dim = geo_feat_dim
        self.encoder, self.in_dim = get_encoder(encoding, desired_resolution=2048 * bound)
        sigma_net = []
        for l in range(num_layers):
            in_dim = self.in_dim if l == 0 else hidden_dim
            out_dim = 1 + self.geo_feat_dim if l == num_layers - 1 else hidden_dim
            sigma_net.append(nn.Linear(in_dim, out_dim, bias=False))
        self.sigma_net = nn.ModuleList(sigma_net)

        self.num_layers_color = num_layers_color
        self.hidden_dim_color = hidden_dim_color
        self.encoder_dir, self.in_dim_dir = get_encoder(encoding_dir)
        color_net = []
        for l in range(num_layers_color):
            in_dim = self.in_dim_dir + self.geo_feat_dim if l == 0 else hidden_dim
            out_dim = 3 if l == num_layers_color - 1 else hidden_dim
            color_net.append(nn.Linear(in_dim, out_dim, bias=False))
        self.color_net = nn.ModuleList(color_net)
        self.bg_net = None
        self.fused_eval = True  # route eval-mode density() through the fused kernels

    # ------------------------------------------------------------------ helpers
    def _use_fused_density(self, x):
        # eval mode == frozen field (simulate.py:191 calls model.eval()): gradients flow to x only.
        # under torch.no_grad() nothing needs parameter gradients either (update_extra_state's 2 x 128^3 density sweep)
        frozen = not self.training or not torch.is_grad_enabled() or not (
            self.encoder.embeddings.requires_grad or self.sigma_net[0].weight.requires_grad
            or self.sigma_net[1].weight.requires_grad)
        on_gpu = x.is_cuda or (x.device.type == "cpu" and x.is_pinned() and self.encoder.embeddings.is_cuda)   # pinned host points
        return (self.fused_eval and frozen and on_gpu and self.num_layers == 2
                and hasattr(self.encoder, "embeddings")
                and fused_supported(self.encoder, self.sigma_net[0].weight, self.sigma_net[1].weight, x))

    def _sigma_mlp(self, h):
        for l in range(self.num_layers):
            h = linear_nobias(self.sigma_net[l], h)
            if l != self.num_layers - 1:
                h = F.relu(h, inplace=True)
        return h

    def _color_mlp(self, d, geo_feat):
        d = self.encoder_dir(d)
        h = torch.cat([d, geo_feat], dim=-1)
        for l in range(self.num_layers_color):
            h = linear_nobias(self.color_net[l], h)
            if l != self.num_layers_color - 1:
                h = F.relu(h, inplace=True)
        return torch.sigmoid(h)

    # ------------------------------------------------------------------ reference API
    def density(self, x):
        """network.py:126-143 -> {'sigma': [...], 'geo_feat': [..., 15]}."""
        if self._use_fused_density(x):
            if x.requires_grad:
                self._warn_eval_params()
            sigma, geo_feat = field_query(x, self.encoder, self.sigma_net[0].weight.detach(),
                                          self.sigma_net[1].weight.detach(), self.bound)
            if torch.is_autocast_enabled():   # the reference's half sigma_net output there (sigma is fp32 either way)
                geo_feat = geo_feat.to(torch.get_autocast_dtype("cuda"))
            return {"sigma": sigma, "geo_feat": geo_feat}
        h = self._sigma_mlp(self.encoder(x, bound=self.bound))
        return {"sigma": trunc_exp(h[..., 0]), "geo_feat": h[..., 1:]}

    def _warn_eval_params(self):
        """eval() mode takes the fused kernels, whose gradients reach the inputs only (the planner / estimator case:
        simulate.py:191 calls model.eval() and differentiates w.r.t. points / poses).  The reference would ALSO fill
        parameter gradients there; say so once instead of silently leaving them None."""
        if (not self.training and torch.is_grad_enabled() and not getattr(self, "_eval_param_warned", False)
                and any(p.requires_grad for p in self.parameters())):
            import warnings
            self._eval_param_warned = True
            warnings.warn("NeRFNetwork in eval() mode uses the fused field kernels: gradients flow to x / d only, parameter "
                          "gradients are not computed.  Call model.train() (or set model.fused_eval = False) to get them, or "
                          "freeze the parameters (requires_grad_(False)) to silence this.", stacklevel=3)

    def _frozen(self):
        return not self.training or not torch.is_grad_enabled() or not any(p.requires_grad for p in self.parameters())

    def forward(self, x, d):
        """network.py:95-124 -> (sigma [...], color [..., 3]).  Frozen field on CUDA: one fused kernel each way."""
        if (self.fused_eval and self._frozen() and x.is_cuda and self.num_layers == 2 and hasattr(self.encoder, "embeddings")
                and fused_color_supported(self, x, d)):
            if x.requires_grad or d.requires_grad:
                self._warn_eval_params()
            sigma, rgb = field_color(self, x, d)
            if torch.is_autocast_enabled():
                rgb = rgb.to(torch.get_autocast_dtype("cuda"))
            return sigma, rgb
        dens = self.density(x)
        return dens["sigma"], self._color_mlp(d, dens["geo_feat"])

    def color(self, x, d, mask=None, geo_feat=None, **kwargs):
        """network.py:163-191 (masked colour query used by the non-cuda_ray renderer)."""
        if mask is not None:
            rgbs = torch.zeros(mask.shape[0], 3, dtype=x.dtype, device=x.device)
            if not mask.any():
                return rgbs
            x, d, geo_feat = x[mask], d[mask], geo_feat[mask]
        h = self._color_mlp(d, geo_feat)
        if mask is not None:
            rgbs[mask] = h.to(rgbs.dtype)
        else:
            rgbs = h
        return rgbs

    def get_params(self, lr):
        """network.py:194-206."""
        return [
            {"params": self.encoder.parameters(), "lr": lr},
            {"params": self.sigma_net.parameters(), "lr": lr},
            {"params": self.encoder_dir.parameters(), "lr": lr},
            {"params": self.color_net.parameters(), "lr": lr},
        ]
