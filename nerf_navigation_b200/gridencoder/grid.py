"""Host-side mirror of the reference's gridencoder/grid.py (same class, ctor kwargs, attributes, autograd contract)
on top of the C ABI (include/ngp_b200.h).  Citations are into the reference tree.

Differences by design (documented in DESIGN.md): outputs are produced directly as [B, L*C] (no [L,B,C] + permute copy,
grid.py:52); the backward re-gathers corners instead of storing dy_dx (grid.py:43-46); the table gradient is only
computed when the embeddings require grad (the reference always scatters, grid.py:74-81) and is deterministic.
"""
import numpy as np
import torch
import torch.nn as nn
from torch.autograd import Function

from .. import _lib

_gridtype_to_id = {"hash": 0, "tiled": 1}


def _dtype_id(t):
    if t.dtype == torch.float32:
        return 0
    if t.dtype == torch.float16:
        return 1
    raise RuntimeError("embeddings must be float32 or float16")  # reference: CHECK_IS_FLOATING, gridencoder.cu:18


class _grid_encode(Function):
    """grid.py:19-87.  inputs [B,D] float in [0,1]; embeddings [rows,C]; offsets [L+1] int32 -> [B, L*C]."""

    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda")
    def forward(ctx, inputs, embeddings, offsets, per_level_scale, base_resolution, calc_grad_inputs=False,
                gridtype=0, align_corners=False):
        _lib.require_cuda(inputs, embeddings, offsets)
        inputs = inputs.contiguous().float()
        embeddings = embeddings.contiguous()
        B, D = inputs.shape
        L = offsets.shape[0] - 1
        C = embeddings.shape[1]
        S = float(np.log2(per_level_scale))
        H = int(base_resolution)
        if torch.is_autocast_enabled() and C % 2 == 0:  # grid.py:36-39
            embeddings = embeddings.to(torch.half)
        if offsets.dtype != torch.int32:
            raise RuntimeError("offsets must be an int tensor")  # gridencoder.cu:434
        outputs = torch.empty(B, L * C, device=inputs.device, dtype=embeddings.dtype)
        with torch.cuda.device(inputs.device):
            _lib.check(_lib.lib().ngp_grid_encode_forward(
                _lib.ptr(inputs), _lib.ptr(embeddings), _dtype_id(embeddings), _lib.ptr(offsets), _lib.ptr(outputs), 1,
                B, D, C, L, S, H, 0, None, gridtype, int(align_corners), None, _lib.stream()), "grid_encode_forward")
        ctx.save_for_backward(inputs, embeddings, offsets)
        ctx.dims = (B, D, C, L, S, H, gridtype)
        ctx.calc_grad_inputs = calc_grad_inputs
        ctx.align_corners = align_corners
        return outputs

    @staticmethod
    @torch.amp.custom_bwd(device_type="cuda")
    def backward(ctx, grad):
        # like the reference (grid.py:61-63) this is not double-differentiable: results are constants
        inputs, embeddings, offsets = ctx.saved_tensors
        B, D, C, L, S, H, gridtype = ctx.dims
        grad = grad.contiguous().to(embeddings.dtype)
        grad_inputs = grad_embeddings = None
        with torch.cuda.device(inputs.device):
            if ctx.needs_input_grad[1]:
                from .table_grad import grid_backward_table
                grad_embeddings = grid_backward_table(grad, inputs, embeddings, offsets, ctx.dims, ctx.align_corners)
            if ctx.calc_grad_inputs:
                grad_inputs = torch.empty(B, D, device=inputs.device, dtype=torch.float32)
                _lib.check(_lib.lib().ngp_grid_encode_backward_input(
                    _lib.ptr(grad), 1, _lib.ptr(inputs), _lib.ptr(embeddings), _dtype_id(embeddings), _lib.ptr(offsets),
                    _lib.ptr(grad_inputs), B, D, C, L, S, H, gridtype, int(ctx.align_corners), _lib.stream()),
                    "grid_encode_backward")
        return grad_inputs, grad_embeddings, None, None, None, None, None, None


grid_encode = _grid_encode.apply


def level_offsets(input_dim, num_levels, per_level_scale, base_resolution, log2_hashmap_size, align_corners):
    """grid.py:109-122 — rows per level (capped at 2^log2_hashmap_size, rounded up to 8) and their prefix sum."""
    offsets, offset = [], 0
    max_params = 2 ** log2_hashmap_size
    for i in range(num_levels):
        resolution = int(np.ceil(base_resolution * per_level_scale ** i))
        n = min(max_params, (resolution if align_corners else resolution + 1) ** input_dim)
        n = int(np.ceil(n / 8) * 8)
        offsets.append(offset)
        offset += n
    offsets.append(offset)
    return np.array(offsets, dtype=np.int32)


class GridEncoder(nn.Module):
    """grid.py:93-156 — same constructor, parameter/buffer names (`embeddings`, `offsets`) and attributes."""

    def __init__(self, input_dim=3, num_levels=16, level_dim=2, per_level_scale=2, base_resolution=16,
                 log2_hashmap_size=19, desired_resolution=None, gridtype="hash", align_corners=False):
        super().__init__()
        if desired_resolution is not None:
            per_level_scale = np.exp2(np.log2(desired_resolution / base_resolution) / (num_levels - 1))
        self.input_dim = input_dim
        self.num_levels = num_levels
        self.level_dim = level_dim
        self.per_level_scale = per_level_scale
        self.log2_hashmap_size = log2_hashmap_size
        self.base_resolution = base_resolution
        self.output_dim = num_levels * level_dim
        self.gridtype = gridtype
        self.gridtype_id = _gridtype_to_id[gridtype]
        self.align_corners = align_corners
        self.max_params = 2 ** log2_hashmap_size
        offsets = level_offsets(input_dim, num_levels, per_level_scale, base_resolution, log2_hashmap_size, align_corners)
        self.register_buffer("offsets", torch.from_numpy(offsets))
        self.n_params = int(offsets[-1]) * level_dim
        self.embeddings = nn.Parameter(torch.empty(int(offsets[-1]), level_dim))
        self.reset_parameters()

    def reset_parameters(self):
        std = 1e-4
        self.embeddings.data.uniform_(-std, std)

    def __repr__(self):
        return (f"GridEncoder: input_dim={self.input_dim} num_levels={self.num_levels} level_dim={self.level_dim} "
                f"resolution={self.base_resolution} -> "
                f"{int(round(self.base_resolution * self.per_level_scale ** (self.num_levels - 1)))} "
                f"per_level_scale={self.per_level_scale:.4f} params={tuple(self.embeddings.shape)} "
                f"gridtype={self.gridtype} align_corners={self.align_corners}")

    def forward(self, inputs, bound=1):
        inputs = (inputs + bound) / (2 * bound)  # grid.py:144
        prefix_shape = list(inputs.shape[:-1])
        inputs = inputs.view(-1, self.input_dim)
        outputs = grid_encode(inputs, self.embeddings, self.offsets, self.per_level_scale, self.base_resolution,
                              inputs.requires_grad, self.gridtype_id, self.align_corners)
        return outputs.view(prefix_shape + [self.output_dim])
