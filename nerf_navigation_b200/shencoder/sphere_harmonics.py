"""SHEncoder — host-side mirror of the reference's shencoder/sphere_harmonics.py:61-87 (degree <= 4 is built, the
only degree nerf-navigation uses: encoding.py:45-47 default degree=4).  fp32 forced like the reference (:16)."""
import torch
import torch.nn as nn
from torch.autograd import Function

from .. import _lib


class _sh_encoder(Function):
    @staticmethod
    @torch.amp.custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, inputs, degree, calc_grad_inputs=False):
        _lib.require_cuda(inputs)
        inputs = inputs.contiguous()
        B, D = inputs.shape
        outputs = torch.empty(B, degree ** 2, dtype=inputs.dtype, device=inputs.device)
        with torch.cuda.device(inputs.device):
            _lib.check(_lib.lib().ngp_sh_encode_forward(_lib.ptr(inputs), _lib.ptr(outputs), B, D, degree, 0, None,
                                                       _lib.stream()), "sh_encode_forward")
        ctx.save_for_backward(inputs)
        ctx.dims = (B, D, degree)
        ctx.calc_grad_inputs = calc_grad_inputs
        return outputs

    @staticmethod
    @torch.amp.custom_bwd(device_type="cuda")
    def backward(ctx, grad):
        if not ctx.calc_grad_inputs:
            return None, None, None
        (inputs,) = ctx.saved_tensors
        B, D, degree = ctx.dims
        grad = grad.contiguous().float()
        grad_inputs = torch.empty_like(inputs)
        with torch.cuda.device(inputs.device):
            _lib.check(_lib.lib().ngp_sh_encode_backward(_lib.ptr(grad), _lib.ptr(inputs), B, D, degree,
                                                        _lib.ptr(grad_inputs), _lib.stream()), "sh_encode_backward")
        return grad_inputs, None, None


sh_encode = _sh_encoder.apply


class SHEncoder(nn.Module):
    def __init__(self, input_dim=3, degree=4):
        super().__init__()
        self.input_dim = input_dim
        self.degree = degree
        self.output_dim = degree ** 2
        assert self.input_dim == 3, "SH encoder only support input dim == 3"
        assert 0 < self.degree <= 8, "SH encoder only supports degree in [1, 8]"
        if self.degree > 4:
            raise NotImplementedError("SHEncoder: only degree <= 4 is built (nerf-navigation uses degree 4)")

    def __repr__(self):
        return f"SHEncoder: input_dim={self.input_dim} degree={self.degree}"

    def forward(self, inputs, size=1):
        inputs = inputs / size  # sphere_harmonics.py:80
        prefix_shape = list(inputs.shape[:-1])
        inputs = inputs.reshape(-1, self.input_dim)
        outputs = sh_encode(inputs, self.degree, inputs.requires_grad)
        return outputs.reshape(prefix_shape + [self.output_dim])
