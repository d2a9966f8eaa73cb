"""FFMLP — host-side mirror of the reference's ffmlp/ffmlp.py:15-169 on the B200 tcgen05 kernels (csrc/ffmlp.cu).

Same class name, constructor arguments, attributes (`.weights` flat fp16-castable parameter with the reference layout and
its `torch.manual_seed(42)` U(+-sqrt(3/hidden)) init, `.padded_output_dim`, `.num_parameters`), same `ffmlp_forward`
Function alias and positional argument order, same autocast decorators (inputs and weights are cast to half).
Differences: the batch is not padded to a multiple of 128 (the kernels take any B; the observable output shape is the
same), no side streams are created (`allocate_splitk` is a no-op kept for API compatibility), and the fused backward
kernel produces the weight gradients itself instead of CUTLASS split-K launches (ffmlp.cu:795-887).
"""
import math

import torch
import torch.nn as nn
from torch.autograd import Function
from torch.amp import custom_bwd, custom_fwd

from .. import _lib


class _ffmlp_forward(Function):
    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.half)
    def forward(ctx, inputs, weights, input_dim, output_dim, hidden_dim, num_layers, activation, output_activation,
                inference=False, calc_grad_inputs=False):
        _lib.require_cuda(inputs, weights)
        if inputs.dtype != torch.half or weights.dtype != torch.half:  # no autocast region: the kernels are fp16
            inputs, weights = inputs.half(), weights.half()
        B = inputs.shape[0]
        inputs = inputs.contiguous()
        weights = weights.contiguous()
        outputs = torch.empty(B, output_dim, device=inputs.device, dtype=inputs.dtype)
        forward_buffer = None
        if not inference:
            forward_buffer = torch.empty(num_layers, B, hidden_dim, device=inputs.device, dtype=inputs.dtype)
        with torch.cuda.device(inputs.device):
            _lib.check(_lib.lib().ngp_ffmlp_forward(
                _lib.ptr(inputs), _lib.ptr(weights), B, input_dim, output_dim, hidden_dim, num_layers, activation,
                output_activation, _lib.ptr(forward_buffer), _lib.ptr(outputs), _lib.stream()), "ffmlp_forward")
        if not inference:
            ctx.save_for_backward(inputs, weights, outputs, forward_buffer)
            ctx.dims = (input_dim, output_dim, hidden_dim, num_layers, activation, output_activation, calc_grad_inputs)
        return outputs

    @staticmethod
    @custom_bwd(device_type="cuda")
    def backward(ctx, grad):
        B = grad.shape[0]
        grad = grad.contiguous().half()
        inputs, weights, outputs, forward_buffer = ctx.saved_tensors
        input_dim, output_dim, hidden_dim, num_layers, activation, output_activation, calc_grad_inputs = ctx.dims
        grad_inputs = torch.empty_like(inputs) if calc_grad_inputs else None
        grad_weights = torch.empty_like(weights)
        L = _lib.lib()
        ws_bytes = L.ngp_ffmlp_backward_workspace_bytes(B, input_dim, num_layers)
        workspace = torch.empty(ws_bytes, device=grad.device, dtype=torch.uint8)
        with torch.cuda.device(grad.device):
            _lib.check(L.ngp_ffmlp_backward(
                _lib.ptr(grad), _lib.ptr(inputs), _lib.ptr(weights), _lib.ptr(forward_buffer), B, input_dim, output_dim,
                hidden_dim, num_layers, activation, output_activation, int(calc_grad_inputs), None,
                _lib.ptr(grad_inputs), _lib.ptr(grad_weights), _lib.ptr(workspace), ws_bytes, _lib.stream()),
                "ffmlp_backward")
        return grad_inputs, grad_weights, None, None, None, None, None, None, None, None


ffmlp_forward = _ffmlp_forward.apply


def convert_activation(act):
    """ffmlp.py:87-94."""
    return {"relu": 0, "exponential": 1, "sine": 2, "sigmoid": 3, "squareplus": 4, "softplus": 5}.get(act, 6)


class FFMLP(nn.Module):
    def __init__(self, input_dim, output_dim, hidden_dim, num_layers, activation="relu"):
        super().__init__()
        self.input_dim = input_dim
        self.output_dim = output_dim
        self.hidden_dim = hidden_dim
        self.num_layers = num_layers
        self.activation = convert_activation(activation)
        self.output_activation = convert_activation("none")
        self.tensorcore_width = 16
        assert hidden_dim in [16, 32, 64, 128, 256], f"FFMLP only support hidden_dim in [16, 32, 64, 128, 256], but got {hidden_dim}"
        assert input_dim > 0 and input_dim % 16 == 0, f"FFMLP input_dim should be 16 * m (m  > 0), but got {input_dim}"
        assert output_dim <= 16, f"FFMLP current only supports output dim <= 16, but got {output_dim}"
        assert num_layers >= 2, f"FFMLP num_layers should be larger than 2 (3 matmuls), but got {num_layers}"
        if hidden_dim != 64:
            raise NotImplementedError("the B200 FFMLP kernels implement the 64-wide network nerf-navigation uses")
        self.padded_output_dim = int(math.ceil(output_dim / 16)) * 16
        self.num_parameters = hidden_dim * (input_dim + hidden_dim * (num_layers - 1) + self.padded_output_dim)
        self.weights = nn.Parameter(torch.zeros(self.num_parameters))
        self.reset_parameters()

    def cleanup(self):
        pass  # no side streams to destroy (ffmlp.py:130-132)

    def __repr__(self):
        return (f"FFMLP: input_dim={self.input_dim} output_dim={self.output_dim} hidden_dim={self.hidden_dim} "
                f"num_layers={self.num_layers} activation={self.activation}")

    def reset_parameters(self):
        torch.manual_seed(42)
        std = math.sqrt(3 / self.hidden_dim)
        self.weights.data.uniform_(-std, std)

    def forward(self, inputs):
        """inputs [B, input_dim] -> [B, output_dim] (fp16)."""
        outputs = ffmlp_forward(inputs, self.weights, self.input_dim, self.padded_output_dim, self.hidden_dim,
                                self.num_layers, self.activation, self.output_activation, not self.training,
                                inputs.requires_grad)
        if self.padded_output_dim != self.output_dim:
            outputs = outputs[:, :self.output_dim]
        return outputs
