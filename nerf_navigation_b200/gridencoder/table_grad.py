"""Table gradient of the hash-grid encoder (csrc/tablegrad.cu).  Replaces the reference's atomicAdd scatter
(gridencoder.cu:227-314) behind grid.py:64-87.  Three methods:

  DETERMINISTIC = True, METHOD = "fixed"   (default) fixed-point integer atomics: bit-reproducible, no sort
  DETERMINISTIC = True, METHOD = "sorted"  stable sort of (row, contribution) + segmented reduction: bit-reproducible
  DETERMINISTIC = False                    the reference's own method (float atomics, sums in arrival order: equal to the
                                           others up to fp32 summation order, not bit-reproducible)"""
import torch

from .. import _lib

DETERMINISTIC = True
METHOD = "fixed"

# training.GradBucket registers, per embeddings tensor (by data_ptr), the slice of its flat gradient bucket that the kernels
# should write into, so the table gradient is produced in place where the all-reduce and the optimiser read it
_OUT = {}


def set_output(embeddings, out):
    """`out`: fp32 tensor of embeddings' shape (a view of a flat bucket) or None to remove."""
    if out is None:
        _OUT.pop(embeddings.data_ptr(), None)
    else:
        _OUT[embeddings.data_ptr()] = out


def _out_for(embeddings, device):
    out = _OUT.get(embeddings.data_ptr())
    if out is not None and out.shape == embeddings.shape and out.device == device:
        return out.detach()   # a fresh alias: autograd adopts it as .grad without copying (the registered view stays ours)
    return torch.empty(embeddings.shape, dtype=torch.float32, device=device)


def grid_backward_table(grad, inputs, embeddings, offsets, dims, align_corners, grad_layout=1):
    """grad [B, L*C] (layout 1) or [L,B,C] (layout 0), same dtype as embeddings -> grad_embeddings like embeddings."""
    B, D, C, L, S, H, gridtype = dims
    L_ = _lib.lib()
    if not DETERMINISTIC:
        out = _out_for(embeddings, inputs.device)
        _lib.check(L_.ngp_grid_encode_backward_table_atomic(
            _lib.ptr(grad), grad_layout, 0 if grad.dtype == torch.float32 else 1, _lib.ptr(inputs), _lib.ptr(offsets),
            int(embeddings.shape[0]), _lib.ptr(out), B, D, C, L, S, H, gridtype, int(align_corners), _lib.stream()),
            "grid_encode_backward(table, atomic)")
        return out if embeddings.dtype == torch.float32 else out.to(embeddings.dtype)
    if METHOD == "fixed":
        ws_bytes = L_.ngp_grid_backward_table_fixed_workspace_bytes(int(embeddings.shape[0]), C)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=inputs.device)
        out = _out_for(embeddings, inputs.device)
        _lib.check(L_.ngp_grid_encode_backward_table_fixed(
            _lib.ptr(grad), grad_layout, 0 if grad.dtype == torch.float32 else 1, _lib.ptr(inputs), _lib.ptr(offsets),
            int(embeddings.shape[0]), _lib.ptr(out), B, D, C, L, S, H, gridtype, int(align_corners), _lib.ptr(ws), ws_bytes,
            _lib.stream()), "grid_encode_backward(table, fixed-point)")
        return out if embeddings.dtype == torch.float32 else out.to(embeddings.dtype)
    ws_bytes = L_.ngp_grid_backward_table_workspace_bytes(B, D, L)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=inputs.device)
    out = _out_for(embeddings, inputs.device)
    _lib.check(L_.ngp_grid_encode_backward_table(
        _lib.ptr(grad), grad_layout, 0 if grad.dtype == torch.float32 else 1, _lib.ptr(inputs), _lib.ptr(offsets),
        int(embeddings.shape[0]), _lib.ptr(out), B, D, C, L, S, H, gridtype, int(align_corners), _lib.ptr(ws), ws_bytes,
        _lib.stream()), "grid_encode_backward(table)")
    return out if embeddings.dtype == torch.float32 else out.to(embeddings.dtype)
