"""Table gradient of the hash-grid encoder: deterministic sorted segmented reduction (csrc/tablegrad.cu).
Replaces the reference's atomicAdd scatter (gridencoder.cu:227-314) behind grid.py:64-87.

DETERMINISTIC = False switches to the reference's own method (float atomics, sums in arrival order: equal to the
deterministic result up to fp32 summation order, not bit-reproducible) for throughput."""
import torch

from .. import _lib

DETERMINISTIC = True

def grid_backward_table(grad, inputs, embeddings, offsets, dims, align_corners, grad_layout=1):
    """grad [B, L*C] (layout 1) or [L,B,C] (layout 0), same dtype as embeddings -> grad_embeddings like embeddings."""
    B, D, C, L, S, H, gridtype = dims
    L_ = _lib.lib()
    if not DETERMINISTIC:
        out = torch.empty(embeddings.shape, dtype=torch.float32, device=inputs.device)
        _lib.check(L_.ngp_grid_encode_backward_table_atomic(
            _lib.ptr(grad), grad_layout, 0 if grad.dtype == torch.float32 else 1, _lib.ptr(inputs), _lib.ptr(offsets),
            int(embeddings.shape[0]), _lib.ptr(out), B, D, C, L, S, H, gridtype, int(align_corners), _lib.stream()),
            "grid_encode_backward(table, atomic)")
        return out if embeddings.dtype == torch.float32 else out.to(embeddings.dtype)
    ws_bytes = L_.ngp_grid_backward_table_workspace_bytes(B, D, L)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=inputs.device)
    out = torch.empty(embeddings.shape, dtype=torch.float32, device=inputs.device)
    _lib.check(L_.ngp_grid_encode_backward_table(
        _lib.ptr(grad), grad_layout, 0 if grad.dtype == torch.float32 else 1, _lib.ptr(inputs), _lib.ptr(offsets),
        int(embeddings.shape[0]), _lib.ptr(out), B, D, C, L, S, H, gridtype, int(align_corners), _lib.ptr(ws), ws_bytes,
        _lib.stream()), "grid_encode_backward(table)")
    return out if embeddings.dtype == torch.float32 else out.to(embeddings.dtype)
