"""ctypes binding of libngp_b200.so (the C ABI in include/ngp_b200.h).

There is no CPU fallback: if the library is missing or a call fails, a RuntimeError is raised
(the reference raises RuntimeError via TORCH_CHECK, e.g. gridencoder.cu:15-18).
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("NGP_B200_LIB") or os.path.join(_HERE, "libngp_b200.so")  # env override: dev experiments only

_lib = None

P = C.c_void_p
I = C.c_int
I64 = C.c_int64
U32 = C.c_uint32
F = C.c_float

_SIGNATURES = {
    "ngp_abi_version": ([], I),
    "ngp_grid_level_scales": ([P, I, F, U32, P], I),
    "ngp_grid_encode_forward": ([P, P, I, P, P, I, I64, I, I, I, F, U32, I, P, U32, I, P, P], I),
    "ngp_grid_encode_backward_input": ([P, I, P, P, I, P, P, I64, I, I, I, F, U32, U32, I, P], I),
    "ngp_grid_backward_table_workspace_bytes": ([I64, I, I], I64),
    "ngp_grid_encode_backward_table": ([P, I, I, P, P, I64, P, I64, I, I, I, F, U32, U32, I, P, I64, P], I),
    "ngp_linear_small": ([P, P, P, I64, I, I, I, P], I),
    "ngp_grid_backward_table_fixed_workspace_bytes": ([I64, I], I64),
    "ngp_grid_encode_backward_table_fixed": ([P, I, I, P, P, I64, P, I64, I, I, I, F, U32, U32, I, P, I64, P], I),
    "ngp_grid_encode_backward_table_atomic": ([P, I, I, P, P, I64, P, I64, I, I, I, F, U32, U32, I, P], I),
    "ngp_field_query_forward": ([P, F, P, P, F, U32, U32, I, P, P, P, P, P, I64, P], I),
    "ngp_field_query_value_grad": ([P, F, P, P, F, U32, U32, I, P, P, P, P, F, F, P, P, P, I64, P], I),
    "ngp_field_query_forward_jacobian": ([P, F, P, P, F, U32, U32, I, P, P, P, P, P, P, I64, P], I),
    "ngp_field_query_backward_jacobian": ([P, P, P, I64, P], I),
    "ngp_field_query_ff_forward": ([P, F, P, P, F, U32, U32, I, P, I, I, P, I64, P], I),
    "ngp_set_field_impl": ([I], I),
    "ngp_set_device_check": ([I], I),
    "ngp_field_color_forward": ([P, P, F, P, P, F, U32, U32, I, P, P, P, P, P, P, P, P, I64, P], I),
    "ngp_field_color_backward": ([P, P, F, P, P, F, U32, U32, I, P, P, P, P, P, P, P, P, P, P, P, P, I64, P], I),
    "ngp_near_far_from_aabb": ([P, P, P, I64, F, P, P, P], I),
    "ngp_morton3D": ([P, I64, P, P], I),
    "ngp_morton3D_invert": ([P, I64, P, P], I),
    "ngp_packbits": ([P, I64, F, P, P], I),
    "ngp_march_rays_train_workspace_bytes": ([I64], I64),
    "ngp_march_rays_train": ([P, P, P, F, F, U32, I64, U32, U32, I64, P, P, P, P, P, P, P, U32, P, I64, P], I),
    "ngp_march_rays_train_write": ([P, P, P, F, F, U32, I64, U32, U32, I64, P, P, P, P, P, P, U32, P, P], I),
    "ngp_march_rays_train_backward": ([P, P, P, P, P, P, P, F, U32, U32, I64, I64, P, P, P], I),
    "ngp_composite_rays_train_forward": ([P, P, P, P, I64, I64, P, P, P, P], I),
    "ngp_composite_rays_train_backward": ([P, P, P, P, P, P, P, P, I64, I64, P, P, P], I),
    "ngp_march_rays": ([I64, U32, P, P, P, P, F, F, U32, U32, U32, P, P, P, P, P, P, U32, P], I),
    "ngp_composite_rays": ([I64, U32, P, P, P, P, P, P, P, P, P], I),
    "ngp_pose_rays_forward": ([P, P, I, I, I, F, F, F, F, P, P, P], I),
    "ngp_pose_rays_backward": ([P, P, P, I, I, I, F, F, F, F, P, P], I),
    "ngp_render_rays_infer_workspace_bytes": ([], I64),
    "ngp_render_rays_infer": ([P, P, I64, P, F, P, F, F, U32, U32, U32, P, P, F, U32, U32, I, P, P, P, P, P, F, P, I, P, P,
                              P, P, P, P], I),
    "ngp_render_plan_workspace_bytes": ([I64], I64),
    "ngp_render_plan": ([P, P, I64, P, F, P, F, F, U32, U32, U32, P, P, P, I64, P], I),
    "ngp_render_rays_infer_planned": ([P, P, I64, P, F, P, F, F, U32, U32, U32, P, P, F, U32, U32, I, P, P, P, P, P, F, P, I, P, P,
                                      P, P, P, P, P, I64, P, P], I),
    "ngp_render_rays_infer_segments_workspace_bytes": ([I64, I], I64),
    "ngp_render_rays_infer_segments": ([P, P, I64, P, F, P, F, F, U32, U32, U32, P, P, F, U32, U32, I, P, P, P, P, P, F, P, I, P, P,
                                       P, P, I, P, I64, P], I),
    "ngp_state_to_pose_forward": ([P, I, F, F, P, P], I),
    "ngp_state_to_pose_backward": ([P, P, I, F, F, P, P], I),
    "ngp_adam_step": ([P, P, P, P, P, I64, F, F, F, F, I64, F, F, P, P, P], I),
    "ngp_linear_wgrad_workspace_bytes": ([I64], I64),
    "ngp_linear_wgrad": ([P, P, P, I64, I, I, P, I64, P], I),
    "ngp_linear_wgrad_static": ([P, P, P, I64, I, I, P], I),
    "ngp_occ_grid_positions": ([I, U32, F, F, U32, P, P], I),
    "ngp_occ_grid_update_blocks": ([I64], I64),
    "ngp_occ_grid_update": ([P, P, I64, F, F, P, P], I),
    "ngp_occ_grid_partial_workspace_bytes": ([U32], I64),
    "ngp_occ_grid_partial_select": ([P, I, U32, F, F, U32, U32, P, P, P, P], I),
    "ngp_occ_grid_scatter_update": ([P, P, P, U32, I64, F, F, P, P, P], I),
    "ngp_occ_grid_mark_untrained": ([P, I, F, F, F, F, U32, U32, F, P, P], I),
    "ngp_lattice_positions": ([P, P, U32, P, P], I),
    "ngp_maxpool_threshold": ([P, U32, U32, F, P, P], I),
    "ngp_sh_encode_forward": ([P, P, I64, I, I, I, P, P], I),
    "ngp_sh_encode_backward": ([P, P, I64, I, I, P, P], I),
    "ngp_ffmlp_forward": ([P, P, I64, I, I, I, I, I, I, P, P, P], I),
    "ngp_ffmlp_backward_workspace_bytes": ([I64, I, I], I64),
    "ngp_ffmlp_backward": ([P, P, P, P, I64, I, I, I, I, I, I, I, P, P, P, P, I64, P], I),
    "ngp_probe_gather": ([P, I64, I, I, I, I, P, P], I),
    "ngp_field_query_backward": ([P, F, P, P, F, U32, U32, I, P, P, P, P, P, P, P, I64, P], I),
}


def declared_symbols():
    """Every entry point include/ngp_b200.h declares (checked by the CPU test-suite)."""
    import re
    hdr = open(os.path.join(_HERE, "..", "include", "ngp_b200.h")).read()
    return sorted(set(re.findall(r"\b(ngp_[a-z0-9_]+)\s*\(", hdr)))


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise RuntimeError(f"{SO_PATH} is missing: build it with `python -m nerf_navigation_b200.build` "
                               "(there is no CPU fallback)")
        L = C.CDLL(SO_PATH)
        L.ngp_last_error.restype = C.c_char_p
        for name, (args, res) in _SIGNATURES.items():
            fn = getattr(L, name)
            fn.argtypes = args
            fn.restype = res
        _lib = L
    return _lib


def check(rc, what=""):
    if rc != 0:
        raise RuntimeError(f"{what}: {lib().ngp_last_error().decode()}")


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def stream():
    return torch.cuda.current_stream().cuda_stream


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("expected a CUDA tensor (nerf_navigation_b200 has no CPU path)")


def require_contiguous(*tensors):
    for t in tensors:
        if t is not None and not t.is_contiguous():
            raise RuntimeError("expected a contiguous tensor")
