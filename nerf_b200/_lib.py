"""ctypes binding of libnerf_b200.so (include/nerf_b200.h).  There is no fallback:
a missing library or a non-zero status raises."""
import ctypes
import os

import torch

_PKG = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_PKG, "libnerf_b200.so")
_lib = None

c_f = ctypes.c_float
c_i = ctypes.c_int
c_l = ctypes.c_int64
c_u64 = ctypes.c_uint64
c_p = ctypes.c_void_p

# name -> (restype, argtypes); must list every symbol declared in include/nerf_b200.h
SIGNATURES = {
    "nerf_abi_version": (c_i, []),
    "nerf_last_error": (ctypes.c_char_p, []),
    "nerf_generate_rays": (c_i, [c_p, c_p, c_i, c_i, c_i, c_f, c_l, c_l, c_p, c_p, c_p]),
    "nerf_gather_pixel_rays": (c_i, [c_p, c_p, c_i, c_p, c_i, c_i, c_i, c_i, c_i, c_f, c_p, c_l,
                                     c_p, c_p, c_p, c_p, c_p, c_p]),
    "nerf_sample_linear": (c_i, [c_f, c_f, c_p, c_i, c_l, c_p, c_i, c_u64, c_l, c_p, c_p]),
    "nerf_sample_geometric": (c_i, [c_f, c_p, c_i, c_l, c_p, c_i, c_u64, c_l, c_p, c_p]),
    "nerf_add_density_noise": (c_i, [c_p, c_i, c_l, c_f, c_u64, c_l, c_p]),
    "nerf_sample_pdf_merge": (c_i, [c_p, c_p, c_i, c_i, c_p, c_l, c_p, c_p]),
    "nerf_encode_g1": (c_i, [c_p, c_p, c_p, c_p, c_i, c_p, c_i, c_l, c_i, c_f, c_i, c_p, c_p, c_p, c_p]),
    "nerf_composite_fwd": (c_i, [c_p, c_p, c_p, c_p, c_p, c_p, c_l, c_i, c_p, c_p, c_p, c_p, c_p]),
    "nerf_composite_bwd": (c_i, [c_p, c_p, c_p, c_p, c_p, c_p, c_l, c_i, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p]),
    "nerf_g1_param_layout": (c_l, [c_i, c_i, ctypes.POINTER(c_l)]),
    "nerf_g1_mlp_workspace_bytes": (c_l, [c_l, c_i, c_i, c_i]),
    "nerf_g1_mlp_fwd": (c_i, [c_p, c_p, c_p, c_l, c_i, c_i, c_i, c_i, c_p, c_p, c_p, c_l, c_p]),
    "nerf_g1_mlp_bwd": (c_i, [c_p, c_p, c_p, c_l, c_i, c_i, c_i, c_p, c_p, c_p, c_p, c_l, c_p]),
    "nerf_g1_packed_bytes": (c_l, [c_i, c_i]),
    "nerf_g1_pack_weights": (c_i, [c_p, c_i, c_i, c_p, c_p]),
    "nerf_g1_render_workspace_bytes": (c_l, [c_l, c_i, c_i, c_l]),
    "nerf_g1_render_fused": (c_i, [c_p, c_i, c_i, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_f, c_l, c_l, c_i,
                                   c_f, c_f, c_p, c_p, c_p, c_i, c_u64, c_l, c_p, c_f, c_f, c_i, c_p, c_l, c_l,
                                   c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_l, c_p]),
    "nerf_g1_packed_x3_bytes": (c_l, []),
    "nerf_g1_pack_weights_x3": (c_i, [c_p, c_p, c_p]),
    "nerf_g1_render_fused_x3": (c_i, [c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_f, c_l, c_l, c_i,
                                      c_f, c_f, c_p, c_p, c_p, c_i, c_u64, c_l, c_f, c_f, c_i,
                                      c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_l, c_p]),
    "nerf_g1_render_static_max_points": (c_l, []),
    "nerf_g1_render_fused_static": (c_i, [c_p, c_p, c_p, c_p, c_p, c_i, c_i, c_i, c_f, c_l, c_l, c_i,
                                          c_f, c_f, c_p, c_p, c_i, c_u64, c_l, c_f, c_f, c_i, c_p, c_p, c_p, c_p]),
    "nerf_g1_mlp_tc_debug": (c_i, [c_p, c_p, c_p, c_p, c_l, c_i, c_f, c_i, c_p, c_p, c_p, c_p, c_i, c_p]),
    "nerf_tc_wgrad_debug": (c_i, [c_p, c_p, c_p, c_p, c_i, c_i, c_p, c_i, c_p]),
    "nerf_g1_train_workspace_bytes": (c_l, [c_l, c_i, c_i]),
    "nerf_g1_packed_bwd_bytes": (c_l, [c_i, c_i]),
    "nerf_g1_pack_weights_bwd": (c_i, [c_p, c_i, c_i, c_p, c_p]),
    "nerf_g1_train_fwd_tc": (c_i, [c_p, c_i, c_i, c_p, c_p, c_l, c_i, c_f, c_f, c_p, c_p, c_p, c_i, c_u64, c_l,
                                   c_f, c_i, c_p, c_p, c_p, c_p, c_p, c_l, c_p]),
    "nerf_g1_train_bwd_tc": (c_i, [c_p, c_p, c_p, c_i, c_i, c_p, c_p, c_l, c_i, c_p, c_p, c_l, c_i, c_p]),
    "nerf_g1_packed_bwd_x3_bytes": (c_l, []),
    "nerf_g1_pack_weights_bwd_x3": (c_i, [c_p, c_p, c_p]),
    "nerf_g1_train_x3_workspace_bytes": (c_l, [c_l, c_i]),
    "nerf_g1_train_fwd_x3": (c_i, [c_p, c_p, c_p, c_l, c_i, c_f, c_f, c_p, c_p, c_p, c_i, c_u64, c_l, c_f, c_i,
                                   c_p, c_p, c_p, c_p, c_l, c_p]),
    "nerf_g1_train_bwd_x3": (c_i, [c_p, c_p, c_p, c_p, c_p, c_l, c_i, c_p, c_p, c_l, c_i, c_p]),
    "nerf_head_encode": (c_i, [c_p, c_p, c_p, c_l, c_i, c_f, c_i, c_i, c_p, c_p, c_p, c_p]),
    "nerf_head_mlp_workspace_bytes": (c_l, [c_l]),
    "nerf_head_mlp_fwd": (c_i, [c_p, c_p, c_l, c_i, c_i, c_p, c_p, c_l, c_p]),
    "nerf_head_packed_bytes": (c_l, []),
    "nerf_head_pack_weights": (c_i, [c_p, c_i, c_i, c_p, c_p]),
    "nerf_head_fwd_tc": (c_i, [c_p, c_p, c_p, c_p, c_l, c_i, c_f, c_i, c_i, c_i, c_i, c_p, c_p, c_p]),
    "nerf_head_composite": (c_i, [c_p, c_i, c_i, c_p, c_p, c_f, c_l, c_i, c_p, c_p, c_p, c_i, c_p]),
    "nerf_head_mlp_train_workspace_bytes": (c_l, [c_l]),
    "nerf_head_mlp_fwd_train": (c_i, [c_p, c_p, c_l, c_i, c_i, c_p, c_p, c_l, c_p]),
    "nerf_head_composite_bwd": (c_i, [c_p, c_i, c_i, c_p, c_p, c_f, c_l, c_i, c_p, c_p, c_p, c_p, c_i, c_p]),
    "nerf_head_mlp_bwd": (c_i, [c_p, c_p, c_l, c_i, c_i, c_p, c_i, c_p, c_p, c_l, c_p]),
    "nerf_adam_step": (c_i, [c_p, c_p, c_p, c_p, c_l, c_l, c_f, c_f, c_f, c_f, c_f, c_p]),
    "nerf_dp_allreduce_adam": (c_i, [c_p, c_p, c_p, c_l, c_l, c_i, c_i, c_u64, c_p, c_p, c_p, c_l, c_l, c_f, c_f, c_f, c_f, c_f, c_p]),
    "nerf_mse_loss": (c_i, [c_p, c_p, c_l, c_p, c_p, c_p]),
    "nerf_composite_mse_train": (c_i, [c_p, c_p, c_p, c_p, c_p, c_p, c_l, c_i, c_f, c_u64, c_l, c_p, c_p, c_p, c_p, c_p]),
}


class NerfB200Error(RuntimeError):
    pass


def library_path():
    return _SO


def lib():
    """Loads the library once.  Raises if it was not built (no CPU fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise NerfB200Error(
                f"{_SO} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nerf_b200 has no CPU or PyTorch fallback)")
        handle = ctypes.CDLL(_SO)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)          # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(status, what):
    if status == 0:
        return
    if status < 0:
        names = {-1: "bad argument", -2: "unsupported configuration", -3: "workspace too small"}
        raise ValueError(f"{what}: {names.get(status, status)}")
    raise NerfB200Error(f"{what}: CUDA error {status} ({lib().nerf_last_error().decode()})")


def ptr(t):
    """Device pointer of a contiguous CUDA fp32/int64 tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise NerfB200Error("nerf_b200 kernels need CUDA tensors (there is no CPU path)")
    if not t.is_contiguous():
        raise ValueError("tensor must be contiguous")
    return t.data_ptr()


def stream():
    return torch.cuda.current_stream().cuda_stream
