"""ctypes binding of libapb200.so (include/apb200.h).

The product path has NO fallback: if the CUDA library is missing or a call returns a non-zero
status this module raises.  Tensors are passed as raw device pointers plus the current torch CUDA
stream, mirroring how the reference's pybind wrappers receive pre-allocated tensors
(openpoints/cpp/pointnet2_batch/src/pointnet2_api.cpp:10-24).
"""
import ctypes
import os

import torch

_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libapb200.so")
_lib = None

_i, _f, _p = ctypes.c_int, ctypes.c_float, ctypes.c_void_p
_d, _ll = ctypes.c_double, ctypes.c_longlong

# name -> argtypes (all return int status unless noted)
_SIGNATURES = {
    "apb_fps_f32": [_p, _p, _p, _i, _i, _i, _p],
    "apb_tune_group": [_i, _i],
    "apb_tune_wgrad": [_i, _i],
    "apb_tune_sa2": [_i],
    "apb_tune_group_grad": [_i],
    "apb_tune_group_seg": [_i],
    "apb_gather_f32": [_p, _p, _p, _i, _i, _i, _i, _p],
    "apb_gather_grad_f32": [_p, _p, _p, _i, _i, _i, _i, _p],
    "apb_ball_query_f32": [_p, _p, _p, _i, _i, _i, _f, _i, _p],
    "apb_ball_query_grid_f32": [_p, _p, _p, _p, _ll, _i, _i, _i, _f, _i, _p],
    "apb_ball_query_grid_static_f32": [_p, _p, _p, _i, _i, _i, _f, _i, _p],
    "apb_knn_f32": [_p, _p, _p, _p, _i, _i, _i, _i, _p],
    "apb_group_f32": [_p, _p, _p, _i, _i, _i, _i, _i, _p],
    "apb_group_bf16": [_p, _p, _p, _i, _i, _i, _i, _i, _p],
    "apb_group_minmax_f32": [_p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p],
    "apb_group_grad_f32": [_p, _p, _p, _i, _i, _i, _i, _i, _p],
    "apb_three_nn_f32": [_p, _p, _p, _p, _i, _i, _i, _p],
    "apb_three_interp_f32": [_p, _p, _p, _p, _i, _i, _i, _i, _p],
    "apb_three_interp_grad_f32": [_p, _p, _p, _p, _i, _i, _i, _i, _p],
    "apb_deform_mask_fwd_f32": [_p, _p, _p, _p, _p, _f, _p, _p, _i, _i, _i, _p],
    "apb_deform_mask_bwd_f32": [_p, _p, _p, _p, _p, _f, _p, _p, _p, _p, _i, _i, _i, _p],
    "apb_pwconv_fwd_bf16": [_p, _p, _p, _p, _i, _p, _i, _i, _i, _i, _i, _p],
    "apb_pwconv_fwd_tf32": [_p, _p, _p, _p, _i, _p, _i, _i, _i, _i, _i, _p],
    "apb_pwconv_fwd_tf32_rows": [_p, _p, _p, _p, _i, _p, _i, _i, _i, _i, _i, _p],
    "apb_pwconv_ex_tf32": [_p, _p, _p, _p, _i, _p, _i, _p, _p, _i, _i, _i, _i, _i, _p],
    "apb_pmax_decode": [_p, _p, _p, _ll, _p],
    "apb_pwconv_wgrad_tf32": [_p, _p, _p, _i, _i, _i, _i, _p],
    "apb_pwconv_wgrad_ld_tf32": [_p, _p, _p, _i, _i, _i, _i, _i, _p],
    "apb_affine_f32": [_p, _p, _p, _p, _i, _p, _i, _i, _ll, _p],
    "apb_bn_stats_f32": [_p, _p, _i, _i, _ll, _p],
    "apb_bn_apply_f32": [_p, _p, _p, _p, _f, _d, _i, _p, _p, _p, _p, _p, _f, _i, _i, _ll, _p],
    "apb_bn_bwd_reduce_f32": [_p, _p, _p, _p, _p, _p, _i, _i, _ll, _p],
    "apb_bn_bwd_apply_f32": [_p, _p, _p, _p, _p, _p, _p, _d, _p, _i, _i, _ll, _p],
    "apb_kmax_stats_f32": [_p, _p, _p, _p, _p, _i, _i, _i, _i, _p],
    "apb_attn_fwd_f32": [_p, _p, _p, _p, _p, _i, _i, _i, _i, _p],
    "apb_attn_bwd_f32": [_p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p],
    "apb_spectral_norm_fwd_f32": [_p, _p, _p, _p, _p, _p, _p, _p, _p, _f, _i, _i, _p],
    "apb_spectral_norm_bwd_f32": [_p, _p, _p, _p, _p, _p, _p, _i, _i, _p],
    "apb_bn_fold_f32": [_p, _p, _p, _p, _f, _p, _p, _p, _i, _p],
    "apb_sums_split_f32": [_p, _p, _p, _i, _p],
    "apb_bn_finalize_f32": [_p, _d, _f, _f, _p, _p, _p, _p, _p, _p, _p, _p, _i, _p],
    "apb_sa2_fwd_tf32": [_p] * 7 + [_i, _p, _p, _p, _p, _p] + [_i] * 7 + [_p],
    "apb_sa2_bwd_rows_tf32": [_p] * 7 + [_i, _p, _p, _p, _i, _i] + [_p] * 9 + [_i] * 6 + [_p],
    "apb_sa2_wgrad_tf32": [_p] * 7 + [_i, _p, _p, _p, _i] + [_i] * 6 + [_p],
    "apb_bnmax_finalize_f32": [_p, _d, _f, _f, _p, _p, _p, _p, _p, _p, _p, _i, _p, _p, _p, _p, _p, _p, _i, _i, _i, _p],
    "apb_bnmax_bwd_reduce_f32": [_p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _p],
    "apb_bnmax_bwd_coef_f32": [_p, _p, _p, _p, _p, _d, _p, _p, _p, _p, _i, _p],
    "apb_group_ystats_f32": [_p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p],
    "apb_sa1_bwd_count_f32": [_p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p],
    "apb_sa1_bwd_sparse_f32": [_p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p],
    "apb_sa1_bwd_dense_f32": [_p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _p],
    "apb_sa_ustats_f32": [_p, _p, _p, _p, _p, _p, _i, _i, _i, _p],
    "apb_attn_small_fwd_f32": [_p, _p, _p, _p, _p, _i, _i, _i, _i, _p],
    "apb_attn_small_bwd_f32": [_p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p],
    "apb_rowmax_f32": [_p, _p, _p, _ll, _i, _p],
    "apb_bnmax_bwd_f32": [_p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p],
    "apb_peer_ws_create": [_p, _p],
    "apb_peer_ws_open": [_p, _p],
    "apb_peer_ws_close": [_p, _i],
    "apb_peer_allreduce": [_p, _i, _i, _p, _i, _i, _p],
    "apb_sa_fused_fwd_bf16": [_p, _p, _p, _p, _p, _p, _p, _p, _i, _p, _p, _p, _i, _p, _p,
                              _i, _i, _i, _i, _i, _i, _i, _i, _i, _p],
}
EXPORTED_SYMBOLS = sorted(list(_SIGNATURES) + ["apb_version", "apb_error_string", "apb_fps_block_size",
                                                  "apb_peer_ws_bytes", "apb_peer_max_elems", "apb_ball_query_grid_ws_bytes", "apb_ball_query_grid_static_ws_bytes", "apb_tune_wgrad"])


class ApbError(RuntimeError):
    pass


def lib_path():
    return _LIB_PATH


def load():
    """Load libapb200.so once; raise loudly if it has not been built (no CPU fallback exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise ApbError(
            f"{_LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(adaptpoint_b200 has no CPU or PyTorch fallback path)")
    lib = ctypes.CDLL(_LIB_PATH)
    for name, argtypes in _SIGNATURES.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = _i
    lib.apb_version.restype = ctypes.c_char_p
    lib.apb_version.argtypes = []
    lib.apb_error_string.restype = ctypes.c_char_p
    lib.apb_error_string.argtypes = [_i]
    lib.apb_fps_block_size.restype = _i
    lib.apb_fps_block_size.argtypes = [_i]
    lib.apb_ball_query_grid_ws_bytes.restype = _ll
    lib.apb_ball_query_grid_ws_bytes.argtypes = [_i, _i]
    lib.apb_ball_query_grid_static_ws_bytes.restype = _ll
    lib.apb_ball_query_grid_static_ws_bytes.argtypes = []
    for name in ("apb_peer_ws_bytes", "apb_peer_max_elems"):
        getattr(lib, name).restype = _i
        getattr(lib, name).argtypes = []
    _lib = lib
    return lib


def version():
    return load().apb_version().decode()


def tune_group(channels_per_cta=0, threads=0):
    """Benchmarks only: override the grouping/gather forward work split (0, 0 = automatic)."""
    status = load().apb_tune_group(int(channels_per_cta), int(threads))
    if status != 0:
        raise ApbError(f"apb_tune_group({channels_per_cta}, {threads}) rejected")


def tune_group_grad(mode=0):
    """Benchmarks only: grouping / three_interpolate backward path (0 automatic, 1 L2 reductions, 2 shared rows)."""
    status = load().apb_tune_group_grad(int(mode))
    if status != 0:
        raise ApbError(f"apb_tune_group_grad({mode}) rejected")


def tune_group_seg(kb=0):
    """Benchmarks only: slice size (KB) of the sliced grouping kernels (rows longer than shared memory)."""
    status = load().apb_tune_group_seg(int(kb))
    if status != 0:
        raise ApbError(f"apb_tune_group_seg({kb}) rejected")


def fps_block_size(n):
    return load().apb_fps_block_size(int(n))


def _ptr(t):
    if t is None:
        return None
    if not t.is_cuda:
        raise ApbError("adaptpoint_b200 ops need CUDA tensors (there is no CPU path)")
    return ctypes.c_void_p(t.data_ptr())


# Launch accounting for bench.py: LAUNCHES counts C-ABI calls (one kernel launch each); when TRACE is a list,
# every call is bracketed by CUDA events on the launching stream: (name, int args, start event, end event).
LAUNCHES = 0
TRACE = None


def call(name, tensors_and_scalars, device):
    """Invoke one C-ABI entry point on `device`'s current stream and raise on a non-zero status."""
    global LAUNCHES
    lib = load()
    with torch.cuda.device(device):
        ts = torch.cuda.current_stream(device)
        stream = ctypes.c_void_p(ts.cuda_stream)
        args = [(_ptr(a) if (a is None or isinstance(a, torch.Tensor)) else a) for a in tensors_and_scalars]
        LAUNCHES += 1
        if TRACE is not None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(ts)
            status = getattr(lib, name)(*args, stream)
            e1.record(ts)
            TRACE.append((name, tuple(a for a in tensors_and_scalars if isinstance(a, int)), e0, e1))
        else:
            status = getattr(lib, name)(*args, stream)
    if status != 0:
        raise ApbError(f"{name} failed: {lib.apb_error_string(status).decode()} (status {status})")
