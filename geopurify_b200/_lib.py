"""ctypes binding of libgeopurify_b200.so (the C ABI declared in include/geopurify_b200.h).

There is no CPU fallback: if the library is missing or a call fails, an exception is raised.
PyTorch is used only for device memory and streams.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_int64, c_size_t, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GEOPURIFY_B200_LIB") or os.path.join(_HERE, "libgeopurify_b200.so")  # env: A/B builds

GP_F32, GP_F64, GP_BF16 = 1, 2, 3


class GeoPurifyError(RuntimeError):
    pass


class StudentDesc(ctypes.Structure):
    _fields_ = [("n_layers", c_int), ("kv", c_int * 16), ("cin", c_int * 16), ("cout", c_int * 16),
                ("bn", c_int * 16), ("relu", c_int * 16), ("residual_from", c_int * 16),
                ("w_off", c_int64 * 16), ("bn_off", c_int64 * 16)]


_P = c_void_p
_SIGS = {
    "gp_version": (c_int, []),
    "gp_last_error": (c_char_p, []),
    "gp_timing_enable": (c_int, [c_int]),
    "gp_timing_read": (c_int, [c_int, POINTER(c_double), POINTER(c_int)]),
    "gp_project_views": (c_int, [_P, c_int, c_int64, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_double, _P, _P, _P]),
    "gp_pix_to_mapping": (c_int, [_P, c_int64, c_int, _P, _P]),
    "gp_splat_depth": (c_int, [_P, c_int, c_int64, _P, _P, c_int, c_int, c_int, c_int, _P, _P]),
    "gp_accept_views": (c_int, [_P, c_int, c_int, c_int, _P, _P]),
    "gp_mask_partition": (c_int, [_P, _P, c_int, c_int, c_int, c_int, c_int, c_float, _P, _P, _P]),
    "gp_visible_csr_workspace_bytes": (c_size_t, [c_int64, c_int]),
    "gp_visible_csr": (c_int, [_P, _P, c_int64, c_int, c_int, _P, _P, _P, _P, _P, _P, c_size_t, _P]),
    "gp_lift_mean_fused": (c_int, [_P, c_int, c_int64, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_double, _P, _P,
                                   c_int, c_int, _P, c_int64, _P, _P]),
    "gp_lift_mean_idmap": (c_int, [_P, c_int, c_int64, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_double, _P, _P,
                                   _P, c_int, c_int, c_int, _P, c_int64, _P, _P]),
    "gp_expand_idmap": (c_int, [_P, _P, c_int, c_int, c_int, c_int, c_int, _P, c_int, _P]),
    "gp_lift_mean_pix": (c_int, [_P, c_int64, c_int, c_int, c_int, _P, _P, c_int, c_int, _P, c_int64, _P, _P]),
    "gp_lift_consensus_workspace_bytes": (c_size_t, [c_int64, c_int, c_int, c_int, c_int]),
    "gp_lift_consensus": (c_int, [_P, c_int64, c_int, c_int, c_int, _P, _P, _P, c_int, c_int, _P, c_int, c_float, _P,
                                  c_float, _P, c_int64, _P, _P, c_size_t, _P]),
    "gp_nn1_fill_workspace_bytes": (c_size_t, [c_int64]),
    "gp_nn1_fill": (c_int, [_P, c_int64, _P, c_float, _P, _P, c_size_t, _P]),
    "gp_copy_rows": (c_int, [_P, c_int64, c_int, c_int64, _P, _P]),
    "gp_voxelize_workspace_bytes": (c_size_t, [c_int64]),
    "gp_voxelize": (c_int, [_P, c_int, c_int64, _P, _P, _P, _P, _P, _P, _P, c_size_t, _P]),
    "gp_voxel_index_workspace_bytes": (c_size_t, [c_int64, c_int, c_int, c_int]),
    "gp_voxel_index_build": (c_int, [_P, c_int64, c_int, c_int, c_int, _P, _P, _P, c_size_t, _P]),
    "gp_voxel_index_check": (c_int, [_P, c_int64, c_int, c_int, c_int, _P]),
    "gp_kernel_map": (c_int, [_P, c_int64, c_int, c_int, c_int, _P, _P]),
    "gp_knn_grid": (c_int, [_P, c_int64, c_int, c_int, c_int, c_int, _P, _P, _P]),
    "gp_knn_points_workspace_bytes": (c_size_t, [c_int64]),
    "gp_knn_points": (c_int, [_P, c_int64, c_int, c_float, _P, _P, c_size_t, _P]),
    "gp_mine_workspace_bytes": (c_size_t, [c_int64, c_int, c_int, c_int]),
    "gp_mine_contrastive": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, c_int, c_int, c_int, _P, _P, _P,
                                    c_size_t, _P]),
    "gp_scatter_mean_workspace_bytes": (c_size_t, [c_int64, c_int64]),
    "gp_scatter_mean_prepare": (c_int, [_P, c_int64, c_int64, _P, c_size_t, _P]),
    "gp_scatter_mean": (c_int, [_P, c_int64, c_int, _P, c_int64, c_int64, _P, c_int64, c_int, _P]),
    "gp_scatter_mean_static": (c_int, [_P, c_int64, c_int64, _P, c_int64, c_int, _P, c_int64, c_int, _P]),
    "gp_student_workspace_bytes": (c_size_t, [POINTER(StudentDesc), c_int64]),
    "gp_student_forward": (c_int, [POINTER(StudentDesc), _P, _P, c_int64, _P, c_int64, _P, c_int, _P, c_size_t, _P]),
    "gp_student_tc_pack_bytes": (c_size_t, [POINTER(StudentDesc)]),
    "gp_student_tc_pack": (c_int, [POINTER(StudentDesc), _P, _P, _P]),
    "gp_student_tc_workspace_bytes": (c_size_t, [POINTER(StudentDesc), c_int64]),
    "gp_student_forward_tc": (c_int, [POINTER(StudentDesc), _P, _P, _P, c_int64, _P, c_int64, _P, c_int, _P, c_size_t, _P]),
    "gp_affinity_softmax": (c_int, [_P, c_int, _P, c_int64, c_int, c_float, _P, _P]),
    "gp_pool": (c_int, [_P, c_int64, c_int, _P, _P, c_int64, c_int, c_int, _P, _P, POINTER(c_void_p), _P]),
    "gp_pool_plan_bytes": (c_size_t, [c_int64, c_int]),
    "gp_pool_plan": (c_int, [_P, _P, c_int64, c_int, _P, c_size_t, _P]),
    "gp_pool_tc_workspace_bytes": (c_size_t, [c_int64]),
    "gp_pool_tc": (c_int, [_P, c_int64, c_int, _P, _P, _P, c_int64, c_int, c_int, _P, _P, c_size_t, _P]),
    "gp_gather_rows": (c_int, [_P, c_int64, c_int, _P, c_int64, _P, _P]),
    "gp_logits_workspace_bytes": (c_size_t, [c_int, c_int]),
    "gp_logits_argmax": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, c_float, _P, _P, _P, _P, c_size_t, _P]),
    "gp_logits_tc_workspace_bytes": (c_size_t, [c_int64, c_int, c_int]),
    "gp_logits_argmax_tc": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, c_float, _P, _P, _P, c_size_t, _P]),
    "gp_iou_hist": (c_int, [_P, _P, _P, c_int64, _P, c_int, c_int, _P, _P, _P]),
}

EXPORTS = tuple(_SIGS)
_lib = None


def load():
    """Load the shared library (once).  Raises GeoPurifyError if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise GeoPurifyError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C geopurify_b200/csrc`.  geopurify_b200 has no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in _SIGS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        msg = load().gp_last_error().decode("utf-8", "replace")
        raise GeoPurifyError(f"libgeopurify_b200 error {rc}: {msg}")


def ptr(t, dtype=None):
    """Device pointer of a CUDA tensor (None -> NULL).  Tensors must be contiguous; with `dtype` the element type
    is checked too (the C ABI takes untyped pointers: an int64 tensor where int32 is expected would be reinterpreted)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise GeoPurifyError("geopurify_b200 kernels need CUDA tensors (no CPU fallback)")
    if not t.is_contiguous():
        raise GeoPurifyError("tensor must be contiguous")
    if dtype is not None and t.dtype != dtype:
        raise GeoPurifyError(f"tensor dtype {t.dtype} where the C ABI expects {dtype}")
    return c_void_p(t.data_ptr())


def stream():
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def require_cuda():
    if not torch.cuda.is_available():
        raise GeoPurifyError("geopurify_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    load()


def dtype_code(t):
    return {torch.float32: GP_F32, torch.float64: GP_F64, torch.bfloat16: GP_BF16}[t.dtype]


launch_count = 0  # kernels launched through this binding (bench.py's gpu_launches claim)


def count(n):
    global launch_count
    launch_count += n
