"""ctypes binding of libb200mp.so (the C ABI declared in include/b200mp.h).

There is no CPU fallback: if the library is missing or a call fails, an exception is raised."""
from __future__ import annotations

import ctypes as C
import os
import re
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200mp.so")
HEADER = os.path.join(os.path.dirname(_HERE), "include", "b200mp.h")

_lib: Optional[C.CDLL] = None

_p, _i, _l, _f, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t
_SIGS = {
    "b200mp_version": [],
    "b200mp_partial_rows": [_p],
    "b200mp_partial_rows_for": [_i, _p],
    "b200mp_tile_edges": [_i, _p],
    "b200mp_csc_workspace_bytes": [_l, _l, _p],
    "b200mp_csc_build": [_p, _p, _l, _l, _p, _p, _p, _p, _p, _p, _sz, _p],
    "b200mp_degree": [_p, _l, _p, _p, _p],
    "b200mp_bincount": [_p, _l, _p, _l, _p],
    "b200mp_relabel_workspace_bytes": [_l, _p],
    "b200mp_relabel": [_p, _l, _p, _p, _p, _p, _sz, _p],
    "b200mp_relabel_static": [_p, _l, _p, _p, _p, _p],
    "b200mp_gather_rows": [_p, _p, _p, _l, _i, _p],
    "b200mp_scatter_rows": [_p, _p, _p, _l, _i, _p],
    "b200mp_edge_linear_fwd": [_i, _i, _l, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _f, _f, _p, _p, _p, _p],
    "b200mp_segment_reduce_fwd": [_i, _i, _l, _p, _p, _p, _p, _l, _l, _l, _p, _p, _l, _p, _p, _p, _p],
    "b200mp_message_bwd": [_i, _i, _l, _p, _p, _p, _p, _p, _p, _p, _p, _i, _p, _p, _p, _i, _p, _p],
    "b200mp_message_reduce_workspace": [_i, _l, _p, _p],
    "b200mp_message_fwd_reduce": [_i, _i, _l, _l, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p],
    "b200mp_pna_scale_fwd": [_i, _l, _p, _p, _p, _i, _p, _p, _p, _p, _p],
    "b200mp_pna_scale_bwd": [_i, _l, _p, _p, _i, _p, _p, _p, _p, _p],
    "b200mp_pna_coef": [_i, _l, _p, _p, _p, _p, _p, _p, _p],
    "b200mp_split_tf32": [_p, _p, _p, _l, _p],
    "b200mp_node_linear": [_p, _l, _i, _i, _p, _i, _p, _p, _p],
    "b200mp_node_linear_ld": [_p, _l, _i, _i, _p, _i, _p, _p, _i, _p],
    "b200mp_node_linear_wgrad": [_p, _i, _i, _p, _i, _i, _l, _p, _i, _p],
    "b200mp_node_linear_wgrad_rows": [_p, _p, _i, _i, _p, _p, _i, _i, _l, _p, _i, _p],
    "b200mp_node_linear64": [_p, _p, _p, _i, _l, _p, _i, _p, _p, _i, _p],
    "b200mp_node_linear64_colsum": [_p, _p, _p, _i, _l, _p, _p, _p, _p, _i, _p],
    "b200mp_node_pair_wgrad": [_p, _p, _p, _l, _p, _i, _p],
    "b200mp_negative_sample_workspace_bytes": [_l, _p],
    "b200mp_negative_sample": [_p, _l, _l, _p, _l, _i, C.c_uint64, _p, _p, _sz, _p],
    "b200mp_sample_hop_count": [_p, _p, _l, _l, _p, _p],
    "b200mp_sample_hop_fill": [_p, _p, _l, _l, _p, C.c_uint64, _p, _p],
    "b200mp_first_occurrence": [_p, _l, _p, _p, _p, _p],
    "b200mp_col_attention_fwd": [_p, _p, _p, _l, _i, _i, _i, _f, C.c_uint64, _p, _p],
    "b200mp_col_attention_bwd": [_p, _p, _p, _p, _l, _i, _i, _i, _f, C.c_uint64, _p, _p, _p, _p],
    "b200mp_bn_stats": [_p, _l, _i, _p, _p],
    "b200mp_bn_relu_res_fwd": [_p, _p, _l, _i, _p, _p, _p, _p, _p],
    "b200mp_bn_relu_res_bwd_reduce": [_p, _p, _l, _i, _p, _p, _p, _p, _p],
    "b200mp_bn_relu_res_bwd_apply": [_p, _p, _l, _i, _p, _p, _p, _p, _p, _p, _p],
    "b200mp_add_dropout_layernorm_fwd": [_p, _p, _l, _i, _f, C.c_uint64, _p, _p, _f, _p, _p, _p],
    "b200mp_add_dropout_layernorm_bwd": [_p, _p, _p, _p, _l, _i, _f, C.c_uint64, _p, _p, _p, _p, _p, _p],
    "b200mp_add_dropout_layernorm_fwd_ex": [_p, _p, _l, _i, _f, C.c_uint64, _p, _p, _f, _f, _p, _f, _p, _f, _p, _p, _p],
    "b200mp_add_dropout_layernorm_bwd_ex": [_p, _f, _p, _f, _p, _f, _p, _p, _p, _l, _i, _f, C.c_uint64, _p, _p, _p, _p, _p, _p],
    "b200mp_scaled_sum": [_p, _p, _i, _l, _p, _p],
    "b200mp_relu_dropout": [_p, _p, _l, _f, C.c_uint64, _p, _p],
    "b200mp_col_encode_fwd": [_p, _i, _p, _l, _i, _p, _p],
    "b200mp_col_encode_bwd": [_p, _i, _i, _l, _i, _p, _p, _p, _p, _p],
    "b200mp_message_bwd_rows": [_i, _l, _p, _p, _p, _p, _p, _p, _p, _i, _p, _p, _p, _i, _p, _p],
    "b200mp_edge_update_bwd_rows": [_l, _p, _p, _p, _p, _p, _f, _f, _p, _p, _p, _p, _p, _p, _i, _p, _p, _p, _p, _p],
    "b200mp_edge_linear_fwd_bf16": [_i, _l, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _f, _f, _p, _p, _p, _p],
    "b200mp_segment_reduce_fwd_bf16": [_i, _l, _p, _p, _p, _p, _l, _l, _l, _p, _p, _l, _p, _p, _p, _p],
    "b200mp_message_bwd_bf16": [_i, _l, _p, _p, _p, _p, _p, _p, _p, _p, _i, _p, _p, _p, _i, _p, _p],
    "b200mp_edge_update_bwd_bf16": [_l, _p, _p, _p, _p, _p, _p, _p, _f, _f, _p, _p, _p, _p, _p, _p, _p, _i, _p, _p, _p],
    "b200mp_pair_gather_act": [_p, _p, _p, _p, _p, _l, _i, _i, _p, _p],
    "b200mp_rank_positive": [_p, _p, _l, _i, _p, _p],
    "b200mp_segment_mean_blend_fwd": [_p, _p, _p, _p, _l, _i, _p, _p],
    "b200mp_segment_mean_blend_bwd": [_p, _p, _p, _l, _l, _i, _p, _p, _p],
    "b200mp_row_linear": [_p, _p, _l, _i, _i, _p, _i, _p, _p, _p, _p, _p, _i, _p, _p, _p, _f, _f, _p, _p, _p, _p],
    "b200mp_message_grad_rows": [_i, _i, _l, _p, _p, _p, _p, _p, _p, _p],
    "b200mp_scatter_add_rows": [_p, _p, _l, _i, _p, _p],
    "b200mp_umma_selftest": [_i, _p, _p, _p, _p],
    "b200mp_umma_probe": [_p, _i, _p, _i, C.c_uint64, C.c_uint64, _i, _i, _i, C.c_uint32, _p, _p],
    "b200mp_edge_update_bwd": [_i, _l, _p, _p, _p, _p, _p, _p, _p, _f, _f, _p, _p, _p, _p, _p, _p, _p, _i, _p, _p, _p],
}


class Column(C.Structure):
    """struct b200mp_column of include/b200mp.h"""
    _fields_ = [("kind", C.c_int32), ("offset", C.c_int32), ("stride", C.c_int64), ("feat", C.c_void_p),
                ("weight", C.c_void_p), ("bias", C.c_void_p), ("mean", C.c_float), ("std", C.c_float),
                ("min_year", C.c_float), ("ncat", C.c_int32)]


COL_NUMERICAL, COL_CATEGORICAL, COL_TIMESTAMP = 0, 1, 2


def declared_symbols():
    """Every entry point include/b200mp.h declares (used by the CPU test that checks the
    library exports all of them)."""
    with open(HEADER) as f:
        return sorted(set(re.findall(r"\bint\s+(b200mp_\w+)\s*\(", f.read())))


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(b200mp has no CPU or PyTorch fallback)")
        l = C.CDLL(LIB_PATH)
        for name, args in _SIGS.items():
            fn = getattr(l, name)
            fn.argtypes = args
            fn.restype = _i
        _lib = l
    return _lib


class B200MPError(RuntimeError):
    pass


def check(rc: int, what: str):
    if rc != 0:
        kind = {-1: "invalid argument", -2: "workspace too small", -3: "size beyond the entry point's fixed maximum"}.get(rc, f"cudaError {rc}" if rc > 0 else "error")
        raise B200MPError(f"{what} failed: {kind} (rc={rc})")


def ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def stream() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise B200MPError("b200mp kernels run on CUDA tensors only (no CPU fallback): got a "
                              f"{t.device} tensor")


_cache = {}


def partial_rows(H: int = 0) -> int:
    if ("pr", H) not in _cache:
        v = C.c_int(0)
        check(lib().b200mp_partial_rows_for(H, C.byref(v)), "b200mp_partial_rows_for")
        _cache[("pr", H)] = v.value
    return _cache[("pr", H)]


def tile_edges(H: int) -> int:
    if ("te", H) not in _cache:
        v = C.c_int(0)
        rc = lib().b200mp_tile_edges(H, C.byref(v))
        if rc != 0:
            raise B200MPError(f"hidden width {H} is not supported by the kernels (16/32/64/128)")
        _cache[("te", H)] = v.value
    return _cache[("te", H)]


def n_part(E: int, H: int) -> int:
    return max(1, min(partial_rows(H), (E + tile_edges(H) - 1) // tile_edges(H)))
