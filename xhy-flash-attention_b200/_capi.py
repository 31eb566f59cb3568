"""ctypes binding of libflashattn_b200.so (include/fa_capi.h, include/fa_b200.h).

This is the thin layer between the PyTorch ``autograd.Function`` wrappers and the C ABI.  PyTorch is used for
device memory and streams only; every attention computation happens in the CUDA kernels of the shared
library.  There is no CPU fallback: if the library is missing or cannot be loaded, importing the operator API
raises immediately.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# FA_B200_LIB: load another build of the same library (A/B timing of two builds on one box); default = the in-tree build
LIB_PATH = os.environ.get("FA_B200_LIB") or os.path.join(_HERE, "lib", "libflashattn_b200.so")

FA_B200_ABI_VERSION = 6


class Tensor(C.Structure):
    _fields_ = [("ptr", C.c_void_p), ("batch_stride", C.c_int64), ("row_stride", C.c_int64),
                ("head_stride", C.c_int64)]


class Params(C.Structure):
    """Mirror of ``fa_b200_params`` (include/fa_b200.h) -- field order and types must match exactly."""
    _fields_ = [
        ("abi_version", C.c_int32), ("is_bf16", C.c_int32),
        ("q", Tensor), ("k", Tensor), ("v", Tensor), ("o", Tensor),
        ("dout", Tensor), ("dq", Tensor), ("dk", Tensor), ("dv", Tensor),
        ("softmax_lse", C.c_void_p), ("softmax_d", C.c_void_p), ("dq_accum", C.c_void_p), ("dkv_accum", C.c_void_p),
        ("b", C.c_int32), ("seqlen_q", C.c_int32), ("seqlen_k", C.c_int32),
        ("h", C.c_int32), ("h_k", C.c_int32), ("d", C.c_int32),
        ("seqlen_q_rounded", C.c_int32), ("seqlen_k_rounded", C.c_int32), ("d_rounded", C.c_int32),
        ("total_q", C.c_int64), ("total_k", C.c_int64),
        ("cu_seqlens_q", C.c_void_p), ("cu_seqlens_k", C.c_void_p), ("seqused_k", C.c_void_p),
        ("varlen_padded_input", C.c_int32),
        ("softmax_scale", C.c_float), ("lse_empty_is_neg_inf", C.c_int32),
        ("is_causal", C.c_int32), ("causal_bottom_right", C.c_int32),
        ("window_left", C.c_int32), ("window_right", C.c_int32),
        ("lt_start", C.c_void_p), ("lt_end", C.c_void_p), ("ut_start", C.c_void_p), ("ut_end", C.c_void_p),
        ("h_mask", C.c_int32), ("flashmask_maxmin", C.c_void_p),
        ("dkv_expanded", C.c_int32), ("deterministic", C.c_int32), ("force_generic", C.c_int32),
        ("num_splits", C.c_int32), ("oaccum", C.c_void_p), ("lseaccum", C.c_void_p),
        ("p_dropout", C.c_float), ("philox_seed", C.c_uint64), ("philox_offset", C.c_uint64), ("s_dmask", C.c_void_p),
        ("attn_mask", C.c_void_p), ("mask_head_mod", C.c_int32), ("mask_seq_mod", C.c_int32),
        ("unscale_softmax", C.c_float),
        ("page_table", C.c_void_p), ("page_table_batch_stride", C.c_int64),
        ("page_size", C.c_int32), ("num_pages", C.c_int32),
        ("kv_batch_idx", C.c_void_p), ("b_k", C.c_int32),
        ("leftpad_k", C.c_void_p), ("softcap", C.c_float),
        ("alibi_slopes", C.c_void_p), ("alibi_batch_stride", C.c_int64),
        ("attn_bias", C.c_void_p), ("bias_batch_mod", C.c_int32), ("bias_rows", C.c_int32), ("ds_out", C.c_void_p),
        ("dq_accum_batch_major", C.c_int32), ("seqused_q", C.c_void_p),
        ("num_splits_dynamic", C.c_void_p),
        ("is_e4m3", C.c_int32), ("q_descale", C.c_void_p), ("k_descale", C.c_void_p), ("v_descale", C.c_void_p),
        ("q_descale_batch_stride", C.c_int64), ("q_descale_head_stride", C.c_int64),
        ("k_descale_batch_stride", C.c_int64), ("k_descale_head_stride", C.c_int64),
        ("v_descale_batch_stride", C.c_int64), ("v_descale_head_stride", C.c_int64),
    ]


_lib: Optional[C.CDLL] = None


def load() -> C.CDLL:
    """Load the shared library (once).  Raises ``RuntimeError`` if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU or PyTorch fallback for the attention kernels)")
    lib = C.CDLL(LIB_PATH)
    pp = C.POINTER(Params)
    for name in ("fa_b200_fwd", "fa_b200_bwd", "fa_b200_fwd_splitkv", "fa_b200_fwd_combine", "fa_b200_fwd_splitkv_noalloc"):
        fn = getattr(lib, name)
        fn.argtypes = [pp, C.c_void_p]
        fn.restype = C.c_bool
    lib.fa_b200_flashmask_prepare.argtypes = [pp, C.c_int, C.c_void_p]
    lib.fa_b200_flashmask_prepare.restype = C.c_bool
    lib.fa_b200_num_splits.argtypes = [C.c_int] * 6
    lib.fa_b200_num_splits.restype = C.c_int
    lib.fa_b200_decode_num_splits.argtypes = [C.c_int] * 4
    lib.fa_b200_decode_num_splits.restype = C.c_int
    lib.fa_b200_append_kv.argtypes = ([C.c_void_p] * 5 + [C.c_int] * 6 + [C.c_int64] * 6 + [C.c_void_p])
    lib.fa_b200_append_kv.restype = C.c_bool
    lib.fa_b200_append_kv_paged.argtypes = [C.POINTER(Params), C.c_void_p, C.c_void_p, C.c_int] + [C.c_int64] * 3 + [C.c_void_p]
    lib.fa_b200_append_kv_paged.restype = C.c_bool
    lib.fa_b200_rotary.argtypes = [C.c_void_p] * 5 + [C.c_int] * 9 + [C.c_int64] * 6 + [C.c_void_p]
    lib.fa_b200_rotary.restype = C.c_bool
    lib.fa_b200_lse_merge.argtypes = ([C.c_void_p] * 4 + [C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int,
                                                         C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_void_p])
    lib.fa_b200_lse_merge.restype = C.c_bool
    lib.fa_b200_lse_merge_batched.argtypes = ([C.c_void_p] * 4 + [C.c_int, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int] +
                                              [C.c_int64] * 8 + [C.c_void_p])
    lib.fa_b200_lse_merge_batched.restype = C.c_bool
    lib.fa_b200_prepare_num_splits.argtypes = [pp, C.c_void_p, C.c_int, C.c_void_p]
    lib.fa_b200_prepare_num_splits.restype = C.c_bool
    lib.fa_b200_abi_version.restype = C.c_int
    lib.fa_b200_last_kernel_kind.restype = C.c_int
    lib.fa_b200_set_forward_sm_reserve.argtypes = [C.c_int]
    lib.fa_b200_set_forward_sm_reserve.restype = None
    lib.fa_b200_launch_count.argtypes = [C.c_int]
    lib.fa_b200_launch_count.restype = C.c_int64
    lib.fa_b200_last_fallback_reason.restype = C.c_char_p
    lib.fa_b200_fallback_count.argtypes = [C.c_int]
    lib.fa_b200_fallback_count.restype = C.c_int64
    lib.fa_b200_profile_enable.argtypes = [C.c_int]
    lib.fa_b200_profile_enable.restype = None
    lib.fa_b200_profile_last_ms.argtypes = [C.c_int]
    lib.fa_b200_profile_last_ms.restype = C.c_float
    lib.flash_attn_error.restype = C.c_char_p
    lib.flash_attn_set_error.argtypes = [C.c_char_p]
    if lib.fa_b200_abi_version() != FA_B200_ABI_VERSION:
        raise RuntimeError("libflashattn_b200.so ABI version mismatch")
    _lib = lib
    return lib


def last_error() -> str:
    return load().flash_attn_error().decode("utf-8", "replace")


def check(ok: bool, what: str) -> None:
    if not ok:
        raise RuntimeError(f"{what} failed: {last_error()}")


def stream_ptr(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def tdesc(t: Optional[torch.Tensor], packed: bool = False) -> Tensor:
    """(batch, row, head, d) strided view -> fa_b200_tensor.  ``packed``: (rows, head, d), batch stride 0."""
    if t is None:
        return Tensor(None, 0, 0, 0)
    if packed:
        assert t.dim() == 3 and t.stride(-1) == 1
        return Tensor(t.data_ptr(), 0, t.stride(0), t.stride(1))
    assert t.dim() == 4 and t.stride(-1) == 1
    return Tensor(t.data_ptr(), t.stride(0), t.stride(1), t.stride(2))


def ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def new_params(dtype: torch.dtype) -> Params:
    p = Params()
    p.abi_version = FA_B200_ABI_VERSION
    if dtype == torch.bfloat16:
        p.is_bf16 = 1
    elif dtype == torch.float16:
        p.is_bf16 = 0
    else:
        raise RuntimeError("FlashAttention only supports fp16 and bf16 data type")
    p.total_q = p.total_k = -1
    p.window_left = p.window_right = -1
    return p


def launch_count(reset: bool = False) -> int:
    return int(load().fa_b200_launch_count(1 if reset else 0))


def last_kernel_kind() -> int:
    return int(load().fa_b200_last_kernel_kind())


def last_fallback_reason() -> str:
    """Why the last call of this thread ran on the generic SIMT kernels ("" if it ran on tcgen05)."""
    return load().fa_b200_last_fallback_reason().decode("utf-8", "replace")


def fallback_count(reset: bool = False) -> int:
    return int(load().fa_b200_fallback_count(1 if reset else 0))
