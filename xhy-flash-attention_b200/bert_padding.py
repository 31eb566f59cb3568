"""Padded <-> packed conversion for the varlen functions (the reference ships the same helpers as
``flash_attn/bert_padding.py:97-135``; its tests build every varlen case with them, tests/test_flash_attn.py:38-108).

``unpad_input`` turns a padded batch (batch, seqlen, ...) and a validity mask into the packed rows
(total_tokens, ...) plus the ``cu_seqlens`` / ``max_seqlen`` that ``flash_attn_varlen_func`` takes; ``pad_input`` is its
inverse.  Both are differentiable.  Row gather / scatter are torch's own indexed kernels -- this is data layout on
either side of the attention call, not part of the attention hot path.
"""
from __future__ import annotations

import torch

__all__ = ["index_first_axis", "index_put_first_axis", "unpad_input", "pad_input"]


def index_first_axis(x: torch.Tensor, indices: torch.Tensor) -> torch.Tensor:
    """``x[indices]`` along dim 0 with integer indices (gradient: scatter-add into zeros)."""
    return torch.index_select(x, 0, indices)


def index_put_first_axis(values: torch.Tensor, indices: torch.Tensor, first_axis_dim: int) -> torch.Tensor:
    """Zeros of shape (first_axis_dim, ...) with ``out[indices] = values`` (gradient: gather)."""
    out = torch.zeros((first_axis_dim,) + tuple(values.shape[1:]), device=values.device, dtype=values.dtype)
    return out.index_copy(0, indices, values)


def unpad_input(hidden_states: torch.Tensor, attention_mask: torch.Tensor):
    """hidden_states (batch, seqlen, ...), attention_mask (batch, seqlen) bool / int, nonzero = valid token.

    Returns ``(packed (total, ...), indices (total,) into the flattened batch*seqlen axis, cu_seqlens (batch + 1,) int32,
    max_seqlen_in_batch int)``."""
    valid = attention_mask.to(torch.bool)
    seqlens = valid.sum(dim=-1, dtype=torch.int32)
    indices = torch.nonzero(valid.reshape(-1), as_tuple=False).reshape(-1)
    cu_seqlens = torch.zeros(seqlens.numel() + 1, dtype=torch.int32, device=seqlens.device)
    cu_seqlens[1:] = torch.cumsum(seqlens, dim=0)
    flat = hidden_states.reshape((-1,) + tuple(hidden_states.shape[2:]))
    return index_first_axis(flat, indices), indices, cu_seqlens, int(seqlens.max().item())


def pad_input(hidden_states: torch.Tensor, indices: torch.Tensor, batch: int, seqlen: int) -> torch.Tensor:
    """packed (total, ...) -> padded (batch, seqlen, ...) with zeros at the invalid positions."""
    out = index_put_first_axis(hidden_states, indices, batch * seqlen)
    return out.reshape((batch, seqlen) + tuple(hidden_states.shape[1:]))
