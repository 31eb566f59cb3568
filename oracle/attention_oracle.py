"""CPU oracle for the fused-attention hot path.  TEST INFRASTRUCTURE ONLY.

This module is a CPU (PyTorch, fp32/fp64) restatement of what the reference's
fused attention computes.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it; the
product package (``xhy-flash-attention_b200``) never does and has no CPU
fallback.

Parity pinning: ``attention_ref`` below follows the reference's own test oracle
(``tests/test_flash_attn.py:111-159``) and is pinned against golden vectors
produced by executing *that very function* (extracted with ``ast`` from the
reference tree) on seeded inputs -- see ``tests/golden/make_golden.py`` and
``tests/test_oracle_golden.py``.  The mask builders restate kernel-side
semantics the reference's tests do not cover (the reference stores no golden
vectors for them); each cites the file:line it follows.
"""
from __future__ import annotations

import math
from typing import Optional, Sequence, Tuple

import torch

NEG_INF = float("-inf")


# --------------------------------------------------------------------------- #
# mask builders (True = masked out)                                           #
# --------------------------------------------------------------------------- #
def causal_mask_top_left(sq: int, sk: int) -> torch.Tensor:
    """FA2 surface: column j is masked for row i iff j > i.

    Follows tests/test_flash_attn.py:144-146 (``triu(ones(sq, sk), 1)``) and the
    kernel rule ``col >= min(sk, row + 1)`` in csrc/flash_attn/src/softmax.h:143-178.
    """
    i = torch.arange(sq).view(sq, 1)
    j = torch.arange(sk).view(1, sk)
    return j > i


def causal_mask_bottom_right(sq: int, sk: int) -> torch.Tensor:
    """v2/v3 + kvcache surface: masked iff ``j >= i + 1 + sk - sq``.

    Follows csrc/flashmask_v2/mask.h:85-97.
    """
    i = torch.arange(sq).view(sq, 1)
    j = torch.arange(sk).view(1, sk)
    return j > i + (sk - sq)


def local_mask(sq: int, sk: int, window_left: int, window_right: int) -> torch.Tensor:
    """Sliding window, bottom-right aligned: keep ``i+sk-sq-wl <= j <= i+sk-sq+wr``.

    Negative window size = unbounded on that side (csrc/flashmask_v2/mask.h:101-118).
    """
    i = torch.arange(sq).view(sq, 1) + (sk - sq)
    j = torch.arange(sk).view(1, sk)
    m = torch.zeros(sq, sk, dtype=torch.bool)
    if window_left >= 0:
        m |= j < i - window_left
    if window_right >= 0:
        m |= j > i + window_right
    return m


def flashmask_dense(sq: int, sk: int, lt_start: torch.Tensor,
                    lt_end: Optional[torch.Tensor] = None,
                    ut_start: Optional[torch.Tensor] = None,
                    ut_end: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Dense (…, sq, sk) bool mask from FlashMask column vectors (…, sk) int32.

    Element rule (csrc/flashmask_v2/mainloop_fwd_sm90_tma_gmma_ws.hpp:1401-1461,
    FA2 form csrc/flash_attn/src/softmax.h:180-298): for key column j, query row i
    is masked iff ``lt_start[j] <= i < lt_end[j]`` or ``ut_start[j] <= i < ut_end[j]``;
    a missing ``lt_end`` means +inf, a missing ``ut_start`` means 0, a missing
    ``ut_end`` means "no upper band".
    """
    lead = lt_start.shape[:-1]
    i = torch.arange(sq).view(*([1] * len(lead)), sq, 1)
    lts = lt_start.long().unsqueeze(-2)
    lte = lt_end.long().unsqueeze(-2) if lt_end is not None else None
    m = i >= lts
    if lte is not None:
        m = m & (i < lte)
    if ut_end is not None:
        ute = ut_end.long().unsqueeze(-2)
        u = i < ute
        if ut_start is not None:
            u = u & (i >= ut_start.long().unsqueeze(-2))
        m = m | u
    return m.expand(*lead, sq, sk)


def document_lt_start(doc_lens: Sequence[int]) -> torch.Tensor:
    """Causal document mask as one FlashMask vector: lt_start[j] = end of j's document."""
    out = []
    end = 0
    for n in doc_lens:
        end += int(n)
        out.extend([end] * int(n))
    return torch.tensor(out, dtype=torch.int32)


# --------------------------------------------------------------------------- #
# reference attention                                                         #
# --------------------------------------------------------------------------- #
def attention_ref(q, k, v, query_padding_mask=None, key_padding_mask=None, dropout_p=0.0,
                  dropout_mask=None, causal=False, upcast=True, reorder_ops=False):
    """Restatement of the reference test oracle, tests/test_flash_attn.py:111-159.

    q (b, sq, h, d); k, v (b, sk, h_k, d).  GQA by repeating K/V heads (:135-136);
    causal is TOP-LEFT aligned (:144-146); padded keys -> -inf (:142-143); padded
    query rows -> zero output (:156-158).  Returns (output, attention) in q.dtype.
    """
    dtype_og = q.dtype
    if upcast:
        q, k, v = q.float(), k.float(), v.float()
    b, sq, h, d = q.shape
    sk, hk = k.shape[1], k.shape[2]
    g = h // hk
    k = k.repeat_interleave(g, dim=2)
    v = v.repeat_interleave(g, dim=2)
    if not reorder_ops:
        scores = torch.einsum("bthd,bshd->bhts", q / math.sqrt(d), k)
    else:
        scores = torch.einsum("bthd,bshd->bhts", q, k / math.sqrt(d))
    if key_padding_mask is not None:
        scores.masked_fill_(~key_padding_mask.view(b, 1, 1, sk), NEG_INF)
    if causal:
        scores.masked_fill_(causal_mask_top_left(sq, sk).to(q.device), NEG_INF)
    attention = torch.softmax(scores, dim=-1)
    dropout_scaling = 1.0 / (1 - dropout_p)
    attention_drop = attention.masked_fill(~dropout_mask, 0.0) if dropout_mask is not None else attention
    output = torch.einsum("bhts,bshd->bthd", attention_drop, v * dropout_scaling)
    if query_padding_mask is not None:
        output.masked_fill_(~query_padding_mask.view(b, sq, 1, 1), 0.0)
        attention = attention.masked_fill(~query_padding_mask.view(b, 1, sq, 1), 0.0)
    return output.to(dtype=dtype_og), attention.to(dtype=dtype_og)


def attention_masked_ref(q, k, v, mask: Optional[torch.Tensor] = None, softmax_scale: Optional[float] = None,
                         upcast: bool = True, reorder_ops: bool = False,
                         lse_empty: float = float("inf"), bias: Optional[torch.Tensor] = None,
                         softcap: float = 0.0) -> Tuple[torch.Tensor, torch.Tensor]:
    """General oracle: arbitrary boolean mask (broadcastable to (b, h, sq, sk); True = masked).

    Returns (out (b,sq,h,d) in q.dtype, lse (b,h,sq) fp32).  Rows with every key
    masked give out = 0 and ``lse = lse_empty``: +inf on the FA2 surface
    (csrc/flash_attn/src/flash_fwd_kernel.h:514), -inf on the v2/v3 surface
    (csrc/flashmask_v2/softmax.h:153).  ``lse = logsumexp(scale * q.k)`` over the
    unmasked keys (natural log).  Same op order as ``attention_ref`` so that the
    low-precision run (upcast=False, reorder_ops=True) is the reference's own
    "PyTorch baseline" for the 2x-error criterion (tests/test_flash_attn.py:373-374).
    """
    dtype_og = q.dtype
    if upcast:
        q, k, v = q.float(), k.float(), v.float()
    b, sq, h, d = q.shape
    hk = k.shape[2]
    g = h // hk
    k = k.repeat_interleave(g, dim=2)
    v = v.repeat_interleave(g, dim=2)
    scale = (1.0 / math.sqrt(d)) if softmax_scale is None else softmax_scale
    if not reorder_ops:
        scores = torch.einsum("bthd,bshd->bhts", q * scale, k)
    else:
        scores = torch.einsum("bthd,bshd->bhts", q, k * scale)
    if softcap > 0.0:      # tanh soft-capping of the scaled scores, before bias and masks
        scores = softcap * torch.tanh(scores / softcap)   # (csrc/flash_attn_v3/mainloop_fwd_sm90_tma_gmma_ws.hpp:537-556)
    if bias is not None:   # dense additive mask, broadcastable to (b, h, sq, sk) (csrc/flash_attn/src/softmax.h:360-379)
        scores = scores + bias.to(scores.dtype)
    if mask is not None:
        scores = scores.masked_fill(mask.to(scores.device), NEG_INF)
    lse = torch.logsumexp(scores.float(), dim=-1)
    empty = torch.isinf(lse) & (lse < 0)
    attention = torch.softmax(scores, dim=-1)
    attention = torch.where(empty.unsqueeze(-1), torch.zeros_like(attention), attention)
    out = torch.einsum("bhts,bshd->bthd", attention, v)
    lse = torch.where(empty, torch.full_like(lse, lse_empty), lse)
    return out.to(dtype_og), lse


def attention_grads_ref(q, k, v, dout, mask=None, softmax_scale=None, upcast=True, reorder_ops=False, bias=None,
                        softcap=0.0):
    """(out, lse, dq, dk, dv) by autograd through ``attention_masked_ref``.

    This is how the reference's tests obtain gradients of the oracle
    (tests/test_flash_attn.py:404-416: ``torch.autograd.grad(out_ref, (q,k,v), g)``).
    """
    q_ = q.detach().clone().requires_grad_(True)
    k_ = k.detach().clone().requires_grad_(True)
    v_ = v.detach().clone().requires_grad_(True)
    out, lse = attention_masked_ref(q_, k_, v_, mask, softmax_scale, upcast, reorder_ops, bias=bias, softcap=softcap)
    dq, dk, dv = torch.autograd.grad(out, (q_, k_, v_), dout.to(out.dtype))
    return out.detach(), lse.detach(), dq, dk, dv


def attention_varlen_ref(q, k, v, cu_seqlens_q, cu_seqlens_k, causal=False, softmax_scale=None,
                         upcast=True, reorder_ops=False, bottom_right=False):
    """Packed (total, h, d) oracle: loop over sequences (csrc/flash_attn/src/block_info.h:12-37).

    Returns (out (total_q,h,d), list of per-sequence lse (h, len_q))."""
    outs, lses = [], []
    nb = len(cu_seqlens_q) - 1
    for i in range(nb):
        qs, qe = int(cu_seqlens_q[i]), int(cu_seqlens_q[i + 1])
        ks, ke = int(cu_seqlens_k[i]), int(cu_seqlens_k[i + 1])
        sq, sk = qe - qs, ke - ks
        mask = None
        if causal:
            mask = causal_mask_bottom_right(sq, sk) if bottom_right else causal_mask_top_left(sq, sk)
        o, l = attention_masked_ref(q[qs:qe].unsqueeze(0), k[ks:ke].unsqueeze(0), v[ks:ke].unsqueeze(0),
                                    mask, softmax_scale, upcast, reorder_ops)
        outs.append(o[0])
        lses.append(l[0])
    return torch.cat(outs, 0), lses


# --------------------------------------------------------------------------- #
# split-KV combine + ring merge                                               #
# --------------------------------------------------------------------------- #
def combine_ref(o_partial: torch.Tensor, lse_partial: torch.Tensor):
    """LSE-combine of split partials (csrc/flash_attn_v3/flash_fwd_combine_kernel.h:340-376,423-453).

    o_partial (S, ..., d) fp32, lse_partial (S, ...) fp32 with -inf for empty splits.
    ``lse = logsumexp_s(lse_s)``, ``o = sum_s exp(lse_s - lse) o_s``; all-empty -> o = 0, lse = -inf.
    """
    lse = torch.logsumexp(lse_partial, dim=0)
    w = torch.exp(lse_partial - torch.where(torch.isinf(lse), torch.zeros_like(lse), lse))
    w = torch.where(torch.isinf(lse_partial) & (lse_partial < 0), torch.zeros_like(w), w)
    o = (w.unsqueeze(-1) * o_partial).sum(0)
    return o, lse


def merge_ref(o_acc, lse_acc, o_step, lse_step):
    """Two-way LSE merge used by ring attention (same algebra as ``combine_ref``); an infinite LSE of either sign
    means "no contribution".  o (..., d) fp32, lse (...) fp32.  Returns (o, lse)."""
    e1, e2 = torch.isinf(lse_acc), torch.isinf(lse_step)
    l1 = torch.where(e1, torch.full_like(lse_acc, NEG_INF), lse_acc)
    l2 = torch.where(e2, torch.full_like(lse_step, NEG_INF), lse_step)
    o, lse = combine_ref(torch.stack([o_acc.float(), o_step.float()]), torch.stack([l1, l2]))
    both = e1 & e2
    return o, torch.where(both, lse_acc, lse)


def attention_block_bwd_ref(dout, q, k, v, out, lse, mask=None, softmax_scale=None):
    """Backward of ONE (query block, key block) pair given the FINAL out / lse of the whole row
    (csrc/flash_attn/src/flash_bwd_kernel.h:847 ``P = exp(S - LSE)``, :112-188 ``dPsum = rowsum(dO o O)``,
    :897-906 ``dS = P o (dP - dPsum)``).  fp32.  q, dout, out (b, sq, h, d); k, v (b, sk, h_k, d); lse (b, h, sq).
    Returns (dq, dk, dv) fp32 with dk/dv summed over the q-heads of each kv-head."""
    b, sq, h, d = q.shape
    hk = k.shape[2]
    g = h // hk
    qf, kf, vf, of, gf = q.float(), k.float().repeat_interleave(g, 2), v.float().repeat_interleave(g, 2), out.float(), dout.float()
    scale = (1.0 / math.sqrt(d)) if softmax_scale is None else softmax_scale
    s = torch.einsum("bthd,bshd->bhts", qf, kf) * scale
    if mask is not None:
        s = s.masked_fill(mask.to(s.device), NEG_INF)
    lse_ = torch.where(torch.isinf(lse), torch.full_like(lse, float("inf")), lse)
    p = torch.exp(s - lse_.unsqueeze(-1))
    dp = torch.einsum("bthd,bshd->bhts", gf, vf)
    dpsum = (gf * of).sum(-1).permute(0, 2, 1)                      # (b, h, sq)
    ds = p * (dp - dpsum.unsqueeze(-1))
    dq = torch.einsum("bhts,bshd->bthd", ds, kf) * scale
    dk = torch.einsum("bhts,bthd->bshd", ds, qf) * scale
    dv = torch.einsum("bhts,bthd->bshd", p, gf)
    dk = dk.view(b, -1, hk, g, d).sum(3)
    dv = dv.view(b, -1, hk, g, d).sum(3)
    return dq, dk, dv


# --------------------------------------------------------------------------- #
# FLOP convention (benchmarks/benchmark_flash_attention.py:27-30)             #
# --------------------------------------------------------------------------- #
def attention_flops(batch, seqlen_q, seqlen_k, nheads, headdim, causal, mode="fwd"):
    assert mode in ("fwd", "bwd", "fwd_bwd")
    f = 4 * batch * seqlen_q * seqlen_k * nheads * headdim // (2 if causal else 1)
    return f if mode == "fwd" else (2.5 * f if mode == "bwd" else 3.5 * f)


# ---------------------------------------------------------------------------------------------------------------------
# KV-cache decode extras (csrc/flash_attn_v3: rotary.h, paged_kv.h, seqlen.h:56-91, mainloop softcap :537-556)
# ---------------------------------------------------------------------------------------------------------------------
def rotary_ref(x, cos, sin, positions, interleaved=True):
    """Rotary embedding of x (b, s, h, d) at integer ``positions`` (b, s): pairs (2j, 2j+1) when ``interleaved`` else
    (j, j + rotary_dim/2); x1' = x1 cos - x2 sin, x2' = x1 sin + x2 cos, computed in fp32 and rounded once
    (csrc/flash_attn_v3/rotary.h apply_rotary_interleaved / apply_rotary_contiguous).  cos, sin: (seqlen_ro, rotary_dim/2)."""
    b, s, h, d = x.shape
    half = cos.shape[-1]
    rdim = 2 * half
    c = cos.float()[positions.long()].unsqueeze(2)          # (b, s, 1, half)
    sn = sin.float()[positions.long()].unsqueeze(2)
    xr = x[..., :rdim].float()
    if interleaved:
        x1, x2 = xr[..., 0::2], xr[..., 1::2]
        o = torch.stack((x1 * c - x2 * sn, x1 * sn + x2 * c), dim=-1).reshape(b, s, h, rdim)
    else:
        x1, x2 = xr[..., :half], xr[..., half:]
        o = torch.cat((x1 * c - x2 * sn, x1 * sn + x2 * c), dim=-1)
    return torch.cat((o.to(x.dtype), x[..., rdim:]), dim=-1)


def alibi_bias(slopes, sq, sk):
    """ALiBi bias -slope * |i + sk - sq - j| for slopes (h,) or (b, h) -> (1 or b, h, sq, sk) fp32."""
    i = torch.arange(sq, device=slopes.device).view(-1, 1)
    j = torch.arange(sk, device=slopes.device).view(1, -1)
    dist = (i + sk - sq - j).abs().float()
    sl = slopes.float()
    sl = sl.view(1, -1, 1, 1) if sl.dim() == 1 else sl.view(sl.shape[0], sl.shape[1], 1, 1)
    return -sl * dist


def kvcache_ref(q, k_cache, v_cache, k=None, v=None, rotary_cos=None, rotary_sin=None, cache_seqlens=None,
                cache_batch_idx=None, cache_leftpad=None, block_table=None, softmax_scale=None, causal=False,
                window_size=(-1, -1), softcap=0.0, rotary_interleaved=True, alibi_slopes=None, upcast=True):
    """flash_attn_with_kvcache restated with plain indexing, one batch element at a time.

    Returns (out (b, sq, h, d), lse (b, h, sq), k_cache', v_cache') -- the caches are updated COPIES.
    Valid keys of element i are the physical rows [leftpad_i, cache_seqlens_i (+ seqlen_new)) of cache row
    cache_batch_idx[i], or of the pages block_table[i] laid end to end."""
    b, sq, h, d = q.shape
    k_cache, v_cache = k_cache.clone(), v_cache.clone()
    if cache_seqlens is None:
        lens = torch.zeros(b, dtype=torch.long)
    elif isinstance(cache_seqlens, int):
        lens = torch.full((b,), cache_seqlens, dtype=torch.long)
    else:
        lens = cache_seqlens.long().cpu()
    lp = torch.zeros(b, dtype=torch.long) if cache_leftpad is None else cache_leftpad.long().cpu()
    local = tuple(window_size) != (-1, -1)
    s_new = 0 if k is None else k.shape[1]
    if rotary_cos is not None:
        pos_k = lens.view(b, 1).to(q.device) + torch.arange(s_new, device=q.device).view(1, -1)
        k = rotary_ref(k, rotary_cos, rotary_sin, pos_k, rotary_interleaved)
        if causal or local:
            pos_q = lens.view(b, 1).to(q.device) + torch.arange(sq, device=q.device).view(1, -1)
        else:
            pos_q = lens.view(b, 1).to(q.device).expand(b, sq)
        q = rotary_ref(q, rotary_cos, rotary_sin, pos_q, rotary_interleaved)
    outs, lses = [], []
    for i in range(b):
        if block_table is not None:
            page_size = k_cache.shape[1]
            rows = lambda lo, hi: (block_table[i].long()[torch.arange(lo, hi) // page_size], torch.arange(lo, hi) % page_size)
        else:
            slab = i if cache_batch_idx is None else int(cache_batch_idx[i])
            rows = lambda lo, hi: (torch.full((hi - lo,), slab, dtype=torch.long), torch.arange(lo, hi))
        n0 = int(lens[i])
        if k is not None:
            a, r = rows(n0, n0 + s_new)
            k_cache[a.to(q.device), r.to(q.device)] = k[i]
            v_cache[a.to(q.device), r.to(q.device)] = v[i]
        a, r = rows(int(lp[i]), n0 + s_new)
        ki = k_cache[a.to(q.device), r.to(q.device)].unsqueeze(0)
        vi = v_cache[a.to(q.device), r.to(q.device)].unsqueeze(0)
        sk = ki.shape[1]
        mask = None
        if local:
            mask = local_mask(sq, sk, window_size[0], window_size[1])
            if causal:
                mask = mask | causal_mask_bottom_right(sq, sk)
        elif causal:
            mask = causal_mask_bottom_right(sq, sk)
        bias = None
        if alibi_slopes is not None:
            sl = alibi_slopes if alibi_slopes.dim() == 1 else alibi_slopes[i]
            bias = alibi_bias(sl, sq, sk).to(q.device)
        o, l = attention_masked_ref(q[i:i + 1], ki, vi, mask, softmax_scale, upcast=upcast, lse_empty=float("-inf"),
                                    bias=bias, softcap=softcap)
        outs.append(o)
        lses.append(l)
    return torch.cat(outs), torch.cat(lses), k_cache, v_cache
