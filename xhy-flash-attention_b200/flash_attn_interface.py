"""Operator API -- drop-in for the reference's ``flash_attn/flash_attn_interface.py``.

Same public names, positional order and defaults as the reference (flash_attn/__init__.py:3-8;
flash_attn_interface.py:282-492), each backed by a ``torch.autograd.Function`` with the same
save/return behaviour (:88-279).  The functions below differ from the reference only in what sits
underneath: instead of the ``flash_attn_2_cuda`` pybind module they call the C ABI of
``libflashattn_b200.so`` (``fa_b200_fwd`` / ``fa_b200_bwd``) through ctypes.

Behaviour kept from the reference:
  * ``softmax_scale=None`` -> ``headdim ** -0.5``                               (:220-221)
  * non-unit last stride -> ``.contiguous()``; other strides are passed through  (:40-41)
  * head dim not a multiple of 8 is zero-padded and sliced back                 (csrc/flash_attn/flash_api.cpp:238-247,318-321)
  * ``return_attn_probs=True`` returns ``(out, softmax_lse, S_dmask)``; ``S_dmask`` is ``None``
    unless dropout is active                                                    (:224,230)
  * causal masks are TOP-LEFT aligned                                            (tests/test_flash_attn.py:144-146)
  * packed inputs get one packed gradient tensor                                (:108-113,167-176)

Dropout (``dropout_p > 0``) runs inside the tcgen05 forward / backward kernels for head dims up to 128 (generic SIMT
kernels beyond) with a Philox-4x32-7 mask and the reference's 8-bit keep threshold
(csrc/flash_attn/src/softmax.h:381-447); the random stream is this library's own (one Philox call per 16 columns of a
row), so masks are reproducible from ``rng_state`` but not bit-identical to the reference's -- the reference's tests
check dropout through ``S_dmask`` and the drop fraction, which is how tests/test_dropout_gpu.py checks it here.
``return_attn_probs`` keeps the forward on the same kernels and adds one tcgen05 pass (csrc/sm100_probs.cu) that writes
``S_dmask`` from Q, K and the finished LSE: plain row-major (batch, heads, seqlen_q_rounded, seqlen_k_rounded), softmax
probabilities, negated where dropped, zero where masked.

Added on top of the reference surface (keyword-only, all default to the reference behaviour):
``window_size``, ``flashmask`` (the fork's column-sparse mask vectors), ``flash_attn_with_kvcache``
(upstream signature; the fork has the capability only at C level, csrc/flash_attn_v3/flash.h:96-125) and
``flashmask_attention``.
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

from . import _capi
from ._capi import Params, new_params, ptr, round_up, tdesc


def _maybe_contiguous(x):
    return x.contiguous() if x is not None and x.stride(-1) != 1 else x


def _pad_head(x, d_pad):
    return x if x.shape[-1] == d_pad else F.pad(x, [0, d_pad - x.shape[-1]])


class _Mask:
    """Masking options beyond ``causal`` (all off by default)."""
    __slots__ = ("window", "bottom_right", "lt_start", "lt_end", "ut_start", "ut_end", "lse_neg_inf", "force_generic")

    def __init__(self, window=(-1, -1), bottom_right=False, flashmask=None, lse_neg_inf=False, force_generic=False):
        self.window = tuple(window)
        self.bottom_right = bottom_right or self.window != (-1, -1)
        self.lt_start = self.lt_end = self.ut_start = self.ut_end = None
        if flashmask is not None:
            vecs = list(flashmask) + [None] * (4 - len(flashmask))
            self.lt_start, self.lt_end, self.ut_start, self.ut_end = vecs[:4]
            if self.lt_start is None:
                raise ValueError("flashmask: lt_start is mandatory")
        self.lse_neg_inf = lse_neg_inf
        self.force_generic = force_generic


_NO_MASK = _Mask()


def _fill_mask(p: Params, mask: _Mask, b: int, sk: int, keep: list):
    p.causal_bottom_right = 1 if mask.bottom_right else 0
    p.window_left, p.window_right = mask.window
    p.lse_empty_is_neg_inf = 1 if mask.lse_neg_inf else 0
    p.force_generic = 1 if mask.force_generic else 0
    if mask.lt_start is not None:
        lts = mask.lt_start
        assert lts.dtype == torch.int32 and lts.dim() == 3 and lts.shape[0] == b and lts.shape[2] == sk, \
            "flashmask vectors must be int32 (batch, mask_heads, seqlen_k)"
        p.h_mask = lts.shape[1]
        for name in ("lt_start", "lt_end", "ut_start", "ut_end"):
            t = getattr(mask, name)
            if t is not None:
                assert t.shape == lts.shape and t.dtype == torch.int32
                t = t.contiguous()
                keep.append(t)
                setattr(p, name, t.data_ptr())
        nblock = (((sk + 63) // 64) + 3) & ~3
        ws = torch.empty(8 * b * p.h_mask * nblock, dtype=torch.int32, device=lts.device)
        keep.append(ws)
        p.flashmask_maxmin = ws.data_ptr()


_SM_COUNT = {}
_FORCE_EXPANDED_HEADS = None      # tests: True / False overrides the wave model below


def _prefer_expanded_heads(b, sk, h, h_k, device) -> bool:
    """Wave model of the backward grid: CTAs = key blocks x heads x batch, one CTA per SM.  True when the per-q-head
    grid loses at least 6 points less to the last, partly filled wave than the per-kv-head grid (small grids only)."""
    if _FORCE_EXPANDED_HEADS is not None:
        return _FORCE_EXPANDED_HEADS
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx not in _SM_COUNT:
        _SM_COUNT[idx] = torch.cuda.get_device_properties(idx).multi_processor_count
    sms = _SM_COUNT[idx]
    nkb = (sk + 127) // 128
    grouped = nkb * h_k * b
    if grouped < sms or grouped >= 8 * sms:      # under one wave nothing is lost; deep grids amortise the last wave
        return False
    eff = lambda n: (n / sms) / -(-n // sms)
    return eff(nkb * h * b) > eff(grouped) + 0.06


def _new_rng_state(device=None):
    """(seed, offset) for the Philox dropout mask, as a small CPU tensor (saving it for backward costs no device sync).

    Taken from the CUDA generator of the tensor's device like the reference does (csrc/flash_attn/flash_api.cpp:300-316):
    seed = the generator's seed, offset = its current Philox offset, which is then advanced -- so ``torch.manual_seed`` /
    ``torch.cuda.manual_seed`` reproduce a run and consecutive calls get different masks.  Falls back to torch's CPU
    generator where the CUDA generator is not reachable."""
    try:
        idx = device.index if device is not None and device.index is not None else torch.cuda.current_device()
        gen = torch.cuda.default_generators[idx]
        seed, offset = int(gen.initial_seed()), int(gen.get_offset())
        gen.set_offset(offset + 4)          # one (seed, offset) pair per launch; cuRAND's convention keeps offsets multiples of 4
        return torch.tensor([seed & (2 ** 63 - 1), offset], dtype=torch.int64)
    except Exception:  # noqa: BLE001
        return torch.randint(0, 2 ** 62, (2,), dtype=torch.int64)


def _flash_attn_forward(q, k, v, dropout_p, softmax_scale, causal, return_softmax, mask: _Mask = _NO_MASK,
                        cu_seqlens_q=None, cu_seqlens_k=None, max_seqlen_q=None, max_seqlen_k=None,
                        seqused_k=None):
    """Equivalent of ``flash_attn_2_cuda.fwd`` / ``varlen_fwd`` (csrc/flash_attn/flash_api.cpp:186-324,326-491).

    Returns ``(out, q, k, v, out_padded, softmax_lse, S_dmask, rng_state)`` like the reference binding.
    """
    if not 0.0 <= dropout_p < 1.0:
        raise ValueError("dropout_p must be in [0, 1)")
    q, k, v = [_maybe_contiguous(x) for x in (q, k, v)]
    if not q.is_cuda:
        raise RuntimeError("q must be on CUDA device")  # flash_api.cpp:209
    varlen = cu_seqlens_q is not None
    d_og = q.shape[-1]
    if d_og > 256:
        raise RuntimeError("FlashAttention forward only supports head dimension at most 256")
    if q.shape[-2] % k.shape[-2] != 0:
        raise RuntimeError("Number of heads in key/value must divide number of heads in query")
    d = round_up(d_og, 8)
    q_p, k_p, v_p = [_pad_head(x, d) for x in (q, k, v)]
    p = new_params(q.dtype)
    keep = []
    if varlen:
        total_q, h, _ = q_p.shape
        h_k = k_p.shape[1]
        b = cu_seqlens_q.numel() - 1
        sq, sk = int(max_seqlen_q), int(max_seqlen_k)
        p.cu_seqlens_q, p.cu_seqlens_k = cu_seqlens_q.data_ptr(), cu_seqlens_k.data_ptr()
        p.total_q, p.total_k = total_q, k_p.shape[0]
        out_p = torch.empty_like(q_p)
    else:
        b, sq, h, _ = q_p.shape
        sk, h_k = k_p.shape[1], k_p.shape[2]
        out_p = torch.empty_like(q_p)
    if seqused_k is not None:
        p.seqused_k = seqused_k.data_ptr()
    lse = torch.empty((b, h, sq), dtype=torch.float32, device=q.device)
    p.q, p.k, p.v, p.o = [tdesc(t, packed=varlen) for t in (q_p, k_p, v_p, out_p)]
    p.softmax_lse = lse.data_ptr()
    p.b, p.seqlen_q, p.seqlen_k, p.h, p.h_k, p.d = b, sq, sk, h, h_k, d
    p.seqlen_q_rounded, p.seqlen_k_rounded, p.d_rounded = round_up(sq, 128), round_up(sk, 128), round_up(d, 32)
    p.softmax_scale = softmax_scale
    p.is_causal = 1 if causal else 0
    _fill_mask(p, mask, b, sk, keep)
    rng_state = S_dmask = None
    if dropout_p > 0.0:
        rng_state = _new_rng_state(q.device)
        p.p_dropout = dropout_p
        p.philox_seed, p.philox_offset = int(rng_state[0]), int(rng_state[1])
        if return_softmax:   # (b, h, sq_rounded, sk_rounded), probabilities negated where dropped (flash_api.cpp:276-279)
            S_dmask = torch.empty((b, h, p.seqlen_q_rounded, p.seqlen_k_rounded), dtype=q.dtype, device=q.device)
            p.s_dmask = S_dmask.data_ptr()
    with torch.cuda.device(q.device):
        _capi.check(_capi.load().fa_b200_fwd(p, _capi.stream_ptr(q.device)), "fa_b200_fwd")
    out = out_p[..., :d_og] if d != d_og else out_p
    return out, q_p, k_p, v_p, out_p, lse, S_dmask, rng_state


def _flash_attn_backward(dout, q, k, v, out, softmax_lse, dq, dk, dv, dropout_p, softmax_scale, causal,
                         rng_state=None, mask: _Mask = _NO_MASK, cu_seqlens_q=None, cu_seqlens_k=None,
                         max_seqlen_q=None, max_seqlen_k=None, deterministic=False):
    """Equivalent of ``flash_attn_2_cuda.bwd`` / ``varlen_bwd`` (csrc/flash_attn/flash_api.cpp:493-694,696-912).

    ``q, k, v, out`` are the (head-padded) tensors saved by forward; ``dq, dk, dv`` are written in place
    (they may be strided slices of a packed gradient).  GQA/MQA: dk/dv have ``h_k`` heads -- the group
    reduction happens inside the kernels, not with an expanded buffer + ``sum`` as in the reference (:628-635,682-686).
    """
    dout, q, k, v, out = [_maybe_contiguous(x) for x in (dout, q, k, v, out)]
    varlen = cu_seqlens_q is not None
    d = q.shape[-1]
    dout = _pad_head(dout, d)
    p = new_params(q.dtype)
    keep = []
    if varlen:
        total_q, h, _ = q.shape
        h_k = k.shape[1]
        b = cu_seqlens_q.numel() - 1
        sq, sk = int(max_seqlen_q), int(max_seqlen_k)
        p.cu_seqlens_q, p.cu_seqlens_k = cu_seqlens_q.data_ptr(), cu_seqlens_k.data_ptr()
        p.total_q, p.total_k = total_q, k.shape[0]
    else:
        b, sq, h, _ = q.shape
        sk, h_k = k.shape[1], k.shape[2]
    sq_r, sk_r, d_r = round_up(sq, 128), round_up(sk, 128), round_up(d, 32)
    softmax_d = torch.empty((b, h, sq_r), dtype=torch.float32, device=q.device)
    # fp32 dQ accumulator: batch mode (b, h, sq_r, d_r); varlen mode h * (total_q + 128 * b) rows (one tile of
    # slack per sequence so tile-granular reduce-adds never cross sequences)
    acc_rows = h * round_up(total_q + 128 * b, 128) if varlen else b * h * sq_r
    dq_accum = torch.empty(acc_rows * d_r, dtype=torch.float32, device=q.device)
    # GQA: one CTA normally owns (key block, kv-head) and walks the q-heads of the group, so dK/dV need no reduction.
    # When that grid is only a few waves deep and the last wave is mostly empty (ring steps, batch 1 per GPU: 512 CTAs
    # = 3.46 waves on 148 SMs), one CTA per (key block, Q-head) fills the machine better; the per-q-head dK/dV are then
    # summed over the group afterwards -- the reference's only mode (csrc/flash_attn/flash_api.cpp:628-635,682-686).
    dk_out, dv_out = dk, dv
    expand = h != h_k and d <= 128 and _prefer_expanded_heads(b, sk, h, h_k, q.device)
    if expand:
        dk = torch.empty(k.shape[:-2] + (h, d), dtype=k.dtype, device=k.device)
        dv = torch.empty_like(dk)
        p.dkv_expanded = 1
    p.q, p.k, p.v, p.o, p.dout, p.dq, p.dk, p.dv = [tdesc(t, packed=varlen) for t in (q, k, v, out, dout, dq, dk, dv)]
    p.softmax_lse, p.softmax_d, p.dq_accum = softmax_lse.data_ptr(), softmax_d.data_ptr(), dq_accum.data_ptr()
    # no dk/dv accumulator: the kernels reduce a GQA group inside one CTA (TMEM), p.dkv_accum stays NULL
    p.b, p.seqlen_q, p.seqlen_k, p.h, p.h_k, p.d = b, sq, sk, h, h_k, d
    p.seqlen_q_rounded, p.seqlen_k_rounded, p.d_rounded = sq_r, sk_r, d_r
    p.softmax_scale = softmax_scale
    p.is_causal = 1 if causal else 0
    p.deterministic = 1 if deterministic else 0
    _fill_mask(p, mask, b, sk, keep)
    if dropout_p > 0.0:
        if rng_state is None:
            raise RuntimeError("backward with dropout needs the rng_state saved by forward")
        p.p_dropout = dropout_p
        p.philox_seed, p.philox_offset = int(rng_state[0]), int(rng_state[1])
    with torch.cuda.device(q.device):
        _capi.check(_capi.load().fa_b200_bwd(p, _capi.stream_ptr(q.device)), "fa_b200_bwd")
    if expand:
        g = h // h_k
        torch.sum(dk.view(dk.shape[:-2] + (h_k, g, d)), dim=-2, out=dk_out) if dk_out.is_contiguous() else \
            dk_out.copy_(dk.view(dk.shape[:-2] + (h_k, g, d)).sum(dim=-2))
        torch.sum(dv.view(dv.shape[:-2] + (h_k, g, d)), dim=-2, out=dv_out) if dv_out.is_contiguous() else \
            dv_out.copy_(dv.view(dv.shape[:-2] + (h_k, g, d)).sum(dim=-2))
    return dq, dk_out, dv_out, softmax_d


def _flash_attn_varlen_forward(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p,
                               softmax_scale, causal, return_softmax, mask: _Mask = _NO_MASK):
    return _flash_attn_forward(q, k, v, dropout_p, softmax_scale, causal, return_softmax, mask,
                               cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k)


def _flash_attn_varlen_backward(dout, q, k, v, out, softmax_lse, dq, dk, dv, cu_seqlens_q, cu_seqlens_k,
                                max_seqlen_q, max_seqlen_k, dropout_p, softmax_scale, causal, rng_state=None,
                                mask: _Mask = _NO_MASK):
    return _flash_attn_backward(dout, q, k, v, out, softmax_lse, dq, dk, dv, dropout_p, softmax_scale, causal,
                                rng_state, mask, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k)


# ----------------------------------------------------------------------------------------------------
# autograd Functions -- same six classes as the reference (flash_attn_interface.py:88-279)
# ----------------------------------------------------------------------------------------------------
class FlashAttnQKVPackedFunc(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, dropout_p, softmax_scale, causal, return_softmax, mask=_NO_MASK):
        if softmax_scale is None:
            softmax_scale = qkv.shape[-1] ** (-0.5)
        out, q, k, v, out_padded, softmax_lse, S_dmask, rng_state = _flash_attn_forward(
            qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], dropout_p, softmax_scale, causal=causal,
            return_softmax=return_softmax and dropout_p > 0, mask=mask)
        ctx.rng_state = rng_state
        ctx.save_for_backward(q, k, v, out_padded, softmax_lse)
        ctx.dropout_p, ctx.softmax_scale, ctx.causal, ctx.mask = dropout_p, softmax_scale, causal, mask
        return out if not return_softmax else (out, softmax_lse, S_dmask)

    @staticmethod
    def backward(ctx, dout, *args):
        q, k, v, out, softmax_lse = ctx.saved_tensors
        qkv_shape = q.shape[:-2] + (3, *q.shape[-2:])
        dqkv = torch.empty(qkv_shape, dtype=q.dtype, device=q.device)
        _flash_attn_backward(dout, q, k, v, out, softmax_lse, dqkv[:, :, 0], dqkv[:, :, 1], dqkv[:, :, 2],
                             ctx.dropout_p, ctx.softmax_scale, ctx.causal, mask=ctx.mask, rng_state=ctx.rng_state)
        dqkv = dqkv[..., :dout.shape[-1]]  # the head dimension may have been padded
        return dqkv, None, None, None, None, None


class FlashAttnVarlenQKVPackedFunc(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, cu_seqlens, max_seqlen, dropout_p, softmax_scale, causal, return_softmax, mask=_NO_MASK):
        if softmax_scale is None:
            softmax_scale = qkv.shape[-1] ** (-0.5)
        out, q, k, v, out_padded, softmax_lse, S_dmask, rng_state = _flash_attn_varlen_forward(
            qkv[:, 0], qkv[:, 1], qkv[:, 2], cu_seqlens, cu_seqlens, max_seqlen, max_seqlen, dropout_p,
            softmax_scale, causal=causal, return_softmax=return_softmax and dropout_p > 0, mask=mask)
        ctx.rng_state = rng_state
        ctx.save_for_backward(q, k, v, out_padded, softmax_lse, cu_seqlens)
        ctx.dropout_p, ctx.max_seqlen, ctx.softmax_scale, ctx.causal, ctx.mask = \
            dropout_p, max_seqlen, softmax_scale, causal, mask
        return out if not return_softmax else (out, softmax_lse, S_dmask)

    @staticmethod
    def backward(ctx, dout, *args):
        q, k, v, out, softmax_lse, cu_seqlens = ctx.saved_tensors
        qkv_shape = q.shape[:-2] + (3, *q.shape[-2:])
        dqkv = torch.empty(qkv_shape, dtype=q.dtype, device=q.device)
        _flash_attn_varlen_backward(dout, q, k, v, out, softmax_lse, dqkv[:, 0], dqkv[:, 1], dqkv[:, 2],
                                    cu_seqlens, cu_seqlens, ctx.max_seqlen, ctx.max_seqlen, ctx.dropout_p,
                                    ctx.softmax_scale, ctx.causal, mask=ctx.mask, rng_state=ctx.rng_state)
        dqkv = dqkv[..., :dout.shape[-1]]
        return dqkv, None, None, None, None, None, None, None


class FlashAttnKVPackedFunc(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, kv, dropout_p, softmax_scale, causal, return_softmax, mask=_NO_MASK):
        if softmax_scale is None:
            softmax_scale = q.shape[-1] ** (-0.5)
        out, q, k, v, out_padded, softmax_lse, S_dmask, rng_state = _flash_attn_forward(
            q, kv[:, :, 0], kv[:, :, 1], dropout_p, softmax_scale, causal=causal,
            return_softmax=return_softmax and dropout_p > 0, mask=mask)
        ctx.rng_state = rng_state
        ctx.save_for_backward(q, k, v, out_padded, softmax_lse)
        ctx.dropout_p, ctx.softmax_scale, ctx.causal, ctx.mask = dropout_p, softmax_scale, causal, mask
        return out if not return_softmax else (out, softmax_lse, S_dmask)

    @staticmethod
    def backward(ctx, dout, *args):
        q, k, v, out, softmax_lse = ctx.saved_tensors
        dq = torch.empty_like(q)
        kv_shape = k.shape[:-2] + (2, *k.shape[-2:])
        dkv = torch.empty(kv_shape, dtype=k.dtype, device=k.device)
        _flash_attn_backward(dout, q, k, v, out, softmax_lse, dq, dkv[:, :, 0], dkv[:, :, 1], ctx.dropout_p,
                             ctx.softmax_scale, ctx.causal, mask=ctx.mask, rng_state=ctx.rng_state)
        dq = dq[..., :dout.shape[-1]]
        dkv = dkv[..., :dout.shape[-1]]
        return dq, dkv, None, None, None, None, None


class FlashAttnVarlenKVPackedFunc(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, kv, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p, softmax_scale,
                causal, return_softmax, mask=_NO_MASK):
        if softmax_scale is None:
            softmax_scale = q.shape[-1] ** (-0.5)
        out, q, k, v, out_padded, softmax_lse, S_dmask, rng_state = _flash_attn_varlen_forward(
            q, kv[:, 0], kv[:, 1], cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p,
            softmax_scale, causal=causal, return_softmax=return_softmax and dropout_p > 0, mask=mask)
        ctx.rng_state = rng_state
        ctx.save_for_backward(q, k, v, out_padded, softmax_lse, cu_seqlens_q, cu_seqlens_k)
        ctx.dropout_p, ctx.max_seqlen_q, ctx.max_seqlen_k = dropout_p, max_seqlen_q, max_seqlen_k
        ctx.softmax_scale, ctx.causal, ctx.mask = softmax_scale, causal, mask
        return out if not return_softmax else (out, softmax_lse, S_dmask)

    @staticmethod
    def backward(ctx, dout, *args):
        q, k, v, out, softmax_lse, cu_seqlens_q, cu_seqlens_k = ctx.saved_tensors
        dq = torch.empty_like(q)
        kv_shape = k.shape[:-2] + (2, *k.shape[-2:])
        dkv = torch.empty(kv_shape, dtype=k.dtype, device=k.device)
        _flash_attn_varlen_backward(dout, q, k, v, out, softmax_lse, dq, dkv[:, 0], dkv[:, 1], cu_seqlens_q,
                                    cu_seqlens_k, ctx.max_seqlen_q, ctx.max_seqlen_k, ctx.dropout_p,
                                    ctx.softmax_scale, ctx.causal, mask=ctx.mask, rng_state=ctx.rng_state)
        dq = dq[..., :dout.shape[-1]]
        dkv = dkv[..., :dout.shape[-1]]
        return dq, dkv, None, None, None, None, None, None, None, None, None


class FlashAttnFunc(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, dropout_p, softmax_scale, causal, return_softmax, mask=_NO_MASK):
        if softmax_scale is None:
            softmax_scale = q.shape[-1] ** (-0.5)
        out, q, k, v, out_padded, softmax_lse, S_dmask, rng_state = _flash_attn_forward(
            q, k, v, dropout_p, softmax_scale, causal=causal, return_softmax=return_softmax and dropout_p > 0,
            mask=mask)
        ctx.rng_state = rng_state
        ctx.save_for_backward(q, k, v, out_padded, softmax_lse)
        ctx.dropout_p, ctx.softmax_scale, ctx.causal, ctx.mask = dropout_p, softmax_scale, causal, mask
        return out if not return_softmax else (out, softmax_lse, S_dmask)

    @staticmethod
    def backward(ctx, dout, *args):
        q, k, v, out, softmax_lse = ctx.saved_tensors
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        _flash_attn_backward(dout, q, k, v, out, softmax_lse, dq, dk, dv, ctx.dropout_p, ctx.softmax_scale,
                             ctx.causal, mask=ctx.mask, rng_state=ctx.rng_state)
        dq = dq[..., :dout.shape[-1]]
        dk = dk[..., :dout.shape[-1]]
        dv = dv[..., :dout.shape[-1]]
        return dq, dk, dv, None, None, None, None, None


class FlashAttnVarlenFunc(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p, softmax_scale,
                causal, return_softmax, mask=_NO_MASK):
        if softmax_scale is None:
            softmax_scale = q.shape[-1] ** (-0.5)
        out, q, k, v, out_padded, softmax_lse, S_dmask, rng_state = _flash_attn_varlen_forward(
            q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p, softmax_scale,
            causal=causal, return_softmax=return_softmax and dropout_p > 0, mask=mask)
        ctx.rng_state = rng_state
        ctx.save_for_backward(q, k, v, out_padded, softmax_lse, cu_seqlens_q, cu_seqlens_k)
        ctx.dropout_p, ctx.max_seqlen_q, ctx.max_seqlen_k = dropout_p, max_seqlen_q, max_seqlen_k
        ctx.softmax_scale, ctx.causal, ctx.mask = softmax_scale, causal, mask
        return out if not return_softmax else (out, softmax_lse, S_dmask)

    @staticmethod
    def backward(ctx, dout, *args):
        q, k, v, out, softmax_lse, cu_seqlens_q, cu_seqlens_k = ctx.saved_tensors
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        _flash_attn_varlen_backward(dout, q, k, v, out, softmax_lse, dq, dk, dv, cu_seqlens_q, cu_seqlens_k,
                                    ctx.max_seqlen_q, ctx.max_seqlen_k, ctx.dropout_p, ctx.softmax_scale,
                                    ctx.causal, mask=ctx.mask, rng_state=ctx.rng_state)
        dq = dq[..., :dout.shape[-1]]
        dk = dk[..., :dout.shape[-1]]
        dv = dv[..., :dout.shape[-1]]
        return dq, dk, dv, None, None, None, None, None, None, None, None, None


# ----------------------------------------------------------------------------------------------------
# public functions (flash_attn_interface.py:282-492) -- positional order and defaults as the reference
# ----------------------------------------------------------------------------------------------------
def _mk_mask(window_size, flashmask, force_generic=False):
    if tuple(window_size) == (-1, -1) and flashmask is None and not force_generic:
        return _NO_MASK
    return _Mask(window=window_size, flashmask=flashmask, force_generic=force_generic)


def flash_attn_qkvpacked_func(qkv, dropout_p=0.0, softmax_scale=None, causal=False, return_attn_probs=False, *,
                              window_size=(-1, -1), flashmask=None, _force_generic=False):
    """qkv: (batch_size, seqlen, 3, nheads, headdim).  Returns out (batch_size, seqlen, nheads, headdim)
    [+ softmax_lse (batch_size, nheads, seqlen) fp32, S_dmask] (reference :282-311)."""
    return FlashAttnQKVPackedFunc.apply(qkv, dropout_p, softmax_scale, causal, return_attn_probs,
                                        _mk_mask(window_size, flashmask, _force_generic))


def flash_attn_kvpacked_func(q, kv, dropout_p=0.0, softmax_scale=None, causal=False, return_attn_probs=False, *,
                             window_size=(-1, -1), flashmask=None, _force_generic=False):
    """q: (batch_size, seqlen_q, nheads, headdim); kv: (batch_size, seqlen_k, 2, nheads_k, headdim) with
    nheads % nheads_k == 0 (MQA/GQA) (reference :314-342)."""
    return FlashAttnKVPackedFunc.apply(q, kv, dropout_p, softmax_scale, causal, return_attn_probs,
                                       _mk_mask(window_size, flashmask, _force_generic))


def flash_attn_func(q, k, v, dropout_p=0.0, softmax_scale=None, causal=False, return_attn_probs=False, *,
                    window_size=(-1, -1), flashmask=None, _force_generic=False):
    """q: (batch_size, seqlen_q, nheads, headdim); k, v: (batch_size, seqlen_k, nheads_k, headdim)
    (reference :345-373).  ``causal`` is top-left aligned, as in the reference."""
    return FlashAttnFunc.apply(q, k, v, dropout_p, softmax_scale, causal, return_attn_probs,
                               _mk_mask(window_size, flashmask, _force_generic))


def flash_attn_varlen_qkvpacked_func(qkv, cu_seqlens, max_seqlen, dropout_p=0.0, softmax_scale=None, causal=False,
                                     return_attn_probs=False, *, _force_generic=False):
    """qkv: (total, 3, nheads, headdim); cu_seqlens: (batch_size + 1,) int32 (reference :376-411)."""
    return FlashAttnVarlenQKVPackedFunc.apply(qkv, cu_seqlens, max_seqlen, dropout_p, softmax_scale, causal,
                                              return_attn_probs, _mk_mask((-1, -1), None, _force_generic))


def flash_attn_varlen_kvpacked_func(q, kv, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p=0.0,
                                    softmax_scale=None, causal=False, return_attn_probs=False, *,
                                    _force_generic=False):
    """q: (total_q, nheads, headdim); kv: (total_k, 2, nheads_k, headdim) (reference :414-451)."""
    return FlashAttnVarlenKVPackedFunc.apply(q, kv, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k,
                                             dropout_p, softmax_scale, causal, return_attn_probs,
                                             _mk_mask((-1, -1), None, _force_generic))


def flash_attn_varlen_func(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p=0.0,
                           softmax_scale=None, causal=False, return_attn_probs=False, *, window_size=(-1, -1),
                           _force_generic=False):
    """q: (total_q, nheads, headdim); k, v: (total_k, nheads_k, headdim); cu_seqlens_*: (batch_size + 1,) int32
    (reference :454-492)."""
    return FlashAttnVarlenFunc.apply(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p,
                                     softmax_scale, causal, return_attn_probs,
                                     _mk_mask(window_size, None, _force_generic))


def flashmask_attention(q, k, v, startend_row_indices, dropout_p=0.0, softmax_scale=None, causal=False, *,
                        _force_generic=False):
    """The fork's FlashMask operator (Paddle ``flashmask_attention``; kernels in csrc/flashmask_v2).

    ``startend_row_indices``: int32 (batch, mask_heads, seqlen_k, {1, 2, 4}) holding, per key column,
    [lt_start], [lt_start, lt_end] (causal) / [lt_start, ut_end] (non-causal), or
    [lt_start, lt_end, ut_start, ut_end].  Query row i is masked for column j iff
    lt_start[j] <= i < lt_end[j] or ut_start[j] <= i < ut_end[j].
    """
    n = startend_row_indices.shape[-1]
    cols = [startend_row_indices[..., i].contiguous() for i in range(n)]
    if n == 1:
        fm = (cols[0], None, None, None)
    elif n == 2:
        fm = (cols[0], cols[1], None, None) if causal else (cols[0], None, None, cols[1])
    elif n == 4:
        fm = (cols[0], cols[1], cols[2], cols[3])
    else:
        raise ValueError("startend_row_indices last dim must be 1, 2 or 4")
    return FlashAttnFunc.apply(q, k, v, dropout_p, softmax_scale, causal, False,
                               _Mask(flashmask=fm, force_generic=_force_generic))


def flash_attn_with_kvcache(q, k_cache, v_cache, k=None, v=None, rotary_cos=None, rotary_sin=None,
                            cache_seqlens=None, cache_batch_idx=None, cache_leftpad=None, block_table=None,
                            softmax_scale=None, causal=False, window_size=(-1, -1), softcap=0.0,
                            rotary_interleaved=True, alibi_slopes=None, num_splits=0, return_softmax_lse=False):
    """Decode-time attention over a KV cache, upstream signature (the fork only has the capability at C level:
    csrc/flash_attn_v3/flash.h:96-125 knew_ptr / seqused_k / num_splits / oaccum_ptr, plus the combine kernel).

    q: (batch, seqlen_q, nheads, headdim); k_cache, v_cache: (batch, seqlen_cache, nheads_k, headdim).
    If ``k``/``v`` are given they are appended IN PLACE at ``cache_seqlens`` before attention.  ``causal`` and
    ``window_size`` are bottom-right aligned (csrc/flashmask_v2/mask.h:85-118).  ``num_splits=0`` picks the split
    count (decode shapes: the wave model of ``fa_b200_decode_num_splits``; otherwise the reference heuristic,
    csrc/flashmask_v2/heuristics.h:22-59); partial results are merged by the
    LSE-combine kernel.  Inference only (no autograd).

    Decode extras (csrc/flash_attn_v3/flash.h:96-140): ``block_table`` (paged cache, k/v_cache
    (num_pages, page_size, nheads_k, headdim)), ``cache_batch_idx``, ``cache_leftpad`` (valid keys are
    [leftpad, cache_seqlens)), ``rotary_cos/sin`` applied to ``k`` (and ``q``) before the append, ``softcap``,
    ``alibi_slopes``.  The TMA decode kernel takes all of them when page_size % 64 == 0; other shapes use the
    generic kernels.
    """
    q, k_cache, v_cache = [_maybe_contiguous(x) for x in (q, k_cache, v_cache)]
    b, sq, h, d = q.shape
    paged = block_table is not None
    if paged:
        if cache_batch_idx is not None:
            raise ValueError("flash_attn_with_kvcache: block_table and cache_batch_idx are exclusive")
        block_table = block_table.to(torch.int32).contiguous()
        num_pages, page_size, h_k = k_cache.shape[0], k_cache.shape[1], k_cache.shape[2]
        sk = block_table.shape[1] * page_size
    else:
        sk, h_k = k_cache.shape[1], k_cache.shape[2]
    if d % 8 != 0:
        raise RuntimeError("flash_attn_with_kvcache: head dimension must be a multiple of 8")
    if softmax_scale is None:
        softmax_scale = d ** (-0.5)
    lib = _capi.load()
    stream = _capi.stream_ptr(q.device)
    is_bf16 = 1 if q.dtype == torch.bfloat16 else 0
    if cache_seqlens is not None and isinstance(cache_seqlens, int):
        cache_seqlens = torch.full((b,), cache_seqlens, dtype=torch.int32, device=q.device)
    if cache_seqlens is not None:
        cache_seqlens = cache_seqlens.to(torch.int32).contiguous()
    keep = []   # tensors whose device pointers sit in `p` until the launches are enqueued
    with torch.cuda.device(q.device):
        p = new_params(q.dtype)
        p.k, p.v = tdesc(k_cache), tdesc(v_cache)
        p.b, p.seqlen_q, p.seqlen_k, p.h, p.h_k, p.d = b, sq, sk, h, h_k, d
        p.seqlen_q_rounded, p.seqlen_k_rounded, p.d_rounded = round_up(sq, 128), round_up(sk, 128), round_up(d, 32)
        if paged:
            p.page_table, p.page_table_batch_stride = block_table.data_ptr(), block_table.stride(0)
            p.page_size, p.num_pages = page_size, num_pages
        if cache_batch_idx is not None:
            cache_batch_idx = cache_batch_idx.to(torch.int32).contiguous()
            p.kv_batch_idx, p.b_k = cache_batch_idx.data_ptr(), k_cache.shape[0]
        if cache_leftpad is not None:
            cache_leftpad = cache_leftpad.to(torch.int32).contiguous()
            p.leftpad_k = cache_leftpad.data_ptr()
        local = tuple(window_size) != (-1, -1)
        if rotary_cos is not None:
            # k_new is rotated at positions cache_seqlens + i; q at the same positions when causal / local, else every
            # query row at cache_seqlens (csrc/flash_attn_v3/mainloop_fwd_sm90_tma_gmma_ws.hpp:1090,1582; rotary.h)
            if k is None or rotary_sin is None:
                raise ValueError("flash_attn_with_kvcache: rotary_cos / rotary_sin need k and v")
            cos = rotary_cos.to(q.dtype).contiguous()
            sin = rotary_sin.to(q.dtype).contiguous()
            rdim = 2 * cos.shape[-1]

            def rot(x, fixed):
                y = torch.empty_like(x)
                _capi.check(lib.fa_b200_rotary(
                    x.data_ptr(), y.data_ptr(), cos.data_ptr(), sin.data_ptr(), ptr(cache_seqlens), x.shape[0],
                    x.shape[1], x.shape[2], d, rdim, cos.shape[0], 1 if rotary_interleaved else 0, 1 if fixed else 0,
                    is_bf16, x.stride(0), x.stride(1), x.stride(2), y.stride(0), y.stride(1), y.stride(2), stream),
                    "fa_b200_rotary")
                return y
            q = rot(q, not (causal or local))
            k = rot(_maybe_contiguous(k), False)
        seqused = cache_seqlens
        if k is not None:
            assert v is not None
            k, v = _maybe_contiguous(k), _maybe_contiguous(v)
            if v.stride() != k.stride():
                v = v.contiguous(); k = k.contiguous()
            s_new = k.shape[1]
            p.seqused_k = ptr(cache_seqlens)        # lengths before the append
            _capi.check(lib.fa_b200_append_kv_paged(p, k.data_ptr(), v.data_ptr(), s_new, k.stride(0), k.stride(1),
                                                    k.stride(2), stream), "fa_b200_append_kv_paged")
            seqused = (cache_seqlens + s_new) if cache_seqlens is not None else \
                torch.full((b,), s_new, dtype=torch.int32, device=q.device)
        out = torch.empty_like(q)
        lse = torch.empty((b, h, sq), dtype=torch.float32, device=q.device)
        p.q, p.o = tdesc(q), tdesc(out)
        p.softmax_lse = lse.data_ptr()
        p.softmax_scale = softmax_scale
        p.is_causal = 1 if causal else 0
        p.causal_bottom_right = 1
        p.window_left, p.window_right = window_size
        p.lse_empty_is_neg_inf = 1
        p.softcap = float(softcap)
        if alibi_slopes is not None:
            alibi_slopes = alibi_slopes.to(device=q.device, dtype=torch.float32).contiguous()
            if alibi_slopes.shape not in ((h,), (b, h)):
                raise ValueError("flash_attn_with_kvcache: alibi_slopes must be (nheads,) or (batch, nheads)")
            p.alibi_slopes = alibi_slopes.data_ptr()
            p.alibi_batch_stride = h if alibi_slopes.dim() == 2 else 0
        p.seqused_k = None
        if seqused is not None:
            seqused = seqused.to(torch.int32).contiguous()
            p.seqused_k = seqused.data_ptr()
        keep += [block_table, cache_batch_idx, cache_leftpad, alibi_slopes, seqused, q]
        if num_splits <= 0:
            num_sm = torch.cuda.get_device_properties(q.device).multi_processor_count
            rows = sq * (h // h_k)                         # packed query rows per (batch, kv-head)
            special = paged or cache_batch_idx is not None or cache_leftpad is not None or softcap != 0.0 or \
                alibi_slopes is not None
            if rows <= 128 and d in (64, 128):             # the TMA-fed decode kernel (csrc/decode.cu), 16 rows per CTA
                num_splits = lib.fa_b200_decode_num_splits(b * ((rows + 15) // 16), h_k, sk, num_sm)
            elif d <= 128 and not special:                 # chunked prefill over a cache: one tcgen05 forward pass
                num_splits = 1
            else:
                num_splits = lib.fa_b200_num_splits(b, h_k, rows, sk, d, num_sm)
        p.num_splits = num_splits
        if num_splits > 1:
            oaccum = torch.empty((num_splits, b, sq, h, d), dtype=torch.float32, device=q.device)
            lseaccum = torch.empty((num_splits, b, h, sq), dtype=torch.float32, device=q.device)
            p.oaccum, p.lseaccum = oaccum.data_ptr(), lseaccum.data_ptr()
            _capi.check(lib.fa_b200_fwd_splitkv(p, stream), "fa_b200_fwd_splitkv")
            _capi.check(lib.fa_b200_fwd_combine(p, stream), "fa_b200_fwd_combine")
        else:
            _capi.check(lib.fa_b200_fwd_splitkv(p, stream), "fa_b200_fwd_splitkv")
    return (out, lse) if return_softmax_lse else out


def flash_attn_fp8_func(q, k, v, q_descale=None, k_descale=None, v_descale=None, softmax_scale=None, causal=False, *,
                        window_size=(-1, -1), cu_seqlens_q=None, cu_seqlens_k=None, max_seqlen_q=None, max_seqlen_k=None):
    """FP8 forward: ``q, k, v`` are ``torch.float8_e4m3fn`` tensors, ``*_descale`` optional fp32 ``(batch, heads_k)``
    dequantisation factors; returns ``(out, softmax_lse)`` with ``out`` in bf16.  Forward only.

    The fork has this capability at C level only (``is_e4m3`` + ``*_descale_ptr`` of the FlashMask-v2 / FA3 handle,
    csrc/flashmask_v2/flash.h:55-63,144; dispatch csrc/flashmask_v2/flash_api.cu:107-131): logits are
    ``softmax_scale * q_descale * k_descale * (q8 . k8)``, the weights are rounded to e4m3 for the P V product, the
    result is multiplied by ``v_descale`` (mainloop_fwd_sm90_tma_gmma_ws.hpp:1612-1614,1868-1869).  Causal masks are
    bottom-right aligned as everywhere on that surface.  Fixed-length ``(b, s, h, d)`` or packed ``(total, h, d)`` with
    ``cu_seqlens_*``; head dim a multiple of 16 up to 256.
    """
    if q.dtype != torch.float8_e4m3fn or k.dtype != q.dtype or v.dtype != q.dtype:
        raise RuntimeError("flash_attn_fp8_func takes float8_e4m3fn q, k, v")
    q, k, v = [_maybe_contiguous(x) for x in (q, k, v)]
    varlen = cu_seqlens_q is not None
    d = q.shape[-1]
    if d % 16 != 0 or d > 256:
        raise RuntimeError("FP8 forward needs a head dimension that is a multiple of 16, at most 256")
    p = new_params(torch.bfloat16)
    p.is_e4m3 = 1
    if varlen:
        total_q, h, _ = q.shape
        h_k = k.shape[1]
        b = cu_seqlens_q.numel() - 1
        sq, sk = int(max_seqlen_q), int(max_seqlen_k)
        p.cu_seqlens_q, p.cu_seqlens_k = cu_seqlens_q.data_ptr(), cu_seqlens_k.data_ptr()
        p.total_q, p.total_k = total_q, k.shape[0]
        out = torch.empty((total_q, h, d), dtype=torch.bfloat16, device=q.device)
    else:
        b, sq, h, _ = q.shape
        sk, h_k = k.shape[1], k.shape[2]
        out = torch.empty((b, sq, h, d), dtype=torch.bfloat16, device=q.device)
    if h % h_k != 0:
        raise RuntimeError("Number of heads in key/value must divide number of heads in query")
    lse = torch.empty((b, h, sq), dtype=torch.float32, device=q.device)
    p.q, p.k, p.v, p.o = [tdesc(t, packed=varlen) for t in (q, k, v, out)]
    p.softmax_lse = lse.data_ptr()
    p.b, p.seqlen_q, p.seqlen_k, p.h, p.h_k, p.d = b, sq, sk, h, h_k, d
    p.seqlen_q_rounded, p.seqlen_k_rounded, p.d_rounded = round_up(sq, 128), round_up(sk, 128), round_up(d, 32)
    p.softmax_scale = d ** -0.5 if softmax_scale is None else softmax_scale
    p.is_causal = 1 if causal else 0
    p.causal_bottom_right = 1
    p.lse_empty_is_neg_inf = 1
    p.window_left, p.window_right = int(window_size[0]), int(window_size[1])
    keep = []
    for name, t in (("q", q_descale), ("k", k_descale), ("v", v_descale)):
        if t is None:
            continue
        if t.dtype != torch.float32 or tuple(t.shape) != (b, h_k):
            raise RuntimeError(f"{name}_descale must be fp32 of shape (batch, heads_k)")
        keep.append(t)
        setattr(p, f"{name}_descale", t.data_ptr())
        setattr(p, f"{name}_descale_batch_stride", t.stride(0))
        setattr(p, f"{name}_descale_head_stride", t.stride(1))
    with torch.cuda.device(q.device):
        _capi.check(_capi.load().fa_b200_fwd(p, _capi.stream_ptr(q.device)), "fa_b200_fwd")
    del keep
    return out, lse
