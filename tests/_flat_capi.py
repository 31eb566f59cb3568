"""ctypes callers for the flat FA2 C ABI (include/fa_capi.h == the reference's csrc/capi/flash_attn.h).

The same functions drive ``libflashattn_b200.so`` and -- in tests/test_ref_kernel_gpu.py -- the reference's own kernels
built by oracle/build_ref.sh (``oracle/_ref/libflashattn_ref.so``): the two libraries export identical signatures, so a
caller that works on one works on the other.  Argument order follows csrc/capi/flash_attn.h:11-55 (fwd), :100-161 (bwd),
:223-247 (calc_reduced_attn_scores); note the reference's FlashMask order: downstart, dims, upend, downend, upstart,
maxmin.
"""
import ctypes as C

import torch


def _vp(t):
    return C.c_void_p(t.data_ptr() if t is not None else None)


def _strides(t):
    """(row, head, batch) element strides of a (b, s, h, d) tensor."""
    return C.c_int(t.stride(1)), C.c_int(t.stride(2)), C.c_int(t.stride(0))


def round_up(x, m):
    return (x + m - 1) // m * m


def _i64(vals):
    arr = (C.c_int64 * len(vals))(*vals)
    return arr


def flashmask_args(fm, b, sk, device):
    """fm = dict(lts=, lte=, uts=, ute=) of int32 (b, h_mask, sk) tensors (lts mandatory) -> the six C arguments
    (downstart, dims, upend, downend, upstart, maxmin) plus the objects to keep alive."""
    if fm is None:
        return (None, None, None, None, None, None), ()
    lts = fm["lts"]
    h_mask = lts.shape[1]
    dims = _i64([b, h_mask, sk])
    nblock = (((sk + 31) // 32) + 3) & ~3        # csrc/flash_attn/src/mask.h:68-79, sized for the smallest kBlockN (32)
    ws = torch.empty(8 * b * h_mask * nblock, dtype=torch.int32, device=device)
    args = (_vp(lts), dims, _vp(fm.get("ute")), _vp(fm.get("lte")), _vp(fm.get("uts")), _vp(ws))
    return args, (dims, ws)


def flat_fwd(lib, q, k, v, causal, scale=None, fm=None, attn_mask=None, err=None, p_dropout=0.0, return_softmax=False,
             seed=0, offset=0):
    """flash_attn_fwd on (b, s, h, d) tensors.  Returns (out, lse), or (out, lse, S_dmask) with return_softmax --
    S_dmask (b, h, sq_rounded, sk_rounded) in the element type, csrc/flash_attn/flash_api.cpp:276-279."""
    b, sq, h, d = q.shape
    sk, hk = k.shape[1], k.shape[2]
    scale = d ** -0.5 if scale is None else scale
    out = torch.empty_like(q)
    lse = torch.empty(b, h, sq, device=q.device, dtype=torch.float32)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    (qr, qh, qb), (kr, kh, kb), (vr, vh, vb), (or_, oh, ob) = _strides(q), _strides(k), _strides(v), _strides(out)
    fm_args, keep = flashmask_args(fm, b, sk, q.device)
    mask_dims = _i64(list(attn_mask.shape)) if attn_mask is not None else None
    # the reference's forward kernel stores (seed, offset) into rng_state UNCONDITIONALLY (flash_fwd_kernel.h:320-324),
    # dropout or not: a caller must always pass two uint64 of device memory
    rng_state = torch.zeros(2, dtype=torch.int64, device=q.device)
    S = torch.full((b, h, round_up(sq, 128), round_up(sk, 128)), float("nan"), device=q.device, dtype=q.dtype) if return_softmax else None
    lib.flash_attn_fwd.restype = C.c_bool
    ok = lib.flash_attn_fwd(
        _vp(q), _vp(k), _vp(v), _vp(rng_state), _vp(out), _vp(S), _vp(lse), C.c_int(b), C.c_int(sq), C.c_int(sk),
        C.c_int(round_up(sq, 128)), C.c_int(round_up(sk, 128)), C.c_int(h), C.c_int(hk), C.c_int(d),
        C.c_int(round_up(d, 32)), C.c_float(p_dropout), C.c_float(scale), C.c_float(1.0 / scale), C.c_bool(causal),
        C.c_bool(return_softmax), C.c_bool(q.dtype == torch.bfloat16), stream, C.c_uint64(seed), C.c_uint64(offset),
        _vp(attn_mask), mask_dims, *fm_args, qr, kr, vr, qh, kh, vh, or_, oh, qb, kb, vb, ob)
    if not ok:
        lib.flash_attn_error.restype = C.c_char_p
        raise RuntimeError((lib.flash_attn_error() or b"").decode())
    del keep
    return (out, lse, S) if return_softmax else (out, lse)


def flat_bwd(lib, g, q, k, v, out, lse, causal, scale=None, fm=None, num_splits=0):
    """flash_attn_bwd; returns (dq, dk, dv) with dk/dv already reduced over the GQA group (the ABI hands back
    per-q-head dk/dv, csrc/capi/flash_attn.h:100-161; the caller sums, flash_api.cpp:628-635)."""
    b, sq, h, d = q.shape
    sk, hk = k.shape[1], k.shape[2]
    scale = d ** -0.5 if scale is None else scale
    sq_r, sk_r, d_r = round_up(sq, 128), round_up(sk, 128), round_up(d, 32)
    dq = torch.empty_like(q)
    dk = torch.empty(b, sk, h, d, device=q.device, dtype=q.dtype)
    dv = torch.empty(b, sk, h, d, device=q.device, dtype=q.dtype)
    softmax_d = torch.empty(b, h, sq_r, device=q.device, dtype=torch.float32)
    dq_accum = torch.zeros(b, h, sq_r, d_r, device=q.device, dtype=torch.float32)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    (qr, qh, qb), (kr, kh, kb), (vr, vh, vb), (or_, oh, ob) = _strides(q), _strides(k), _strides(v), _strides(out)
    (dqr, dqh, dqb), (dkr, dkh, dkb), (dvr, dvh, dvb), (gr, gh, gb) = _strides(dq), _strides(dk), _strides(dv), _strides(g)
    fm_args, keep = flashmask_args(fm, b, sk, q.device)
    rng_state = torch.zeros(2, dtype=torch.int64, device=q.device)
    lib.flash_attn_bwd.restype = C.c_bool
    ok = lib.flash_attn_bwd(
        _vp(g), _vp(q), _vp(k), _vp(v), _vp(out), _vp(softmax_d), _vp(lse), _vp(rng_state), _vp(dq), _vp(dk), _vp(dv),
        _vp(dq_accum), C.c_int(b), C.c_int(sq), C.c_int(sk), C.c_int(sq_r), C.c_int(sk_r), C.c_int(h), C.c_int(hk),
        C.c_int(d), C.c_int(d_r), C.c_float(0.0), C.c_float(scale), C.c_float(1.0 / scale), C.c_bool(causal),
        C.c_bool(q.dtype == torch.bfloat16), C.c_int(num_splits), stream, C.c_uint64(0), C.c_uint64(0), None, None,
        *fm_args, qr, kr, vr, qh, kh, vh, or_, oh, qb, kb, vb, ob, dqr, dkr, dvr, dqh, dkh, dvh, gr, gh, dqb, dkb, dvb, gb)
    if not ok:
        lib.flash_attn_error.restype = C.c_char_p
        raise RuntimeError((lib.flash_attn_error() or b"").decode())
    del keep
    g_ = h // hk
    dk = dk.view(b, sk, hk, g_, d).float().sum(3).to(q.dtype)
    dv = dv.view(b, sk, hk, g_, d).float().sum(3).to(q.dtype)
    return dq, dk, dv


def reduced_scores(lib, q, k, lse, scale=None, return_softmax=False):
    """calc_reduced_attn_scores (csrc/capi/flash_attn.h:223-247): fp32 (b, h, 1, sk) column sums of softmax(QK^T),
    ADDED to a zero-initialised buffer; with return_softmax also the fp32 probabilities (b, h, sq, sk)."""
    b, sq, h, d = q.shape
    sk, hk = k.shape[1], k.shape[2]
    scale = d ** -0.5 if scale is None else scale
    red = torch.zeros(b, h, 1, sk, device=q.device, dtype=torch.float32)
    probs = torch.zeros(b, h, sq, sk, device=q.device, dtype=torch.float32) if return_softmax else None
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    (qr, qh, qb), (kr, kh, kb) = _strides(q), _strides(k)
    lib.calc_reduced_attn_scores.restype = C.c_bool
    ok = lib.calc_reduced_attn_scores(
        _vp(q), _vp(k), _vp(lse), _vp(red), _vp(probs), C.c_int(b), C.c_int(sq), C.c_int(sk), C.c_int(h), C.c_int(hk),
        C.c_int(d), C.c_float(scale), C.c_bool(return_softmax), C.c_bool(q.dtype == torch.bfloat16), C.c_int(0), stream,
        qr, kr, C.c_int(sk), qh, kh, C.c_int(sk), qb, kb, C.c_int(h * sk))
    if not ok:
        lib.flash_attn_error.restype = C.c_char_p
        raise RuntimeError((lib.flash_attn_error() or b"").decode())
    return red, probs
