"""GPU parity for the FP8 (e4m3) forward -- the reference's ``is_e4m3`` path (csrc/flashmask_v2/flash_api.cu:107-131,
descale factors csrc/flashmask_v2/mainloop_fwd_sm90_tma_gmma_ws.hpp:1612-1614,1868-1869; sm90-only kernels, so no
reference binary can run here).

Oracle: the fp32 ``attention_masked_ref`` on the DEQUANTISED inputs (e4m3 value x descale): that is exact for the
Q K^T product (e4m3 products accumulate in fp32).  The only approximation the FP8 path adds on top of the 16-bit one is
the rounding of the weights P to e4m3 before P V (3 mantissa bits, relative step 2^-4); it is modelled by a second
oracle that rounds the unnormalised weights exp(s - max) to e4m3 the same way.  Tolerances (written here, as the task
asks): |out - exact| <= 2 x |emulated - exact| + 2e-2 * max|exact| (the emulation error is the yardstick, like the
reference's 2x rule), and LSE, which never sees the rounding, to 2e-3.
"""
import ctypes as C
import math

import pytest
import torch

from oracle import attention_oracle as orc

pytestmark = pytest.mark.gpu
DEV = "cuda"
E4M3_MAX = 448.0


def quantise(x, hk):
    """(b, s, h, d) fp32 -> (e4m3 tensor, descale (b, hk) fp32): one factor per (batch, kv-head group)."""
    b, s, h, d = x.shape
    g = h // hk
    amax = x.view(b, s, hk, g, d).abs().amax(dim=(1, 3, 4)).clamp_min(1e-6)        # (b, hk)
    descale = (amax / E4M3_MAX).float()
    scale = (1.0 / descale).repeat_interleave(g, dim=1).view(b, 1, h, 1)
    x8 = (x * scale).to(torch.float8_e4m3fn)
    return x8, descale.contiguous()


def dequant(x8, descale):
    b, s, h, d = x8.shape
    g = h // descale.shape[1]
    return x8.float() * descale.repeat_interleave(g, dim=1).view(b, 1, h, 1)


def mask_for(sq, sk, causal, window):
    m = None
    if causal:
        m = orc.causal_mask_bottom_right(sq, sk)
    if window != (-1, -1):
        w = orc.local_mask(sq, sk, window[0], window[1])
        m = w if m is None else (m | w)
    return m


def emulated_p_e4m3(q, k, v, mask, scale):
    """Same mathematics with the unnormalised weights exp(s - rowmax) rounded to e4m3 before the P V product."""
    b, sq, h, d = q.shape
    g = h // k.shape[2]
    kk, vv = k.repeat_interleave(g, dim=2), v.repeat_interleave(g, dim=2)
    s = torch.einsum("bthd,bshd->bhts", q, kk) * scale
    if mask is not None:
        s = s.masked_fill(mask.to(s.device), float("-inf"))
    mx = s.amax(dim=-1, keepdim=True)
    mx = torch.where(torch.isinf(mx), torch.zeros_like(mx), mx)
    p = torch.exp(s - mx)
    l = p.sum(-1, keepdim=True)
    p8 = p.to(torch.float8_e4m3fn).float()
    o = torch.einsum("bhts,bshd->bthd", p8, vv) / l.transpose(1, 2).clamp_min(1e-30)
    return torch.where(l.transpose(1, 2) > 0, o, torch.zeros_like(o))


def check(out, lse, exact, lse_ref, emu):
    err = (out.float() - exact).abs().max().item()
    yard = (emu - exact).abs().max().item()
    bound = 2 * yard + 2e-2 * exact.abs().max().item()
    assert err <= bound, (err, yard, bound)
    fin = torch.isfinite(lse_ref)
    assert torch.equal(torch.isfinite(lse), fin)
    assert (lse[fin] - lse_ref[fin]).abs().max().item() <= 2e-3 * max(1.0, lse_ref[fin].abs().max().item())
    assert (lse[~fin] == lse_ref[~fin]).all()


@pytest.mark.parametrize("causal,window", [(False, (-1, -1)), (True, (-1, -1)), (False, (100, 30))])
@pytest.mark.parametrize("sq,sk,h,hk,d", [(128, 128, 4, 4, 128), (256, 512, 8, 2, 128), (333, 1000, 4, 1, 64), (1024, 1024, 2, 2, 96),
                                          (200, 200, 4, 2, 256), (512, 640, 2, 1, 192), (77, 300, 2, 2, 32)])
def test_fp8_forward(fa, capi, sq, sk, h, hk, d, causal, window):
    torch.manual_seed(0)
    b = 2
    q = torch.randn(b, sq, h, d, device=DEV)
    k = torch.randn(b, sk, hk, d, device=DEV)
    v = torch.randn(b, sk, hk, d, device=DEV)
    q8, qd = quantise(q, hk)
    k8, kd = quantise(k, hk)
    v8, vd = quantise(v, hk)
    n0 = capi.fallback_count()
    out, lse = fa.flash_attn_fp8_func(q8, k8, v8, qd, kd, vd, causal=causal, window_size=window)
    assert out.dtype == torch.bfloat16 and capi.last_kernel_kind() == 2 and capi.fallback_count() == n0
    qf, kf, vf = dequant(q8, qd), dequant(k8, kd), dequant(v8, vd)
    mask = mask_for(sq, sk, causal, window)
    exact, lse_ref = orc.attention_masked_ref(qf, kf, vf, mask, lse_empty=float("-inf"))
    emu = emulated_p_e4m3(qf, kf, vf, mask, d ** -0.5)
    check(out, lse, exact, lse_ref, emu)


def test_fp8_no_descale_and_scale(fa):
    """descale pointers are optional (= 1) and softmax_scale is honoured."""
    torch.manual_seed(1)
    b, s, h, d = 1, 384, 2, 128
    q8, k8, v8 = [(torch.randn(b, s, h, d, device=DEV) * 0.5).to(torch.float8_e4m3fn) for _ in range(3)]
    out, lse = fa.flash_attn_fp8_func(q8, k8, v8, softmax_scale=0.05, causal=True)
    qf, kf, vf = q8.float(), k8.float(), v8.float()
    mask = orc.causal_mask_bottom_right(s, s)
    exact, lse_ref = orc.attention_masked_ref(qf, kf, vf, mask, softmax_scale=0.05, lse_empty=float("-inf"))
    check(out, lse, exact, lse_ref, emulated_p_e4m3(qf, kf, vf, mask, 0.05))


def test_fp8_varlen(fa):
    torch.manual_seed(2)
    lens_q, lens_k, h, hk, d = [100, 257, 31], [180, 257, 400], 4, 2, 128
    cu_q = torch.tensor([0] + list(torch.tensor(lens_q).cumsum(0)), device=DEV, dtype=torch.int32)
    cu_k = torch.tensor([0] + list(torch.tensor(lens_k).cumsum(0)), device=DEV, dtype=torch.int32)
    q8 = torch.randn(sum(lens_q), h, d, device=DEV).to(torch.float8_e4m3fn)
    k8 = torch.randn(sum(lens_k), hk, d, device=DEV).to(torch.float8_e4m3fn)
    v8 = torch.randn(sum(lens_k), hk, d, device=DEV).to(torch.float8_e4m3fn)
    sd = torch.tensor([[0.5, 2.0]] * 3, device=DEV)          # (b, hk)
    out, lse = fa.flash_attn_fp8_func(q8, k8, v8, sd, None, sd, causal=True, cu_seqlens_q=cu_q, cu_seqlens_k=cu_k,
                                      max_seqlen_q=max(lens_q), max_seqlen_k=max(lens_k))
    for i in range(3):
        a, e, ka, ke = int(cu_q[i]), int(cu_q[i + 1]), int(cu_k[i]), int(cu_k[i + 1])
        qf = dequant(q8[a:e].unsqueeze(0), sd[i:i + 1])
        kf = k8[ka:ke].unsqueeze(0).float()
        vf = dequant(v8[ka:ke].unsqueeze(0), sd[i:i + 1])
        mask = orc.causal_mask_bottom_right(e - a, ke - ka)
        exact, lse_ref = orc.attention_masked_ref(qf, kf, vf, mask, lse_empty=float("-inf"))
        check(out[a:e].unsqueeze(0), lse[i:i + 1, :, :e - a], exact, lse_ref, emulated_p_e4m3(qf, kf, vf, mask, d ** -0.5))


def test_fp8_through_the_handle(fa, capi):
    """The reference-facing surface: is_e4m3 + *_descale_ptr on the FlashMask-v2 handle, flashmaskv2_run_mha_fwd."""
    lib = capi.load()
    torch.manual_seed(3)
    b, s, h, hk, d = 2, 512, 4, 2, 128
    q = torch.randn(b, s, h, d, device=DEV)
    k = torch.randn(b, s, hk, d, device=DEV)
    v = torch.randn(b, s, hk, d, device=DEV)
    q8, qd = quantise(q, hk)
    k8, kd = quantise(k, hk)
    v8, vd = quantise(v, hk)
    out = torch.empty(b, s, h, d, device=DEV, dtype=torch.bfloat16)
    lse = torch.empty(b, h, s, device=DEV, dtype=torch.float32)
    lib.flashmaskv2_create_fwd_params_handle.restype = C.c_void_p
    hd = C.c_void_p(lib.flashmaskv2_create_fwd_params_handle())

    def put(name, val, ctype):
        fn = getattr(lib, f"flashmaskv2_fwd_params_set_{name}")
        fn.argtypes = [C.c_void_p, ctype]
        fn.restype = None
        fn(hd, val)

    for name, t in (("q", q8), ("k", k8), ("v", v8), ("o", out)):
        put(f"{name}_ptr", t.data_ptr(), C.c_void_p)
        put(f"{name}_batch_stride", t.stride(0), C.c_int64)
        put(f"{name}_row_stride", t.stride(1), C.c_int64)
        put(f"{name}_head_stride", t.stride(2), C.c_int64)
    for name, t in (("q", qd), ("k", kd), ("v", vd)):
        put(f"{name}_descale_ptr", t.data_ptr(), C.c_void_p)
        put(f"{name}_descale_batch_stride", t.stride(0), C.c_int64)
        put(f"{name}_descale_head_stride", t.stride(1), C.c_int64)
    put("softmax_lse_ptr", lse.data_ptr(), C.c_void_p)
    for name, val in (("b", b), ("seqlen_q", s), ("seqlen_k", s), ("h", h), ("h_k", hk), ("d", d), ("dv", d)):
        put(name, val, C.c_int)
    put("scale_softmax", d ** -0.5, C.c_float)
    put("is_e4m3", True, C.c_bool)
    put("is_causal", True, C.c_bool)
    lib.flashmaskv2_run_mha_fwd.argtypes = [C.c_void_p, C.c_void_p]
    lib.flashmaskv2_run_mha_fwd.restype = None
    lib.flashmaskv2_run_mha_fwd(hd, C.c_void_p(torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    lib.flashmaskv2_destroy_fwd_params_handle.argtypes = [C.c_void_p]
    lib.flashmaskv2_destroy_fwd_params_handle(hd)
    qf, kf, vf = dequant(q8, qd), dequant(k8, kd), dequant(v8, vd)
    mask = orc.causal_mask_bottom_right(s, s)
    exact, lse_ref = orc.attention_masked_ref(qf, kf, vf, mask, lse_empty=float("-inf"))
    check(out, lse, exact, lse_ref, emulated_p_e4m3(qf, kf, vf, mask, d ** -0.5))


def test_fp8_rejects_what_it_cannot_do(fa):
    q8 = torch.zeros(1, 128, 2, 72, device=DEV).to(torch.float8_e4m3fn)
    with pytest.raises(RuntimeError):
        fa.flash_attn_fp8_func(q8, q8, q8)                  # head dim not a multiple of 16
    with pytest.raises(RuntimeError):
        fa.flash_attn_fp8_func(q8.to(torch.bfloat16), q8, q8)
