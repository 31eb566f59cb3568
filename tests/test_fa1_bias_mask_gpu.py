"""GPU parity for the two legacy symbols flash_attn_fwd_with_bias_and_mask / flash_attn_bwd_with_bias_and_mask
(include/fa_capi.h; reference csrc/capi/flash_attn.h:247-311, csrc/flash_attn_with_bias_and_mask/
flash_attn_with_bias_mask.cu:268-541), called through ctypes exactly as a dlsym caller would: packed (total, h, d)
tensors + cu_seqlens, additive mask (b, 1|h, 1|sq, sk) and bias (1|b, h, sq, sk), softmax_scale 1, dbias (b, h, sq, sk).
"""
import ctypes as C

import pytest
import torch

from oracle import attention_oracle as orc
from tests._util import check_2x, check_lse

pytestmark = pytest.mark.gpu
DEV = "cuda"
vp = lambda t: None if t is None else C.c_void_p(t.data_ptr())
i32 = C.c_int


def _dims(t):
    return None if t is None else (C.c_int64 * 4)(*t.shape)


def _fwd(lib, q, k, v, out, cu_q, cu_k, b, h, d, msq, msk, scale, zero, lse, mask, bias, p_dropout=0.0):
    lib.flash_attn_fwd_with_bias_and_mask.restype = C.c_bool
    ws = C.c_uint64(123)
    return lib.flash_attn_fwd_with_bias_and_mask(
        vp(q), vp(k), vp(v), vp(out), vp(cu_q), vp(cu_k), i32(q.shape[0]), i32(k.shape[0]), i32(b), i32(h), i32(d),
        i32(msq), i32(msk), C.c_float(p_dropout), C.c_float(scale), C.c_bool(zero), C.c_bool(q.dtype == torch.bfloat16),
        i32(0), vp(lse), None, C.byref(ws), None, C.c_uint64(1), C.c_uint64(2), vp(mask), vp(bias), _dims(mask),
        _dims(bias))


def _bwd(lib, q, k, v, dq, dk, dv, out, dout, cu_q, cu_k, b, h, d, msq, msk, scale, zero, lse, dsm, dbias, mask, bias):
    lib.flash_attn_bwd_with_bias_and_mask.restype = C.c_bool
    ws = C.c_uint64(123)
    return lib.flash_attn_bwd_with_bias_and_mask(
        vp(q), vp(k), vp(v), vp(dq), vp(dk), vp(dv), vp(out), vp(dout), vp(cu_q), vp(cu_k), i32(q.shape[0]),
        i32(k.shape[0]), i32(b), i32(h), i32(d), i32(msq), i32(msk), C.c_float(0.0), C.c_float(scale), C.c_bool(zero),
        C.c_bool(q.dtype == torch.bfloat16), i32(0), vp(lse), vp(dsm), vp(dbias), None, C.byref(ws), None,
        C.c_uint64(1), C.c_uint64(2), vp(mask), vp(bias), _dims(mask), _dims(bias))


def _oracle_seq(q, k, v, g, add, scale, upcast):
    """One sequence: (out, lse, dq, dk, dv, dadd) with the additive term `add` (1, h, sq, sk) or None."""
    q_, k_, v_ = (t.detach().clone().unsqueeze(0).requires_grad_(True) for t in (q, k, v))
    a_ = None
    if add is not None:
        a_ = (add.float() if upcast else add).detach().clone().requires_grad_(True)
    out, lse = orc.attention_masked_ref(q_, k_, v_, None, scale, upcast, not upcast, lse_empty=float("-inf"), bias=a_)
    ins = (q_, k_, v_) + ((a_,) if a_ is not None else ())
    grads = torch.autograd.grad(out, ins, g.unsqueeze(0).to(out.dtype))
    return (out[0].detach(), lse[0], grads[0][0], grads[1][0], grads[2][0], grads[3] if a_ is not None else None)


def _run_case(capi, lens_q, lens_k, h, d, dtype, mask_shape, bias_batch, seed=0):
    lib = capi.load()
    torch.manual_seed(seed)
    b = len(lens_q)
    msq, msk = max(lens_q), max(lens_k)
    tq, tk = sum(lens_q), sum(lens_k)
    q = (torch.randn(tq, h, d, device=DEV) * d ** -0.25).to(dtype)   # scale is 1 on this path: keep q.k at unit variance
    k = (torch.randn(tk, h, d, device=DEV) * d ** -0.25).to(dtype)
    v = torch.randn(tk, h, d, device=DEV, dtype=dtype)
    g = torch.randn(tq, h, d, device=DEV, dtype=dtype)
    cu_q = torch.tensor([0] + list(torch.tensor(lens_q).cumsum(0)), device=DEV, dtype=torch.int32)
    cu_k = torch.tensor([0] + list(torch.tensor(lens_k).cumsum(0)), device=DEV, dtype=torch.int32)
    mask = bias = None
    if mask_shape is not None:
        mh, ms = mask_shape
        keep = torch.rand(b, mh, msq if ms else 1, msk, device=DEV) > 0.25
        keep[..., 0] = True                                            # at least one visible key per row
        mask = torch.where(keep, 0.0, -1e4).to(dtype).contiguous()
    if bias_batch:
        bias = torch.randn(bias_batch, h, msq, msk, device=DEV).to(dtype).contiguous()
    sq_r = (msq + 15) // 16 * 16
    out = torch.full((tq, h, d), float("nan"), device=DEV, dtype=dtype)
    lse = torch.full((b, h, sq_r), float("nan"), device=DEV, dtype=torch.float32)
    assert _fwd(lib, q, k, v, out, cu_q, cu_k, b, h, d, msq, msk, 1.0, True, lse, mask, bias), capi.last_error()
    dq, dk, dv = (torch.full_like(t, float("nan")) for t in (q, k, v))
    dsm = torch.empty(b, h, sq_r, device=DEV, dtype=torch.float32)
    dbias = torch.full((b, h, msq, msk), float("nan"), device=DEV, dtype=dtype) if bias is not None else None
    assert _bwd(lib, q, k, v, dq, dk, dv, out, g, cu_q, cu_k, b, h, d, msq, msk, 1.0, True, lse, dsm, dbias, mask,
                bias), capi.last_error()
    torch.cuda.synchronize()
    for i in range(b):
        sq, sk = lens_q[i], lens_k[i]
        rq, rk = slice(int(cu_q[i]), int(cu_q[i + 1])), slice(int(cu_k[i]), int(cu_k[i + 1]))
        add = None
        if mask is not None:
            add = mask[i:i + 1, :, :sq if mask.shape[2] > 1 else 1, :sk].float().expand(1, h, sq, sk)
        if bias is not None:
            bi = bias[i % bias.shape[0]][None, :, :sq, :sk].float()
            add = bi if add is None else add + bi
        add_lp = None if add is None else add.to(dtype)
        ref = _oracle_seq(q[rq], k[rk], v[rk], g[rq], add, 1.0, True)
        pt = _oracle_seq(q[rq], k[rk], v[rk], g[rq], add_lp, 1.0, False)
        check_2x(f"out[{i}]", out[rq], ref[0], pt[0])
        check_lse(lse[i, :, :sq], ref[1])
        assert torch.isinf(lse[i, :, sq:]).all() and (lse[i, :, sq:] < 0).all()      # zero_tensors fill
        check_2x(f"dq[{i}]", dq[rq], ref[2], pt[2], atol=1e-3)
        check_2x(f"dk[{i}]", dk[rk], ref[3], pt[3], atol=1e-3)
        check_2x(f"dv[{i}]", dv[rk], ref[4], pt[4], atol=1e-3)
        if bias is not None:
            check_2x(f"dbias[{i}]", dbias[i, :, :sq, :sk], ref[5][0], pt[5][0], atol=1e-3)
            assert (dbias[i, :, sq:, :] == 0).all() and (dbias[i, :, :, sk:] == 0).all()
    return out


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("d", [32, 64, 128])
@pytest.mark.parametrize("mask_shape,bias_batch", [((1, 0), 1), ((4, 1), 0), (None, 1), ((1, 1), 2)])
def test_bias_and_mask_uniform(capi, d, dtype, mask_shape, bias_batch):
    """Uniform lengths (what the padded caller passes): cu_seqlens = arange * s."""
    _run_case(capi, [100, 100], [136, 136], 4, d, dtype, mask_shape, bias_batch)


def test_bias_and_mask_ragged(capi):
    """Ragged batch: rows / columns of mask and bias are indexed within the (max_sq, max_sk) dense tensors."""
    _run_case(capi, [57, 128, 1], [200, 64, 33], 2, 64, torch.bfloat16, (2, 1), 3)


def test_plain_fa1_call_and_workspace_query(capi):
    """No mask, no bias: any softmax_scale; out == NULL asks for the workspace size (0 here)."""
    lib = capi.load()
    torch.manual_seed(1)
    b, s, h, d = 2, 77, 3, 64
    q, k, v = (torch.randn(b * s, h, d, device=DEV, dtype=torch.float16) for _ in range(3))
    cu = torch.arange(0, (b + 1) * s, s, device=DEV, dtype=torch.int32)
    lib.flash_attn_fwd_with_bias_and_mask.restype = C.c_bool
    ws = C.c_uint64(99)
    assert lib.flash_attn_fwd_with_bias_and_mask(
        vp(q), vp(k), vp(v), None, vp(cu), vp(cu), i32(b * s), i32(b * s), i32(b), i32(h), i32(d), i32(s), i32(s),
        C.c_float(0.0), C.c_float(0.125), C.c_bool(False), C.c_bool(False), i32(0), None, None, C.byref(ws), None,
        C.c_uint64(0), C.c_uint64(0), None, None, None, None)
    assert ws.value == 0
    out = torch.empty_like(q)
    lse = torch.empty(b, h, 80, device=DEV, dtype=torch.float32)
    assert _fwd(lib, q, k, v, out, cu, cu, b, h, d, s, s, 0.125, False, lse, None, None), capi.last_error()
    q4, k4, v4 = (t.view(b, s, h, d) for t in (q, k, v))
    ref, lse_ref = orc.attention_masked_ref(q4, k4, v4, None, 0.125)
    pt, _ = orc.attention_masked_ref(q4, k4, v4, None, 0.125, upcast=False, reorder_ops=True)
    check_2x("out", out.view(b, s, h, d), ref, pt)
    check_lse(lse[:, :, :s], lse_ref)


def test_bias_needs_unit_scale(capi):
    lib = capi.load()
    b, s, h, d = 1, 16, 1, 32
    q = torch.zeros(b * s, h, d, device=DEV, dtype=torch.bfloat16)
    cu = torch.tensor([0, s], device=DEV, dtype=torch.int32)
    bias = torch.zeros(1, h, s, s, device=DEV, dtype=torch.bfloat16)
    lse = torch.zeros(b, h, s, device=DEV)
    assert not _fwd(lib, q, q, q, torch.empty_like(q), cu, cu, b, h, d, s, s, 0.5, False, lse, None, bias)
    assert "softmax_scale must be 1" in capi.last_error()
