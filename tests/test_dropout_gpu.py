"""GPU parity for dropout and return_attn_probs, the way the reference's tests do it (tests/test_flash_attn.py:
376-427): take the dropout mask the kernel actually used from S_dmask, give the SAME mask to the fp32 oracle
``attention_ref``, compare out / dq / dk / dv with the 2x criterion, check the normalised probabilities and that the
dropped fraction is within 0.01 of dropout_p."""
import math

import pytest
import torch

from oracle import attention_oracle as orc
from tests._util import check_2x

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _ref_with_mask(q, k, v, g, keep, causal, p, upcast=True, reorder_ops=False):
    q_, k_, v_ = [t.detach().clone().requires_grad_(True) for t in (q, k, v)]
    out, attn = orc.attention_ref(q_, k_, v_, None, None, p, keep, causal=causal, upcast=upcast, reorder_ops=reorder_ops)
    dq, dk, dv = torch.autograd.grad(out, (q_, k_, v_), g.to(out.dtype))
    return out.detach(), attn.detach(), dq, dk, dv


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "fp16"])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("h,hk", [(4, 4), (6, 2)])
@pytest.mark.parametrize("sq,sk,d", [(128, 128, 64), (113, 203, 128), (256, 512, 32)])
def test_dropout_output_and_grads(fa, sq, sk, d, h, hk, causal, dtype):
    torch.manual_seed(0)
    b, p = 3, 0.17
    q = torch.randn(b, sq, h, d, device=DEV, dtype=dtype, requires_grad=True)
    k = torch.randn(b, sk, hk, d, device=DEV, dtype=dtype, requires_grad=True)
    v = torch.randn(b, sk, hk, d, device=DEV, dtype=dtype, requires_grad=True)
    g = torch.randn(b, sq, h, d, device=DEV, dtype=dtype)
    out, lse, S = fa.flash_attn_func(q, k, v, p, None, causal, True)
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), g)
    assert S.shape == (b, h, (sq + 127) // 128 * 128, (sk + 127) // 128 * 128)
    S = S[:, :, :sq, :sk].float()
    keep = S >= 0                                   # dropped weights are stored negated
    ref = _ref_with_mask(q, k, v, g, keep, causal, p)
    pt = _ref_with_mask(q, k, v, g, keep, causal, p, upcast=False, reorder_ops=True)
    check_2x("out", out, ref[0], pt[0])
    check_2x("dq", dq, ref[2], pt[2])
    check_2x("dk", dk, ref[3], pt[3])
    check_2x("dv", dv, ref[4], pt[4])
    # the probabilities themselves (reference: attn vs attn_ref, :420-421)
    attn = S.abs()
    assert (attn - ref[1].float()).abs().max().item() <= 2 * (pt[1].float() - ref[1].float()).abs().max().item() + 1e-3
    # dropped fraction over the visible entries (reference: :422-424, tolerance 0.01)
    vis = torch.ones(sq, sk, dtype=torch.bool, device=DEV) if not causal else ~orc.causal_mask_top_left(sq, sk).to(DEV)
    frac = (~keep)[:, :, vis].float().mean().item()
    assert abs(frac - p) < 0.01, frac
    # LSE is independent of dropout
    _, lse0, _ = fa.flash_attn_func(q, k, v, 0.0, None, causal, True)
    torch.testing.assert_close(lse, lse0, atol=2e-3, rtol=1e-3)


def test_dropout_is_reproducible_and_seeded(fa):
    b, s, h, d = 2, 256, 4, 64
    q, k, v = [torch.randn(b, s, h, d, device=DEV, dtype=torch.bfloat16) for _ in range(3)]
    torch.manual_seed(7)
    a1 = fa.flash_attn_func(q, k, v, 0.3)
    torch.manual_seed(7)
    a2 = fa.flash_attn_func(q, k, v, 0.3)
    a3 = fa.flash_attn_func(q, k, v, 0.3)
    assert torch.equal(a1, a2) and not torch.equal(a1, a3)
    with pytest.raises(ValueError):
        fa.flash_attn_func(q, k, v, 1.0)


def test_dropout_varlen_and_packed(fa):
    torch.manual_seed(3)
    lens, h, d, p = [77, 200, 130], 4, 64, 0.25
    total = sum(lens)
    qkv = torch.randn(total, 3, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    cu = torch.tensor([0, 77, 277, 407], device=DEV, dtype=torch.int32)
    out, lse, S = fa.flash_attn_varlen_qkvpacked_func(qkv, cu, max(lens), p, None, True, True)
    g = torch.randn_like(out)
    (dqkv,) = torch.autograd.grad(out, qkv, g)
    for i, n in enumerate(lens):
        a, e = int(cu[i]), int(cu[i + 1])
        keep = S[i:i + 1, :, :n, :n] >= 0
        q, k, v = [qkv[a:e, j].unsqueeze(0) for j in range(3)]
        ref = _ref_with_mask(q, k, v, g[a:e].unsqueeze(0), keep, True, p)
        pt = _ref_with_mask(q, k, v, g[a:e].unsqueeze(0), keep, True, p, upcast=False, reorder_ops=True)
        check_2x("out", out[a:e].unsqueeze(0), ref[0], pt[0])
        for j, name in enumerate(("dq", "dk", "dv")):
            check_2x(name, dqkv[a:e, j].unsqueeze(0), ref[2 + j], pt[2 + j])


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "fp16"])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("sq,sk,d,h,hk", [(256, 256, 128, 4, 4), (300, 515, 128, 6, 2), (128, 384, 64, 4, 1), (1000, 1000, 96, 2, 2),
                                          (256, 384, 256, 4, 2), (515, 300, 192, 2, 2), (333, 1000, 160, 4, 1)])
def test_dropout_tcgen05_path(fa, capi, sq, sk, d, h, hk, causal, dtype):
    """Dropout without return_attn_probs takes the tcgen05 kernels (kDrop instantiations).  The mask is a pure function
    of (seed, offset, b, h, row, col), so the same seed through the generic kernels (which also emit S_dmask) gives the
    same mask: the fast path is checked against the fp32 oracle fed with THAT mask, forward and backward."""
    b, p = 2, 0.17
    torch.manual_seed(1)
    q = torch.randn(b, sq, h, d, device=DEV, dtype=dtype, requires_grad=True)
    k = torch.randn(b, sk, hk, d, device=DEV, dtype=dtype, requires_grad=True)
    v = torch.randn(b, sk, hk, d, device=DEV, dtype=dtype, requires_grad=True)
    g = torch.randn(b, sq, h, d, device=DEV, dtype=dtype)
    torch.manual_seed(7)
    out = fa.flash_attn_func(q, k, v, p, None, causal)
    assert capi.last_kernel_kind() == 2, "dropout forward did not take the tcgen05 kernel"
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), g)
    kind_bwd = capi.last_kernel_kind()
    torch.manual_seed(7)
    out_g, _, S = fa.flash_attn_func(q, k, v, p, None, causal, True)      # generic kernels, same seed
    keep = S[:, :, :sq, :sk].float() >= 0
    ref = _ref_with_mask(q, k, v, g, keep, causal, p)
    pt = _ref_with_mask(q, k, v, g, keep, causal, p, upcast=False, reorder_ops=True)
    check_2x("out", out, ref[0], pt[0])
    check_2x("dq", dq, ref[2], pt[2])
    check_2x("dk", dk, ref[3], pt[3])
    check_2x("dv", dv, ref[4], pt[4])
    torch.testing.assert_close(out.float(), out_g.float(), atol=2e-2, rtol=2e-2)
    assert kind_bwd == 2, capi.last_fallback_reason()


def test_dropout_tcgen05_varlen_and_flashmask(fa, capi):
    """The kDrop kernels on a packed batch and under a FlashMask: same seed through the generic kernels recovers the
    mask (S_dmask), the oracle with that mask judges the fast path."""
    torch.manual_seed(3)
    lens, h, d, p = [77, 200, 130], 4, 64, 0.25
    total = sum(lens)
    qkv = torch.randn(total, 3, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    cu = torch.tensor([0, 77, 277, 407], device=DEV, dtype=torch.int32)
    torch.manual_seed(11)
    out = fa.flash_attn_varlen_qkvpacked_func(qkv, cu, max(lens), p, None, True)
    assert capi.last_kernel_kind() == 2
    g = torch.randn_like(out)
    (dqkv,) = torch.autograd.grad(out, qkv, g)
    assert capi.last_kernel_kind() == 2
    torch.manual_seed(11)
    _, _, S = fa.flash_attn_varlen_qkvpacked_func(qkv, cu, max(lens), p, None, True, True)
    for i, n in enumerate(lens):
        a, e = int(cu[i]), int(cu[i + 1])
        keep = S[i:i + 1, :, :n, :n] >= 0
        q, k, v = [qkv[a:e, j].unsqueeze(0) for j in range(3)]
        ref = _ref_with_mask(q, k, v, g[a:e].unsqueeze(0), keep, True, p)
        pt = _ref_with_mask(q, k, v, g[a:e].unsqueeze(0), keep, True, p, upcast=False, reorder_ops=True)
        check_2x("out", out[a:e].unsqueeze(0), ref[0], pt[0])
        for j, name in enumerate(("dq", "dk", "dv")):
            check_2x(name, dqkv[a:e, j].unsqueeze(0), ref[2 + j], pt[2 + j])
    # FlashMask (document mask) + dropout: fast path against the generic kernels with the same seed
    s_, b = 640, 2
    q = torch.randn(b, s_, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(b, s_, 2, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(b, s_, 2, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    g = torch.randn(b, s_, h, d, device=DEV, dtype=torch.bfloat16)
    lts = orc.document_lt_start([200, 300, 140]).view(1, 1, s_).expand(b, 1, s_).contiguous().to(DEV)
    res = []
    for generic in (False, True):
        torch.manual_seed(5)
        o = fa.flash_attn_func(q, k, v, p, None, True, flashmask=(lts, None, None, None), _force_generic=generic)
        res.append((o,) + torch.autograd.grad(o, (q, k, v), g))
    for name, a_, b_ in zip(("out", "dq", "dk", "dv"), res[0], res[1]):
        torch.testing.assert_close(a_.float(), b_.float(), atol=3e-2, rtol=3e-2, msg=lambda m: f"{name}: {m}")


def _probs_ref(q, k, causal, fm_dense=None):
    """fp32 softmax(scale * q k^T) over (b, h, sq, sk) with the top-left causal rule; fully masked rows -> 0."""
    b, sq, h, d = q.shape
    sk, hk = k.shape[1], k.shape[2]
    kk = k.float().repeat_interleave(h // hk, dim=2)
    s = torch.einsum("bqhd,bkhd->bhqk", q.float(), kk) * d ** -0.5
    if causal:
        s = s.masked_fill(orc.causal_mask_top_left(sq, sk).to(q.device), float("-inf"))
    if fm_dense is not None:
        s = s.masked_fill(fm_dense, float("-inf"))
    return torch.nan_to_num(torch.softmax(s, dim=-1), nan=0.0)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "fp16"])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("sq,sk,d,h,hk", [(128, 128, 64, 4, 4), (113, 203, 128, 6, 2), (384, 1000, 96, 4, 1), (515, 300, 40, 2, 2),
                                          (256, 256, 192, 4, 2), (200, 333, 256, 2, 1)])
def test_s_dmask_tcgen05_pass(fa, capi, sq, sk, d, h, hk, causal, dtype):
    """return_softmax through the flat C ABI, with and without dropout: the forward stays on the tcgen05 kernels and the
    S_dmask pass (sm100_probs.cu) gives the normalised probabilities, zeros at masked / padded positions, negative where
    the weight was dropped -- and the dropped set is exactly the one the generic kernels derive from the same seed."""
    from tests._flat_capi import flat_fwd
    lib = capi.load()
    torch.manual_seed(5)
    b = 2
    q = torch.randn(b, sq, h, d, device=DEV, dtype=dtype)
    k = torch.randn(b, sk, hk, d, device=DEV, dtype=dtype)
    v = torch.randn(b, sk, hk, d, device=DEV, dtype=dtype)
    ref = _probs_ref(q, k, causal)
    tol = 2e-2 if dtype == torch.bfloat16 else 4e-3
    n0 = capi.fallback_count()
    out0, lse0, S0 = flat_fwd(lib, q, k, v, causal, return_softmax=True)
    assert capi.last_kernel_kind() == 2 and capi.fallback_count() == n0, capi.last_fallback_reason()
    assert not torch.isnan(S0[:, :, :sq]).any()                      # every column of every real row is written
    assert (S0[:, :, :sq, sk:] == 0).all()
    assert (S0[:, :, :sq, :sk].float() - ref).abs().max().item() <= tol
    out1, lse1 = flat_fwd(lib, q, k, v, causal)
    assert torch.equal(out0, out1) and torch.equal(lse0, lse1)
    p = 0.17
    out_d, lse_d, S = flat_fwd(lib, q, k, v, causal, return_softmax=True, p_dropout=p, seed=1234, offset=8)
    assert capi.last_kernel_kind() == 2 and capi.fallback_count() == n0, capi.last_fallback_reason()
    S = S[:, :, :sq, :sk].float()
    assert (S.abs() - ref).abs().max().item() <= tol
    nz = ref > 1e-4
    frac = ((S < 0) & nz).float().sum().item() / nz.float().sum().item()
    assert abs(frac - p) < 0.015, frac
    # the generic kernels with the same seed (Python surface, seed drawn from torch's generator): same dropped set,
    # same values up to rounding, same output
    torch.manual_seed(7)
    out_f, _, S_f = fa.flash_attn_func(q, k, v, p, None, causal, True)
    assert capi.last_kernel_kind() == 2 and capi.fallback_count() == n0, capi.last_fallback_reason()
    torch.manual_seed(7)
    out_g, _, S_g = fa.flash_attn_func(q, k, v, p, None, causal, True, _force_generic=True)
    assert capi.last_kernel_kind() == 1
    S_f, S_g = S_f[:, :, :sq, :sk].float(), S_g[:, :, :sq, :sk].float()
    assert torch.equal((S_f < 0) & nz, (S_g < 0) & nz)
    assert (S_f - S_g).abs().max().item() <= tol
    torch.testing.assert_close(out_f.float(), out_g.float(), atol=2e-2, rtol=2e-2)


def test_s_dmask_tcgen05_varlen_flashmask(fa, capi):
    """The S_dmask pass on a packed batch (per-sequence bounds) and under FlashMask vectors (element-wise rule)."""
    torch.manual_seed(3)
    lens, h, d, p = [77, 200, 130], 4, 64, 0.25
    total = sum(lens)
    qkv = torch.randn(total, 3, h, d, device=DEV, dtype=torch.bfloat16)
    cu = torch.tensor([0, 77, 277, 407], device=DEV, dtype=torch.int32)
    n0 = capi.fallback_count()
    torch.manual_seed(11)
    out, lse, S = fa.flash_attn_varlen_qkvpacked_func(qkv, cu, max(lens), p, None, True, True)
    assert capi.last_kernel_kind() == 2 and capi.fallback_count() == n0, capi.last_fallback_reason()
    torch.manual_seed(11)
    out_g, _, S_g = fa.flash_attn_varlen_qkvpacked_func(qkv, cu, max(lens), p, None, True, True, _force_generic=True)
    for i, n in enumerate(lens):
        a, e = int(cu[i]), int(cu[i + 1])
        ref = _probs_ref(qkv[a:e, 0].unsqueeze(0), qkv[a:e, 1].unsqueeze(0), True)
        Si, Sgi = S[i:i + 1, :, :n, :n].float(), S_g[i:i + 1, :, :n, :n].float()
        assert (Si.abs() - ref).abs().max().item() <= 2e-2
        assert torch.equal((Si < 0) & (ref > 1e-4), (Sgi < 0) & (ref > 1e-4))
        assert (S[i, :, :n, n:] == 0).all()
    torch.testing.assert_close(out.float(), out_g.float(), atol=2e-2, rtol=2e-2)
    # FlashMask (document mask: LT start = end of the column's document), dense
    b, s = 2, 384
    q, k, v = [torch.randn(b, s, h, d, device=DEV, dtype=torch.bfloat16) for _ in range(3)]
    doc_end = torch.tensor([100, 100 + 150, s], device=DEV)
    col = torch.arange(s, device=DEV)
    lts = doc_end[torch.bucketize(col, doc_end, right=True)].to(torch.int32).view(1, 1, s).expand(b, 1, s).contiguous()
    torch.manual_seed(13)
    out, lse, S = fa.flash_attn_func(q, k, v, p, None, True, True, flashmask=(lts, None, None, None))
    assert capi.last_kernel_kind() == 2 and capi.fallback_count() == n0, capi.last_fallback_reason()
    row = torch.arange(s, device=DEV).view(s, 1)
    dense = (row >= lts[0, 0].view(1, s)).view(1, 1, s, s)
    ref = _probs_ref(q, k, True, dense)
    Sf = S[:, :, :s, :s].float()
    assert (Sf.abs() - ref).abs().max().item() <= 2e-2
    assert (Sf[dense.expand_as(Sf)] == 0).all()
