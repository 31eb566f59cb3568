"""GPU parity at BASELINE.json's FULL sizes (configs 2-5b), where an oracle pass over the whole problem is too large:

* sampled slices: attention is independent per (batch element, kv-head group), so a few (b, kv head) slices of the
  full-size result are compared with the fp32 oracle under the reference's 2x rule (out, LSE, dQ, dK, dV);
* size-independent identities over the WHOLE result: rows of P sum to one, hence sum_j dV[j] = sum_i dO[i] per
  (b, kv head); the scores only see q.k, hence <q, dQ> = <k, dK> per (b, kv head);
* encodings of the same problem agree (varlen cu_seqlens vs one-row document FlashMask; split-KV vs one split).
"""
import pytest
import torch

from oracle import attention_oracle as orc
from tests._util import check_2x, check_lse, oracle_all

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _identities(q, k, dq, dk, dv, dout, g):
    """q, dq, dout: (b, s, h, d); k, dk, dv: (b, s, hk, d).  Returns nothing, asserts the two whole-tensor identities."""
    b, sq, h, d = q.shape
    hk = k.shape[2]
    # sum over keys of dV == sum over queries (and the heads of the group) of dO
    lhs = dv.double().sum(1)                                             # (b, hk, d)
    rhs = dout.double().view(b, sq, hk, g, d).sum((1, 3))
    scale = dout.double().abs().view(b, sq, hk, g, d).sum((1, 3))
    assert ((lhs - rhs).abs() <= 1e-3 * scale + 1e-2).all(), f"sum dV != sum dO (max diff {(lhs - rhs).abs().max():.3e})"
    # <q, dQ> == <k, dK> per (b, kv head)
    qdq = (q.double() * dq.double()).view(b, sq, hk, g, d).sum((1, 3, 4))
    kdk = (k.double() * dk.double()).sum((1, 3))
    mag = (q.double() * dq.double()).abs().view(b, sq, hk, g, d).sum((1, 3, 4))
    assert ((qdq - kdk).abs() <= 1e-3 * mag).all(), f"<q,dQ> != <k,dK> (max rel {((qdq - kdk).abs() / mag).max():.3e})"


def _check_slice(tag, q, k, v, dout, out, lse, dq, dk, dv, bi, hki, g, mask):
    hs = slice(hki * g, hki * g + g)
    ref = oracle_all(q[bi:bi + 1, :, hs], k[bi:bi + 1, :, hki:hki + 1], v[bi:bi + 1, :, hki:hki + 1],
                     dout[bi:bi + 1, :, hs], mask)
    check_2x(f"{tag} out", out[bi:bi + 1, :, hs], *ref["out"])
    check_lse(lse[bi:bi + 1, hs], ref["lse"])
    check_2x(f"{tag} dq", dq[bi:bi + 1, :, hs], *ref["dq"])
    check_2x(f"{tag} dk", dk[bi:bi + 1, :, hki:hki + 1], *ref["dk"], atol=2e-3)
    check_2x(f"{tag} dv", dv[bi:bi + 1, :, hki:hki + 1], *ref["dv"], atol=2e-3)


def _run_dense(fa, b, s, h, hk, d, causal):
    torch.manual_seed(0)
    q = torch.randn(b, s, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(b, s, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(b, s, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    dout = torch.randn(b, s, h, d, device=DEV, dtype=torch.bfloat16)
    out, lse, _ = fa.flash_attn_func(q, k, v, 0.0, None, causal, True)
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), dout)
    return q.detach(), k.detach(), v.detach(), dout, out.detach(), lse, dq, dk, dv


def test_c3_llama8k_causal(fa, capi):
    """BASELINE config 3 (the bench workload): b8 s8192 GQA 32q/8kv d128 causal bf16 fwd+bwd."""
    b, s, h, hk, d = 8, 8192, 32, 8, 128
    q, k, v, dout, out, lse, dq, dk, dv = _run_dense(fa, b, s, h, hk, d, True)
    assert capi.last_kernel_kind() == 2
    _identities(q, k, dq, dk, dv, dout, h // hk)
    mask = orc.causal_mask_top_left(s, s).to(DEV)
    for bi, hki in ((0, 0), (7, 5)):
        _check_slice(f"c3[{bi},{hki}]", q, k, v, dout, out, lse, dq, dk, dv, bi, hki, h // hk, mask)
        torch.cuda.empty_cache()


def test_c2_2k_noncausal(fa, capi):
    """BASELINE config 2: b8 s2048 h16 d128 non-causal bf16 fwd+bwd."""
    b, s, h, d = 8, 2048, 16, 128
    q, k, v, dout, out, lse, dq, dk, dv = _run_dense(fa, b, s, h, h, d, False)
    assert capi.last_kernel_kind() == 2
    _identities(q, k, dq, dk, dv, dout, 1)
    for bi, hki in ((0, 3), (5, 15)):
        _check_slice(f"c2[{bi},{hki}]", q, k, v, dout, out, lse, dq, dk, dv, bi, hki, 1, None)


def test_c4_varlen_and_flashmask(fa):
    """BASELINE config 4: 64 packed sequences (128..16k, 262144 tokens), GQA 32q/8kv d128 causal, in both encodings."""
    import bench
    lens = bench.c4_lengths()
    total, h, hk, d = sum(lens), 32, 8, 128
    assert total == 262144 and len(lens) == 64
    torch.manual_seed(0)
    q = torch.randn(total, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(total, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(total, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    dout = torch.randn(total, h, d, device=DEV, dtype=torch.bfloat16)
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), device=DEV, dtype=torch.int32)
    out, lse, _ = fa.flash_attn_varlen_func(q, k, v, cu, cu, max(lens), max(lens), 0.0, None, True, True)
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), dout)
    # encoding (ii): ONE 262144-token row, causal document FlashMask (lt_start[j] = end of j's document)
    ends = torch.repeat_interleave(cu[1:], torch.tensor(lens, device=DEV)).view(1, 1, total).int().contiguous()
    q4, k4, v4 = (t.detach().view(1, total, -1, d).requires_grad_(True) for t in (q, k, v))
    out4 = fa.flash_attn_func(q4, k4, v4, 0.0, None, True, flashmask=(ends, None, None, None))
    dq4, dk4, dv4 = torch.autograd.grad(out4, (q4, k4, v4), dout.view(1, total, h, d))
    for name, a, b_ in (("out", out, out4), ("dq", dq, dq4), ("dk", dk, dk4), ("dv", dv, dv4)):
        diff = (a.float() - b_.view_as(a).float()).abs().max().item()
        assert diff <= 2e-2 * max(1.0, a.float().abs().max().item()), f"encodings disagree on {name}: {diff:.3e}"
    _identities(q.detach()[None], k.detach()[None], dq[None], dk[None], dv[None], dout[None], h // hk)
    order = sorted(range(64), key=lambda i: lens[i])
    for i in (order[0], order[32], order[-1]):                         # shortest, median, longest sequence
        n, r = lens[i], slice(int(cu[i]), int(cu[i + 1]))
        mask = orc.causal_mask_top_left(n, n).to(DEV)
        hki = i % hk
        hs = slice(hki * 4, hki * 4 + 4)
        ref = oracle_all(q.detach()[r][None, :, hs], k.detach()[r][None, :, hki:hki + 1], v.detach()[r][None, :, hki:hki + 1],
                         dout[r][None, :, hs], mask)
        for tag, o_, dq_, dk_, dv_ in (("varlen", out, dq, dk, dv), ("flashmask", out4[0], dq4[0], dk4[0], dv4[0])):
            check_2x(f"c4 {tag} out[{i}]", o_[r][None, :, hs], *ref["out"])
            check_2x(f"c4 {tag} dq[{i}]", dq_[r][None, :, hs], *ref["dq"])
            check_2x(f"c4 {tag} dk[{i}]", dk_[r][None, :, hki:hki + 1], *ref["dk"], atol=2e-3)
            check_2x(f"c4 {tag} dv[{i}]", dv_[r][None, :, hki:hki + 1], *ref["dv"], atol=2e-3)
        check_lse(lse[i, hs, :n][None], ref["lse"])
        del ref, mask
        torch.cuda.empty_cache()


def test_c5b_decode_32k(fa, capi):
    """BASELINE config 5b: split-KV decode, batch 64, 32k-token cache, GQA 32q/8kv d128, one query token."""
    b, sk, h, hk, d = 64, 32768, 32, 8, 128
    torch.manual_seed(0)
    q = torch.randn(b, 1, h, d, device=DEV, dtype=torch.bfloat16)
    kc = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    vc = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    gen = torch.Generator().manual_seed(1)
    seqlens = torch.randint(sk // 2, sk + 1, (b,), generator=gen, dtype=torch.int32)
    seqlens[0], seqlens[1] = sk, 1
    seqlens = seqlens.to(DEV)
    out, lse = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=seqlens, causal=True, return_softmax_lse=True)
    assert capi.last_kernel_kind() in (1, 3)                           # decode kernel (+ combine when it split)
    one, lse1 = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=seqlens, causal=True, num_splits=1,
                                           return_softmax_lse=True)
    assert capi.last_kernel_kind() == 3                                # the TMA decode kernel itself
    assert (out.float() - one.float()).abs().max().item() <= 2e-3
    torch.testing.assert_close(lse, lse1, atol=2e-3, rtol=1e-3)
    pad = torch.arange(sk, device=DEV)[None, :] >= seqlens[:, None]    # (b, sk) True = beyond the cache length
    for lo in range(0, b, 16):                                         # the whole batch, 16 elements at a time
        sl = slice(lo, lo + 16)
        mask = pad[sl, None, None, :]
        ref, lse_ref = orc.attention_masked_ref(q[sl], kc[sl], vc[sl], mask, lse_empty=float("-inf"))
        pt, _ = orc.attention_masked_ref(q[sl], kc[sl], vc[sl], mask, upcast=False, reorder_ops=True)
        check_2x(f"c5b out[{lo}:{lo + 16}]", out[sl], ref, pt)
        check_lse(lse[sl], lse_ref)
