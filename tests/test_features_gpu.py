"""GPU parity tests for the rest of the boundary: varlen (cu_seqlens), FlashMask column-sparse masks, sliding
window, KV-cache decode with split-KV + LSE combine, ring-merge, and the flat FA2 C symbols called through ctypes.
The reference repository holds no tests for flashmask / split-KV / the C ABI: parity here is
against the dense-mask restatement in oracle/attention_oracle.py."""
import ctypes as C

import pytest
import torch

from oracle import attention_oracle as orc
from tests._util import check_2x, check_lse, oracle_all

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _lens_to_cu(lens):
    cu = torch.zeros(len(lens) + 1, dtype=torch.int32)
    cu[1:] = torch.tensor(lens).cumsum(0)
    return cu.to(DEV)


@pytest.mark.parametrize("generic", [False, True], ids=["tc", "generic"])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", [64, 128])
def test_varlen(fa, d, causal, generic):
    torch.manual_seed(0)
    lens_q = [76, 34, 146, 1, 300, 129]     # incl. the reference's cu_q=[0,76,110,256] prefix (:942-943)
    lens_k = [76, 34, 146, 1, 300, 129] if causal else [1, 97, 200, 3, 260, 128]
    h, hk = 6, 2
    tq, tk = sum(lens_q), sum(lens_k)
    q = torch.randn(tq, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(tk, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(tk, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    g = torch.randn(tq, h, d, device=DEV, dtype=torch.bfloat16)
    cu_q, cu_k = _lens_to_cu(lens_q), _lens_to_cu(lens_k)
    out, lse, _ = fa.flash_attn_varlen_func(q, k, v, cu_q, cu_k, max(lens_q), max(lens_k), 0.0, None, causal, True,
                                            _force_generic=generic)
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), g)
    oq = ok = 0
    for i, (lq, lk) in enumerate(zip(lens_q, lens_k)):
        sl_q, sl_k = slice(oq, oq + lq), slice(ok, ok + lk)
        mask = orc.causal_mask_top_left(lq, lk).to(DEV) if causal else None
        ref = oracle_all(q.detach()[sl_q][None], k.detach()[sl_k][None], v.detach()[sl_k][None], g[sl_q][None], mask)
        check_2x(f"out[{i}]", out[sl_q][None], *ref["out"])
        check_lse(lse[i, :, :lq][None], ref["lse"])
        check_2x(f"dq[{i}]", dq[sl_q][None], *ref["dq"])
        check_2x(f"dk[{i}]", dk[sl_k][None], *ref["dk"])
        check_2x(f"dv[{i}]", dv[sl_k][None], *ref["dv"])
        oq += lq; ok += lk


def test_varlen_bwd_overflow(fa):
    """tests/test_flash_attn.py:929-959: cu_q=[0,76,110,256], cu_k=[0,1,2,3] -> no NaN."""
    torch.manual_seed(0)
    h, d = 2, 128
    q = (torch.randn(256, h, d, device=DEV, dtype=torch.bfloat16) * 5).requires_grad_()
    k, v = [(torch.randn(3, h, d, device=DEV, dtype=torch.bfloat16) * 3).requires_grad_() for _ in range(2)]
    cu_q = torch.tensor([0, 76, 110, 256], device=DEV, dtype=torch.int32)
    cu_k = torch.tensor([0, 1, 2, 3], device=DEV, dtype=torch.int32)
    out = fa.flash_attn_varlen_func(q, k, v, cu_q, cu_k, 150, 1)
    g = torch.randn_like(out)
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), g)
    for t in (out, dq, dk, dv):
        assert not t.isnan().any()


def _rand_flashmask(kind, b, hm, sq, sk, gen):
    """Random legal vectors for each flashmask flavour."""
    j = torch.arange(sk).view(1, 1, sk)
    r = lambda lo, hi: torch.randint(lo, hi, (b, hm, sk), generator=gen)
    if kind == "doc_causal":   # lt_start[j] = end of j's document
        lens = [sk // 3, sk // 4, sk - sk // 3 - sk // 4]
        lts = orc.document_lt_start(lens).view(1, 1, sk).expand(b, hm, sk).clone()
        return (lts, None, None, None), True
    if kind == "lt_band":      # causal + masked band [lts, lte)
        lts = torch.maximum(r(0, sq + 1), j + 1)
        lte = torch.minimum(lts + r(0, sq // 2 + 1), torch.tensor(sq))
        return (lts.int(), lte.int(), None, None), True
    if kind == "bidir_doc":    # non-causal: rows >= lts masked, rows < ute masked
        lens = [sk // 2, sk - sk // 2]
        ends = orc.document_lt_start(lens).view(1, 1, sk)
        starts = torch.cat([torch.full((n,), s) for n, s in zip(lens, [0, lens[0]])]).view(1, 1, sk)
        return (ends.expand(b, hm, sk).int().clone(), None, None, starts.expand(b, hm, sk).int().clone()), False
    if kind == "four":         # two arbitrary bands
        lts = r(sq // 2, sq + 1); lte = torch.minimum(lts + r(0, sq // 3 + 1), torch.tensor(sq))
        uts = r(0, sq // 4 + 1); ute = torch.minimum(uts + r(0, sq // 4 + 1), lts)
        return (lts.int(), lte.int(), uts.int(), ute.int()), False
    raise ValueError(kind)


@pytest.mark.parametrize("generic", [False, True], ids=["tc", "generic"])
@pytest.mark.parametrize("kind", ["doc_causal", "lt_band", "bidir_doc", "four"])
@pytest.mark.parametrize("sq,sk", [(256, 256), (300, 300), (192, 448)])
def test_flashmask(fa, sq, sk, kind, generic):
    gen = torch.Generator().manual_seed(0)
    torch.manual_seed(0)
    b, h, hk, hm, d = 2, 4, 2, 2, 128
    (vecs, causal) = _rand_flashmask(kind, b, hm, sq, sk, gen)
    if causal and sq != sk:
        pytest.skip("causal flashmask flavours are defined for sq == sk")
    q = torch.randn(b, sq, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    g = torch.randn(b, sq, h, d, device=DEV, dtype=torch.bfloat16)
    dvecs = tuple(None if x is None else x.to(DEV) for x in vecs)
    out, lse, _ = fa.flash_attn_func(q, k, v, 0.0, None, causal, True, flashmask=dvecs, _force_generic=generic)
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), g)
    dense = orc.flashmask_dense(sq, sk, *vecs)                       # (b, hm, sq, sk)
    dense = dense.repeat_interleave(h // hm, dim=1)
    if causal:
        dense = dense | orc.causal_mask_top_left(sq, sk)
    ref = oracle_all(q.detach(), k.detach(), v.detach(), g, dense.to(DEV))
    check_2x("out", out, *ref["out"])
    check_lse(lse, ref["lse"])
    check_2x("dq", dq, *ref["dq"])
    check_2x("dk", dk, *ref["dk"])
    check_2x("dv", dv, *ref["dv"])


def test_flashmask_attention_api(fa):
    torch.manual_seed(0)
    b, s, h, d = 1, 384, 2, 64
    q, k, v = [torch.randn(b, s, h, d, device=DEV, dtype=torch.bfloat16) for _ in range(3)]
    lts = orc.document_lt_start([100, 184, 100]).view(1, 1, s, 1).to(DEV)
    out = fa.flashmask_attention(q, k, v, lts, causal=True)
    dense = orc.flashmask_dense(s, s, lts[..., 0].cpu()) | orc.causal_mask_top_left(s, s)
    ref, _ = orc.attention_masked_ref(q, k, v, dense.to(DEV))
    pt, _ = orc.attention_masked_ref(q, k, v, dense.to(DEV), upcast=False, reorder_ops=True)
    check_2x("out", out, ref, pt)


@pytest.mark.parametrize("generic", [False, True], ids=["tc", "generic"])
@pytest.mark.parametrize("window", [(64, 0), (100, 30), (-1, 17), (300, -1)])
@pytest.mark.parametrize("sq,sk", [(256, 256), (200, 333)])
def test_local_window(fa, sq, sk, window, generic):
    q, k, v, g = [None] * 4
    torch.manual_seed(0)
    b, h, hk, d = 2, 4, 4, 128
    q = torch.randn(b, sq, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    g = torch.randn(b, sq, h, d, device=DEV, dtype=torch.bfloat16)
    out, lse, _ = fa.flash_attn_func(q, k, v, 0.0, None, False, True, window_size=window, _force_generic=generic)
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), g)
    mask = orc.local_mask(sq, sk, *window).to(DEV)
    ref = oracle_all(q.detach(), k.detach(), v.detach(), g, mask)
    check_2x("out", out, *ref["out"])
    check_lse(lse, ref["lse"])
    check_2x("dq", dq, *ref["dq"])
    check_2x("dk", dk, *ref["dk"])
    check_2x("dv", dv, *ref["dv"])


def _big_flashmask(kind, b, hm, s, gen):
    """Masks whose masked regions cover whole 128x128 blocks, so that block skipping in the tcgen05 kernels is exercised
    (fully masked blocks dropped, partially masked blocks take the element rule, clear blocks take neither)."""
    if kind == "doc_causal":
        outs = []
        for _ in range(b * hm):
            cuts = sorted(torch.randint(1, s, (4,), generator=gen).tolist())
            lens = [cuts[0]] + [cuts[i + 1] - cuts[i] for i in range(3)] + [s - cuts[3]]
            outs.append(orc.document_lt_start([n for n in lens if n > 0]))
        return (torch.stack(outs).view(b, hm, s).int(), None, None, None), True
    if kind == "bidir_doc":
        cuts = [0, s // 5, s // 2 + 37, s]
        ends = torch.cat([torch.full((cuts[i + 1] - cuts[i],), cuts[i + 1]) for i in range(3)])
        starts = torch.cat([torch.full((cuts[i + 1] - cuts[i],), cuts[i]) for i in range(3)])
        e = ends.view(1, 1, s).expand(b, hm, s).int().clone()
        st = starts.view(1, 1, s).expand(b, hm, s).int().clone()
        return (e, None, None, st), False
    if kind == "sliding_in_doc":   # causal, rows [j + w, end of doc) visible again is NOT expressible; band = [lts, lte)
        j = torch.arange(s).view(1, 1, s)
        w = 200
        lts = (j + w).clamp(max=s).expand(b, hm, s).int().clone()          # rows >= j + w masked ...
        lte = torch.full((b, hm, s), s, dtype=torch.int32)                 # ... to the end: causal sliding window
        return (lts, lte, None, None), True
    if kind == "four_blocks":      # two wide bands per column, piecewise constant over 96-column groups
        grp = (torch.arange(s) // 96).view(1, 1, s)
        ngrp = int(grp.max()) + 1
        base = torch.randint(s // 2, s, (b, hm, ngrp), generator=gen)
        lts = torch.gather(base, 2, grp.expand(b, hm, s))
        lte = torch.minimum(lts + 300, torch.tensor(s))
        ub = torch.randint(0, s // 4, (b, hm, ngrp), generator=gen)
        uts = torch.gather(ub, 2, grp.expand(b, hm, s))
        ute = torch.minimum(uts + 280, lts)
        return (lts.int(), lte.int(), uts.int(), ute.int()), False
    raise ValueError(kind)


@pytest.mark.parametrize("kind", ["doc_causal", "bidir_doc", "sliding_in_doc", "four_blocks"])
@pytest.mark.parametrize("s,h,hk,hm", [(1536, 4, 2, 1), (1100, 4, 4, 4), (2048, 8, 2, 2)])
def test_flashmask_block_skipping(fa, capi, kind, s, h, hk, hm):
    gen = torch.Generator().manual_seed(3)
    torch.manual_seed(3)
    b, d = 2, 128
    vecs, causal = _big_flashmask(kind, b, hm, s, gen)
    q = torch.randn(b, s, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(b, s, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(b, s, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    g = torch.randn(b, s, h, d, device=DEV, dtype=torch.bfloat16)
    dvecs = tuple(None if x is None else x.to(DEV) for x in vecs)
    out, lse, _ = fa.flash_attn_func(q, k, v, 0.0, None, causal, True, flashmask=dvecs)
    assert capi.last_kernel_kind() == 2
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), g)
    dense = orc.flashmask_dense(s, s, *vecs).repeat_interleave(h // hm, dim=1)
    if causal:
        dense = dense | orc.causal_mask_top_left(s, s)
    ref = oracle_all(q.detach(), k.detach(), v.detach(), g, dense.to(DEV))
    check_2x("out", out, *ref["out"])
    check_lse(lse, ref["lse"])
    check_2x("dq", dq, *ref["dq"])
    check_2x("dk", dk, *ref["dk"])
    check_2x("dv", dv, *ref["dv"])
    # the generic SIMT kernels (no block skipping) must agree with the tcgen05 kernels
    out_g = fa.flash_attn_func(q, k, v, 0.0, None, causal, flashmask=dvecs, _force_generic=True)
    torch.testing.assert_close(out.float(), out_g.float(), atol=2e-2, rtol=2e-2)


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", [64, 128, 40])
def test_varlen_packed_variants(fa, d, causal):
    """flash_attn_varlen_qkvpacked_func / flash_attn_varlen_kvpacked_func (reference tests/test_flash_attn.py:441-514):
    inputs are strided views of one packed tensor, gradients come back as ONE packed tensor; results are the same
    bits as the unpacked call on the same data (same kernels, same strides), and follow the oracle."""
    torch.manual_seed(0)
    lens = [76, 34, 146, 200]
    h, total = 4, sum(lens)
    cu = _lens_to_cu(lens)
    qkv = torch.randn(total, 3, h, d, device=DEV, dtype=torch.float16, requires_grad=True)
    g = torch.randn(total, h, d, device=DEV, dtype=torch.float16)
    out = fa.flash_attn_varlen_qkvpacked_func(qkv, cu, max(lens), 0.0, None, causal)
    (dqkv,) = torch.autograd.grad(out, qkv, g)
    assert dqkv.shape == qkv.shape
    q, k, v = [qkv.detach()[:, i].clone().requires_grad_(True) for i in range(3)]
    out2 = fa.flash_attn_varlen_func(q, k, v, cu, cu, max(lens), max(lens), 0.0, None, causal)
    dq, dk, dv = torch.autograd.grad(out2, (q, k, v), g)
    assert torch.equal(out, out2)
    for i, gref in enumerate((dq, dk, dv)):
        torch.testing.assert_close(dqkv[:, i], gref, atol=2e-3, rtol=1e-2)   # dQ accumulation order is not fixed
    kv = torch.stack((k.detach(), v.detach()), dim=1).requires_grad_(True)
    q3 = q.detach().clone().requires_grad_(True)
    out3 = fa.flash_attn_varlen_kvpacked_func(q3, kv, cu, cu, max(lens), max(lens), 0.0, None, causal)
    dq3, dkv = torch.autograd.grad(out3, (q3, kv), g)
    assert torch.equal(out3, out2) and dkv.shape == kv.shape
    torch.testing.assert_close(dkv[:, 0], dk, atol=2e-3, rtol=1e-2)
    torch.testing.assert_close(dkv[:, 1], dv, atol=2e-3, rtol=1e-2)
    o = 0
    for n in lens:   # and the oracle, per sequence
        sl = slice(o, o + n)
        mask = orc.causal_mask_top_left(n, n).to(DEV) if causal else None
        ref = oracle_all(q.detach()[sl][None], k.detach()[sl][None], v.detach()[sl][None], g[sl][None], mask)
        check_2x("out", out[sl][None], *ref["out"])
        check_2x("dq", dqkv[sl, 0][None], *ref["dq"])
        check_2x("dk", dqkv[sl, 1][None], *ref["dk"])
        check_2x("dv", dqkv[sl, 2][None], *ref["dv"])
        o += n


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("mha_type", ["mha", "gqa", "mqa"])
def test_varlen_from_padded_batch(fa, mha_type, causal):
    """The reference's own varlen recipe (tests/test_flash_attn.py:657-783): padded batch + random padding masks ->
    ``unpad_input`` -> ``flash_attn_varlen_func`` -> ``pad_input``, compared with ``attention_ref`` on the padded batch
    with the same masks (outputs and input gradients, the 2x rule)."""
    bp = fa.bert_padding
    torch.manual_seed(0)
    b, s, h, d = 5, 239, 6, 64
    hk = {"mha": 6, "gqa": 2, "mqa": 1}[mha_type]
    q = torch.randn(b, s, h, d, device=DEV, dtype=torch.float16, requires_grad=True)
    k = torch.randn(b, s, hk, d, device=DEV, dtype=torch.float16, requires_grad=True)
    v = torch.randn(b, s, hk, d, device=DEV, dtype=torch.float16, requires_grad=True)
    g = torch.randn(b, s, h, d, device=DEV, dtype=torch.float16)
    lens = torch.tensor([239, 1, 100, 128, 57], device=DEV)
    pad = torch.arange(s, device=DEV).view(1, s) < lens.view(b, 1)     # one mask for q and k: causal stays well defined
    qu, idx_q, cu_q, max_q = bp.unpad_input(q, pad)
    ku, _, cu_k, max_k = bp.unpad_input(k, pad)
    vu, _, _, _ = bp.unpad_input(v, pad)
    out_u = fa.flash_attn_varlen_func(qu, ku, vu, cu_q, cu_k, max_q, max_k, 0.0, None, causal)
    out = bp.pad_input(out_u, idx_q, b, s)
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), g)
    qd, kd, vd = [t.detach().clone().requires_grad_(True) for t in (q, k, v)]
    ref, _ = orc.attention_ref(qd, kd, vd, pad, pad, 0.0, None, causal=causal)
    rq, rk, rv = torch.autograd.grad(ref, (qd, kd, vd), g)
    qd2, kd2, vd2 = [t.detach().clone().requires_grad_(True) for t in (q, k, v)]
    pt, _ = orc.attention_ref(qd2, kd2, vd2, pad, pad, 0.0, None, causal=causal, upcast=False, reorder_ops=True)
    pq, pk, pv = torch.autograd.grad(pt, (qd2, kd2, vd2), g)
    check_2x("out", out, ref, pt)
    check_2x("dq", dq, rq, pq)
    check_2x("dk", dk, rk, pk)
    check_2x("dv", dv, rv, pv)
    assert (out[~pad] == 0).all() and (dq[~pad] == 0).all() and (dk[~pad] == 0).all()


def test_more_than_2g_elements(fa):
    """Maximum sizes: q / out / dq with more than 2^31 elements (4.9 GB each in bf16), so every element offset in the
    kernels has to be 64-bit.  Sliding-window attention keeps the arithmetic small; the far end of the tensors (offsets
    beyond 2^31 elements and 2^32 bytes) is checked against the oracle on a slice that contains its whole window."""
    free, _ = torch.cuda.mem_get_info()
    if free < 60 * 2 ** 30:
        pytest.skip("needs ~45 GB of device memory")
    torch.manual_seed(0)
    b, s, h, hk, d, w = 1, 300032, 64, 2, 128, 128
    assert b * s * h * d > 2 ** 31
    q = torch.randn(b, s, h, d, device=DEV, dtype=torch.bfloat16).requires_grad_(True)
    k = torch.randn(b, s, hk, d, device=DEV, dtype=torch.bfloat16).requires_grad_(True)
    v = torch.randn(b, s, hk, d, device=DEV, dtype=torch.bfloat16).requires_grad_(True)
    out = fa.flash_attn_func(q, k, v, causal=True, window_size=(w, 0))
    g = torch.zeros_like(out)
    lo, hi = s - 1024, s                      # the slice that is checked: the last 1024 rows / keys
    g[:, lo:hi] = torch.randn(b, hi - lo, h, d, device=DEV, dtype=torch.bfloat16)
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), g)
    torch.cuda.synchronize()
    qs, ks, vs = [t.detach()[:, lo - w:hi].clone() for t in (q, k, v)]     # rows lo-w.. so that every window is complete
    n = hi - lo + w
    mask = orc.causal_mask_bottom_right(n, n) | orc.local_mask(n, n, w, 0)
    ref = oracle_all(qs, ks, vs, g[:, lo - w:hi], mask.to(DEV))
    sl = slice(w, n)                          # the rows / keys of the slice proper
    check_2x("out", out[:, lo:hi], ref["out"][0][:, sl], ref["out"][1][:, sl])
    check_2x("dq", dq[:, lo:hi], ref["dq"][0][:, sl], ref["dq"][1][:, sl])
    check_2x("dk", dk[:, lo:hi], ref["dk"][0][:, sl], ref["dk"][1][:, sl])
    check_2x("dv", dv[:, lo:hi], ref["dv"][0][:, sl], ref["dv"][1][:, sl])
    assert (dq[:, : lo - w] == 0).all() and (dk[:, : lo - 2 * w] == 0).all()   # nothing leaks to the untouched front


@pytest.mark.parametrize("generic", [False, True], ids=["tc", "generic"])
def test_varlen_empty_sequences(fa, generic):
    """Empty inputs: sequences of length 0 inside a packed batch (cu_seqlens with repeated offsets), on the query side,
    the key side or both.  A query with no key gives out = 0 and no gradient; a key with no query gets dk = dv = 0."""
    torch.manual_seed(0)
    lens_q = [5, 0, 130, 0, 64]
    lens_k = [7, 3, 0, 0, 64]
    h, hk, d = 4, 2, 64
    tq, tk = sum(lens_q), sum(lens_k)
    q = torch.randn(tq, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(tk, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(tk, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    g = torch.randn(tq, h, d, device=DEV, dtype=torch.bfloat16)
    cu_q, cu_k = _lens_to_cu(lens_q), _lens_to_cu(lens_k)
    out = fa.flash_attn_varlen_func(q, k, v, cu_q, cu_k, max(lens_q), max(lens_k), 0.0, None, False, _force_generic=generic)
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), g)
    assert not any(torch.isnan(t.float()).any() for t in (out, dq, dk, dv))
    oq = ok = 0
    for lq, lk in zip(lens_q, lens_k):
        sq_, sk_ = slice(oq, oq + lq), slice(ok, ok + lk)
        if lq and lk:
            ref = oracle_all(q.detach()[sq_][None], k.detach()[sk_][None], v.detach()[sk_][None], g[sq_][None])
            check_2x("out", out[sq_][None], *ref["out"])
            check_2x("dq", dq[sq_][None], *ref["dq"])
            check_2x("dk", dk[sk_][None], *ref["dk"])
            check_2x("dv", dv[sk_][None], *ref["dv"])
        else:
            assert (out[sq_] == 0).all() and (dq[sq_] == 0).all() and (dk[sk_] == 0).all() and (dv[sk_] == 0).all()
        oq += lq; ok += lk
