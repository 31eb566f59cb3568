"""CPU: pin the oracle restatement against golden vectors produced by the reference's own ``attention_ref``
(tests/golden/make_golden.py ran the function extracted from /root/reference/tests/test_flash_attn.py:111-159)."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import attention_oracle as orc

_ALL = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))
GOLDEN = [p for p in _ALL if not os.path.basename(p).startswith("rotary_")]
GOLDEN_ROTARY = [p for p in _ALL if os.path.basename(p).startswith("rotary_")]


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_oracle_matches_reference_golden(path):
    g = np.load(path)
    q, k, v, dout = [torch.from_numpy(g[n]).requires_grad_(n != "g") for n in ("q", "k", "v", "g")]
    qpm = torch.from_numpy(g["query_padding_mask"]) if "query_padding_mask" in g else None
    kpm = torch.from_numpy(g["key_padding_mask"]) if "key_padding_mask" in g else None
    causal = bool(g["causal"])
    out, attn = orc.attention_ref(q, k, v, qpm, kpm, 0.0, None, causal=causal)
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), dout)
    for name, got in (("out", out), ("attn", attn), ("dq", dq), ("dk", dk), ("dv", dv)):
        np.testing.assert_allclose(got.detach().numpy(), g[name], rtol=1e-5, atol=1e-6, err_msg=name)


@pytest.mark.parametrize("path", [p for p in GOLDEN if "padded" not in p],
                         ids=[os.path.basename(p)[:-4] for p in GOLDEN if "padded" not in p])
def test_masked_oracle_matches_golden(path):
    """attention_masked_ref (the general oracle used for every mask kind) agrees with the pinned attention_ref."""
    g = np.load(path)
    q, k, v = [torch.from_numpy(g[n]) for n in ("q", "k", "v")]
    sq, sk = q.shape[1], k.shape[1]
    mask = orc.causal_mask_top_left(sq, sk) if bool(g["causal"]) else None
    out, lse = orc.attention_masked_ref(q, k, v, mask)
    np.testing.assert_allclose(out.numpy(), g["out"], rtol=1e-5, atol=1e-6)
    # lse against logsumexp of the golden attention: attn = exp(s - lse)  =>  sum(attn) == 1 on non-empty rows
    assert torch.isfinite(lse).all()


def test_mask_builders():
    assert orc.causal_mask_top_left(3, 5).tolist() == [[False, True, True, True, True],
                                                       [False, False, True, True, True],
                                                       [False, False, False, True, True]]
    assert orc.causal_mask_bottom_right(2, 4).tolist() == [[False, False, False, True], [False, False, False, False]]
    m = orc.local_mask(4, 4, 1, 0)
    assert m.tolist() == [[False, True, True, True], [False, False, True, True],
                          [True, False, False, True], [True, True, False, False]]
    lts = orc.document_lt_start([2, 3])
    assert lts.tolist() == [2, 2, 5, 5, 5]
    fm = orc.flashmask_dense(5, 5, lts)
    # causal document mask: row i masked for col j iff i >= end(doc(j))
    assert fm[2, 0].item() and not fm[1, 0].item() and not fm[4, 2].item()


def test_combine_ref_matches_full():
    torch.manual_seed(0)
    q = torch.randn(1, 3, 2, 16); k = torch.randn(1, 64, 2, 16); v = torch.randn(1, 64, 2, 16)
    full, lse_full = orc.attention_masked_ref(q, k, v)
    parts = [orc.attention_masked_ref(q, k[:, a:b], v[:, a:b], lse_empty=float("-inf")) for a, b in ((0, 20), (20, 64))]
    o = torch.stack([p[0].float() for p in parts]).permute(0, 1, 3, 2, 4)  # (S,b,h,sq,d)
    l = torch.stack([p[1] for p in parts])                                  # (S,b,h,sq)
    oc, lc = orc.combine_ref(o, l)
    torch.testing.assert_close(oc.permute(0, 2, 1, 3), full, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(lc, lse_full, rtol=1e-5, atol=1e-6)


def test_flop_convention():
    # benchmarks/benchmark_flash_attention.py:27-30
    assert orc.attention_flops(8, 2048, 2048, 16, 128, False, "fwd") == 4 * 8 * 2048 * 2048 * 16 * 128
    assert orc.attention_flops(1, 8192, 8192, 32, 128, True, "fwd_bwd") == 3.5 * 4 * 8192 * 8192 * 32 * 128 // 2


@pytest.mark.parametrize("path", GOLDEN_ROTARY, ids=[os.path.basename(p)[:-4] for p in GOLDEN_ROTARY])
def test_rotary_oracle_matches_reference_golden(path):
    """rotary_ref against the reference's apply_rotary_emb_torch (flash_attn/layers/rotary.py:22-32), positions 0..s-1,
    and with an offset: rotating rows at positions o..o+s-1 equals the golden of the shifted table."""
    g = np.load(path)
    x, cos, sin = [torch.from_numpy(g[n]) for n in ("x", "cos", "sin")]
    b, s = x.shape[:2]
    pos = torch.arange(s).view(1, s).expand(b, s)
    out = orc.rotary_ref(x, cos, sin, pos, interleaved=bool(g["interleaved"]))
    np.testing.assert_allclose(out.numpy(), g["out"], rtol=1e-5, atol=1e-6)
    half = s // 2    # second half of the rows alone, at offset `half`
    out2 = orc.rotary_ref(x[:, half:], cos, sin, pos[:, half:], interleaved=bool(g["interleaved"]))
    np.testing.assert_allclose(out2.numpy(), g["out"][:, half:], rtol=1e-5, atol=1e-6)


def test_kvcache_ref_reduces_to_plain_attention():
    """kvcache_ref with no extras is attention over the first cache_seqlens keys; paged / batch-idx / leftpad
    addressing returns the same numbers as the equivalent contiguous cache."""
    torch.manual_seed(0)
    b, sq, h, hk, d, sk = 3, 2, 4, 2, 32, 64
    q = torch.randn(b, sq, h, d)
    kc, vc = torch.randn(b, sk, hk, d), torch.randn(b, sk, hk, d)
    lens = torch.tensor([64, 17, 40], dtype=torch.int32)
    out, lse, _, _ = orc.kvcache_ref(q, kc, vc, cache_seqlens=lens, causal=True)
    for i in range(b):
        n = int(lens[i])
        o_i, l_i = orc.attention_masked_ref(q[i:i + 1], kc[i:i + 1, :n], vc[i:i + 1, :n],
                                            orc.causal_mask_bottom_right(sq, n), lse_empty=float("-inf"))
        np.testing.assert_allclose(out[i:i + 1].numpy(), o_i.numpy(), rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose(lse[i:i + 1].numpy(), l_i.numpy(), rtol=1e-6, atol=1e-6)
    # paged: 16-row pages in a shuffled pool
    ps = 16
    perm = torch.randperm(b * sk // ps)
    table = perm.view(b, sk // ps).to(torch.int32)
    pool_k, pool_v = torch.empty(b * sk // ps, ps, hk, d), torch.empty(b * sk // ps, ps, hk, d)
    pool_k[perm] = kc.view(b * sk // ps, ps, hk, d)
    pool_v[perm] = vc.view(b * sk // ps, ps, hk, d)
    out_p, lse_p, _, _ = orc.kvcache_ref(q, pool_k, pool_v, cache_seqlens=lens, block_table=table, causal=True)
    np.testing.assert_allclose(out_p.numpy(), out.numpy(), rtol=1e-6, atol=1e-6)
    # cache_batch_idx: reversed cache rows
    idx = torch.tensor([2, 1, 0], dtype=torch.int32)
    out_i, _, _, _ = orc.kvcache_ref(q, kc.flip(0), vc.flip(0), cache_seqlens=lens, cache_batch_idx=idx, causal=True)
    np.testing.assert_allclose(out_i.numpy(), out.numpy(), rtol=1e-6, atol=1e-6)
    # leftpad: shift the valid keys right by 5
    lp = torch.tensor([0, 5, 3], dtype=torch.int32)
    kc2, vc2 = torch.randn(b, sk + 8, hk, d), torch.randn(b, sk + 8, hk, d)
    lens2 = lens.clone()
    for i in range(b):
        n, o = int(lens[i]), int(lp[i])
        n = min(n, sk + 8 - o)
        kc2[i, o:o + n], vc2[i, o:o + n] = kc[i, :n], vc[i, :n]
        lens2[i] = o + n
    out_l, _, _, _ = orc.kvcache_ref(q, kc2, vc2, cache_seqlens=lens2, cache_leftpad=lp, causal=True)
    np.testing.assert_allclose(out_l.numpy(), out.numpy(), rtol=1e-6, atol=1e-6)
