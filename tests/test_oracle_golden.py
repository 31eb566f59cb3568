"""CPU: pin the oracle restatement against golden vectors produced by the reference's own ``attention_ref``
(tests/golden/make_golden.py ran the function extracted from /root/reference/tests/test_flash_attn.py:111-159)."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import attention_oracle as orc

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


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
