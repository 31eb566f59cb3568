"""CPU: the padded <-> packed helpers (reference flash_attn/bert_padding.py:97-135) against a plain boolean-mask
restatement, including gradients, and against the padding masks stored with the golden vectors."""
import glob
import importlib
import os

import numpy as np
import torch

bp = importlib.import_module("xhy-flash-attention_b200.bert_padding")


def test_unpad_pad_roundtrip_and_grads():
    torch.manual_seed(0)
    b, s, h, d = 4, 9, 3, 8
    x = torch.randn(b, s, h, d, requires_grad=True)
    lens = torch.tensor([9, 1, 5, 0])
    mask = torch.arange(s).view(1, s) < lens.view(b, 1)
    packed, idx, cu, mx = bp.unpad_input(x, mask)
    assert packed.shape == (int(lens.sum()), h, d) and mx == 9
    assert cu.dtype == torch.int32 and cu.tolist() == [0, 9, 10, 15, 15]
    assert torch.equal(packed, x[mask])                       # same rows, same order as boolean indexing
    back = bp.pad_input(packed, idx, b, s)
    assert torch.equal(back, x * mask.view(b, s, 1, 1))
    g = torch.randn_like(back)
    (gx,) = torch.autograd.grad(back, x, g)
    assert torch.equal(gx, g * mask.view(b, s, 1, 1))         # gradient flows through valid positions only


def test_unpad_matches_golden_padding_masks():
    """The golden fixtures of the padded cases carry the reference's own key/query padding masks."""
    here = os.path.join(os.path.dirname(__file__), "golden")
    for path in sorted(glob.glob(os.path.join(here, "padded_*.npz"))):
        gld = np.load(path)
        q, m = torch.from_numpy(gld["q"]), torch.from_numpy(gld["query_padding_mask"])
        packed, idx, cu, mx = bp.unpad_input(q, m)
        assert cu[-1].item() == int(m.sum()) and mx == int(m.sum(-1).max())
        assert torch.equal(bp.pad_input(packed, idx, q.shape[0], q.shape[1]), q * m.view(*m.shape, 1, 1))
