"""GPU: the sharded varlen path over the CUDA kernels.  The ranks of a 4-way plan run one after another on cuda:0 (the
path has no collective, so ranks never wait for one another); the assembled result must equal the unsharded kernel
call bit for bit (same kernels, same per-sequence arithmetic) and the oracle under the 2x rule."""
import importlib

import pytest
import torch

from oracle import attention_oracle as orc
from tests._util import check_2x

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("world", [2, 4])
def test_sharded_varlen_matches_unsharded(fa, world, causal):
    vs = importlib.import_module("xhy-flash-attention_b200.varlen_shard")
    torch.manual_seed(0)
    lens = [300, 17, 129, 0, 512, 64, 1, 255, 384]
    cu = torch.tensor([0] + torch.tensor(lens).cumsum(0).tolist(), dtype=torch.int32, device=DEV)
    total, h, hk, d = sum(lens), 8, 2, 128
    q = torch.randn(total, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(total, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(total, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    g = torch.randn(total, h, d, device=DEV, dtype=torch.bfloat16)
    out = torch.zeros_like(q)
    for rank in range(world):
        out_l, shard = fa.flash_attn_varlen_sharded(q, k, v, cu, cu, rank, world, causal=causal)
        out_l.backward(shard.take_q(g))
        shard.put_q(out, out_l.detach())
    dq, dk, dv = q.grad.clone(), k.grad.clone(), v.grad.clone()
    q.grad = k.grad = v.grad = None
    whole = fa.flash_attn_varlen_func(q, k, v, cu, cu, max(lens), max(lens), 0.0, None, causal)
    wq, wk, wv = torch.autograd.grad(whole, (q, k, v), g)
    assert torch.equal(out, whole.detach())
    for a, b in ((dq, wq), (dk, wk), (dv, wv)):
        torch.testing.assert_close(a.float(), b.float(), atol=2e-2, rtol=2e-2)    # fp32 dQ adds land in another order
    ref = orc.attention_varlen_ref(q.detach(), k.detach(), v.detach(), cu.tolist(), cu.tolist(), causal=causal)[0]
    pt = orc.attention_varlen_ref(q.detach(), k.detach(), v.detach(), cu.tolist(), cu.tolist(), causal=causal,
                                  upcast=False, reorder_ops=True)[0]
    check_2x("out", out, ref, pt)
