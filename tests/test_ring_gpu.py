"""GPU: ring attention over the CUDA kernels + the LSE-merge kernel, 2 ranks.  With >= 2 GPUs the ranks use one device
each and NCCL send/recv; on a single-GPU box both ranks share cuda:0 and the blocks are staged through host memory
over gloo (NCCL refuses two ranks on one device) -- the kernels, the schedule and the merge are the same."""
import importlib
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, causal, b, ret):
    try:
        from oracle import attention_oracle as orc
        from tests._util import check_2x, check_lse
        fa = importlib.import_module("xhy-flash-attention_b200")
        ring = importlib.import_module("xhy-flash-attention_b200.ring_attention")
        ndev = torch.cuda.device_count()
        dev = torch.device("cuda", rank % ndev)
        torch.cuda.set_device(dev)
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        if ndev >= world:
            dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
            assert dist.get_backend() == "nccl"       # >= 2 GPUs visible: the K/V blocks must travel over NCCL
        else:
            dist.init_process_group("gloo", rank=rank, world_size=world)
        torch.manual_seed(0)
        s, h, hk, d = 1024 * world, 8, 2, 128
        q = torch.randn(b, s, h, d, device=dev, dtype=torch.bfloat16)
        k = torch.randn(b, s, hk, d, device=dev, dtype=torch.bfloat16)
        v = torch.randn(b, s, hk, d, device=dev, dtype=torch.bfloat16)
        g = torch.randn(b, s, h, d, device=dev, dtype=torch.bfloat16)
        shard = (lambda x: ring.zigzag_shard(x, world, rank)) if causal else (lambda x: ring.contiguous_shard(x, world, rank))
        ql, kl, vl = [shard(x).requires_grad_(True) for x in (q, k, v)]
        out, lse, _ = ring.ring_flash_attn_func(ql, kl, vl, causal=causal, return_attn_probs=True)
        dq, dk, dv = torch.autograd.grad(out, (ql, kl, vl), shard(g))
        torch.cuda.synchronize()
        mask = orc.causal_mask_top_left(s, s).to(dev) if causal else None
        ref = orc.attention_grads_ref(q, k, v, g, mask)
        pt = orc.attention_grads_ref(q, k, v, g, mask, upcast=False, reorder_ops=True)
        names = ("out", "lse", "dq", "dk", "dv")
        for n, got, r_, p_ in zip(names, (out, lse, dq, dk, dv), ref, pt):
            if n == "lse":
                check_lse(got, shard(r_.transpose(1, 2)).transpose(1, 2))
            else:
                check_2x(n, got, shard(r_), shard(p_), factor=2.0, atol=2e-3 if n != "out" else 1e-4)
        # and against the single-device kernel on the whole sequence
        qf, kf, vf = [x.clone().requires_grad_(True) for x in (q, k, v)]
        of = fa.flash_attn_func(qf, kf, vf, causal=causal)
        torch.testing.assert_close(out.float(), shard(of.detach()).float(), atol=2e-2, rtol=2e-2)
        ret[rank] = "ok"
    except Exception as e:
        import traceback
        ret[rank] = f"{type(e).__name__}: {e}\n{traceback.format_exc()}"
    finally:
        if dist.is_initialized():
            dist.destroy_process_group()


@pytest.mark.parametrize("b", [1, 2])
@pytest.mark.parametrize("causal", [True, False])
def test_ring_attention_two_ranks(fa, causal, b):
    # b = 2: the causal "second half" steps merge into strided views of the running output (one batched launch)
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), causal, b, ret), nprocs=world, join=True)
    assert dict(ret) == {r: "ok" for r in range(world)}, "\n".join(f"[{r}] {m}" for r, m in dict(ret).items())
