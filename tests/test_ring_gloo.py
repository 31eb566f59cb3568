"""CPU, world_size 2 and 4 over gloo: the HOST logic of ring attention (zig-zag sharding, which K/V block is visited
at which step with which mask, the LSE merge order, dK/dV accumulators travelling with their block) checked against
single-process attention.  The device operations are replaced by the oracle here (RingOps hooks); on the GPU box the
same schedule runs over the CUDA kernels and NCCL (tests/test_ring_gpu.py)."""
import importlib
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import attention_oracle as orc

ring = importlib.import_module("xhy-flash-attention_b200.ring_attention")


def _oracle_ops():
    def block_fwd(q, k, v, scale, causal):
        mask = orc.causal_mask_top_left(q.shape[1], k.shape[1]) if causal else None
        return orc.attention_masked_ref(q, k, v, mask, scale)

    def block_bwd(dout, q, k, v, out, lse, scale, causal):
        mask = orc.causal_mask_top_left(q.shape[1], k.shape[1]) if causal else None
        return orc.attention_block_bwd_ref(dout, q, k, v, out, lse, mask, scale)

    def merge(o_acc, lse_acc, o_step, lse_step):
        o, l = orc.merge_ref(o_acc, lse_acc, o_step, lse_step.transpose(1, 2))
        o_acc.copy_(o)
        lse_acc.copy_(l)

    return ring.RingOps(block_fwd, block_bwd, merge)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, causal, hk, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        b, s, h, d = 2, 16 * world, 4, 32
        q = torch.randn(b, s, h, d)
        k = torch.randn(b, s, hk, d)
        v = torch.randn(b, s, hk, d)
        g = torch.randn(b, s, h, d)
        shard = (lambda x: ring.zigzag_shard(x, world, rank)) if causal else (lambda x: ring.contiguous_shard(x, world, rank))
        ql, kl, vl = [shard(x).requires_grad_(True) for x in (q, k, v)]
        out, lse, _ = ring.ring_flash_attn_func(ql, kl, vl, causal=causal, return_attn_probs=True, ops=_oracle_ops())
        dq, dk, dv = torch.autograd.grad(out, (ql, kl, vl), shard(g))
        mask = orc.causal_mask_top_left(s, s) if causal else None
        ro, rl, rdq, rdk, rdv = orc.attention_grads_ref(q, k, v, g, mask)
        for name, got, ref in (("out", out, ro), ("dq", dq, rdq), ("dk", dk, rdk), ("dv", dv, rdv)):
            torch.testing.assert_close(got, shard(ref), atol=2e-5, rtol=1e-4, msg=lambda m: f"{name} rank {rank}: {m}")
        lse_ref = shard(rl.transpose(1, 2)).transpose(1, 2)
        torch.testing.assert_close(lse, lse_ref, atol=1e-5, rtol=1e-5)
        ret[rank] = "ok"
    except Exception as e:  # surfaced by the parent
        ret[rank] = f"{type(e).__name__}: {e}"
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,causal,hk", [(2, True, 4), (2, False, 1), (4, True, 1), (4, False, 4)])
def test_ring_schedule_matches_single_process(world, causal, hk):
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), causal, hk, ret), nprocs=world, join=True)
    assert dict(ret) == {r: "ok" for r in range(world)}, dict(ret)


def test_zigzag_roundtrip():
    x = torch.arange(2 * 24 * 3).view(2, 24, 3)
    for world in (1, 2, 3, 4):
        shards = [ring.zigzag_shard(x, world, r) for r in range(world)]
        assert torch.equal(ring.zigzag_unshard(shards), x)
    with pytest.raises(ValueError):
        ring.zigzag_shard(x, 5, 0)


def test_step_plan_covers_every_causal_pair_once():
    """Union over ranks and steps of the (q chunk, k chunk) pairs the plan visits == the causal lower triangle."""
    for world in (2, 3, 8):
        seen = set()
        for rank in range(world):
            qc = {"first": [rank], "second": [2 * world - 1 - rank], "all": [rank, 2 * world - 1 - rank]}
            for t in range(world):
                src = (rank - t) % world
                kc = {"first": [src], "second": [2 * world - 1 - src], "all": [src, 2 * world - 1 - src]}
                qp, kp, cz = ring._step_plan(True, rank, src)
                for a in qc[qp]:
                    for b in kc[kp]:
                        if cz and b > a:
                            continue
                        assert (a, b) not in seen
                        seen.add((a, b))
        assert seen == {(a, b) for a in range(2 * world) for b in range(2 * world) if b <= a}
