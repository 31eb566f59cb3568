"""CPU: the multi-GPU partition of a packed varlen batch (xhy-flash-attention_b200/varlen_shard.py):
LPT balance, exactly-once coverage, rank-local cu_seqlens, and -- world_size 2 over gloo with the oracle standing in
for the kernels -- that the ranks' results assemble into the single-process result, forward and backward."""
import importlib
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import attention_oracle as orc

vs = importlib.import_module("xhy-flash-attention_b200.varlen_shard")


def _cu(lens):
    out = [0]
    for n in lens:
        out.append(out[-1] + n)
    return torch.tensor(out, dtype=torch.int32)


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_plan_covers_every_sequence_once_and_balances(world):
    g = torch.Generator().manual_seed(0)
    lens = torch.randint(1, 400, (37,), generator=g).tolist() + [0, 0, 1]
    plans = vs.shard_varlen(_cu(lens), world)
    assert len(plans) == world
    seen = sorted(i for p in plans for i in p.seq_ids)
    assert seen == list(range(len(lens)))
    allq = torch.cat([p.q_index for p in plans]).sort().values
    assert torch.equal(allq, torch.arange(sum(lens)))
    for p in plans:
        assert p.cu_seqlens_q.tolist() == _cu([lens[i] for i in p.seq_ids]).tolist()
        assert p.max_seqlen_q == max([lens[i] for i in p.seq_ids], default=0)
        assert p.q_index.numel() == int(p.cu_seqlens_q[-1])
    # LPT guarantee: max load <= mean load + largest item
    costs = [n * (n + 1) // 2 for n in lens]
    assert max(p.cost for p in plans) <= sum(costs) / world + max(costs)
    assert sum(p.cost for p in plans) == sum(costs)


def test_lpt_is_deterministic_and_longest_first():
    a = vs.lpt_assign([5, 9, 9, 1, 7], 2)
    assert a == vs.lpt_assign([5, 9, 9, 1, 7], 2)
    assert a[0][0] == 1 and a[1][0] == 2          # the two 9s open the two ranks, input order breaks the tie
    assert sorted(a[0] + a[1]) == [0, 1, 2, 3, 4]


def test_c4_lengths_balance():
    """BASELINE config 4 (64 sequences, 262144 tokens): LPT on len^2 keeps 8 ranks within a few percent."""
    import bench
    lens = bench.c4_lengths()
    plans = vs.shard_varlen(_cu(lens), 8)
    loads = [p.cost for p in plans]
    assert max(loads) / (sum(loads) / 8) < 1.05


def test_rejects_bad_cu_seqlens():
    with pytest.raises(ValueError):
        vs.shard_varlen(torch.tensor([1, 2, 3]), 2)
    with pytest.raises(ValueError):
        vs.shard_varlen(torch.tensor([0, 5, 3]), 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _oracle_varlen(q, k, v, cu_q, cu_k, max_q, max_k, dropout_p, scale, causal):
    return orc.attention_varlen_ref(q, k, v, cu_q.tolist(), cu_k.tolist(), causal=causal, softmax_scale=scale)[0]


def _worker(rank, world, port, causal, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        lens = [17, 3, 40, 0, 25, 8, 31]
        cu = _cu(lens)
        total, h, hk, d = sum(lens), 4, 2, 16
        q = torch.randn(total, h, d, requires_grad=True)
        k = torch.randn(total, hk, d, requires_grad=True)
        v = torch.randn(total, hk, d, requires_grad=True)
        g = torch.randn(total, h, d)
        out_l, shard = vs.flash_attn_varlen_sharded(q, k, v, cu, cu, rank, world, causal=causal, attn_fn=_oracle_varlen)
        out_l.backward(shard.take_q(g))
        # assemble: every rank contributes its rows (disjoint), sum over ranks == the whole result
        out = torch.zeros(total, h, d)
        shard.put_q(out, out_l.detach())
        bufs = [out, q.grad.clone(), k.grad.clone(), v.grad.clone()]
        for t in bufs:
            dist.all_reduce(t)
        qr, kr, vr = [x.detach().clone().requires_grad_(True) for x in (q, k, v)]
        ref = orc.attention_varlen_ref(qr, kr, vr, cu.tolist(), cu.tolist(), causal=causal)[0]
        ref.backward(g)
        for name, got, want in zip(("out", "dq", "dk", "dv"), bufs, (ref.detach(), qr.grad, kr.grad, vr.grad)):
            torch.testing.assert_close(got, want, atol=1e-5, rtol=1e-5, msg=lambda m: f"{name}: {m}")
        ret[rank] = "ok"
    except Exception as e:  # noqa: BLE001
        import traceback
        ret[rank] = f"{type(e).__name__}: {e}\n{traceback.format_exc()}"
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("causal", [True, False])
def test_sharded_varlen_two_ranks_gloo(causal):
    world = 2
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), causal, ret), nprocs=world, join=True)
    assert dict(ret) == {r: "ok" for r in range(world)}, "\n".join(f"[{r}] {m}" for r, m in dict(ret).items())
