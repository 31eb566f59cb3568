"""Multi-GPU partition of a packed variable-length batch (BASELINE config 4 "sharded").

A ``cu_seqlens`` batch shards by whole sequences: every sequence is an independent unit for forward and backward, so
each rank runs the single-GPU ``flash_attn_varlen_func`` on its own sequences with its own ``cu_seqlens`` and there is
no data-path collective.  The reference has no multi-GPU form of the varlen path at all (its only multi-GPU attention
is head sharding in flash_attn/modules/mha.py:564-711); the contract here is the survey's: balance ranks by LPT
(longest processing time first) on the causal cost ``len_q * len_k`` -- the same longest-first idea the reference uses
INSIDE one GPU to order tiles (csrc/flashmask_v2/tile_scheduler.hpp:618-765).

Pure host logic (numpy-free, torch only for the index tensors); covered on CPU by tests/test_varlen_shard.py
(world-2 gloo with the oracle) and on the GPU by tests/test_varlen_shard_gpu.py and ``bench.py --workload c4 --gpus N``.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch


@dataclass
class VarlenShard:
    """What one rank owns: which sequences, where their tokens sit in the global packed tensors, and the rank-local
    ``cu_seqlens`` / ``max_seqlen`` that ``flash_attn_varlen_func`` takes."""
    rank: int
    seq_ids: List[int]                  # global sequence indices, in local order (longest first)
    cu_seqlens_q: torch.Tensor          # (n_local + 1,) int32, local packed offsets
    cu_seqlens_k: torch.Tensor
    max_seqlen_q: int
    max_seqlen_k: int
    q_index: torch.Tensor               # (local_total_q,) int64: rows of the global packed q / out / dq
    k_index: torch.Tensor               # (local_total_k,) int64: rows of the global packed k / v / dk / dv
    cost: int                           # sum of len_q * len_k over the local sequences (the LPT weight)

    def take_q(self, x: torch.Tensor) -> torch.Tensor:
        """Local packed tensor from the global packed one (q-like: q, dout)."""
        return x.index_select(0, self.q_index.to(x.device))

    def take_k(self, x: torch.Tensor) -> torch.Tensor:
        return x.index_select(0, self.k_index.to(x.device))

    def put_q(self, dst: torch.Tensor, local: torch.Tensor) -> None:
        """Scatter a local result (out, dq) back into the global packed tensor."""
        dst.index_copy_(0, self.q_index.to(dst.device), local)

    def put_k(self, dst: torch.Tensor, local: torch.Tensor) -> None:
        dst.index_copy_(0, self.k_index.to(dst.device), local)


def _lens(cu: Sequence[int]) -> List[int]:
    cu = [int(x) for x in (cu.tolist() if isinstance(cu, torch.Tensor) else cu)]
    if len(cu) < 1 or cu[0] != 0 or any(b < a for a, b in zip(cu, cu[1:])):
        raise ValueError("cu_seqlens must start at 0 and be non-decreasing")
    return [b - a for a, b in zip(cu, cu[1:])]


def lpt_assign(costs: Sequence[int], world: int) -> List[List[int]]:
    """Longest-processing-time-first: items by decreasing cost, each to the currently lightest rank (ties: lowest rank;
    equal costs keep their input order, so the plan is deterministic and identical on every rank)."""
    if world < 1:
        raise ValueError("world must be >= 1")
    order = sorted(range(len(costs)), key=lambda i: (-int(costs[i]), i))
    loads = [0] * world
    out: List[List[int]] = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda x: (loads[x], x))
        out[r].append(i)
        loads[r] += int(costs[i])
    return out


def shard_varlen(cu_seqlens_q, world: int, cu_seqlens_k=None, causal: bool = True,
                 device: Optional[torch.device] = None) -> List[VarlenShard]:
    """Plan for all ranks.  Cost of a sequence: the visible (query, key) pairs -- ``len_q * len_k`` halved on the causal
    triangle when ``len_q == len_k`` (the constant factor does not change the assignment; the triangle matters only
    for sequences with len_q != len_k).  Empty sequences (length 0) cost nothing and go wherever LPT puts them; every
    sequence appears on exactly one rank."""
    lq = _lens(cu_seqlens_q)
    lk = _lens(cu_seqlens_k) if cu_seqlens_k is not None else lq
    if len(lq) != len(lk):
        raise ValueError("cu_seqlens_q and cu_seqlens_k describe different batch sizes")
    cuq = [0]
    for n in lq:
        cuq.append(cuq[-1] + n)
    cuk = [0]
    for n in lk:
        cuk.append(cuk[-1] + n)

    def cost(i):
        a, b = lq[i], lk[i]
        if not causal:
            return a * b
        # top-left causal: row r sees min(b, r + 1) keys
        full = max(a - b, 0)
        tri = min(a, b)
        return full * b + tri * (tri + 1) // 2

    costs = [cost(i) for i in range(len(lq))]
    plans = []
    for r, ids in enumerate(lpt_assign(costs, world)):
        lcq, lck = [0], [0]
        qi, ki = [], []
        for i in ids:
            lcq.append(lcq[-1] + lq[i]); lck.append(lck[-1] + lk[i])
            qi.append(torch.arange(cuq[i], cuq[i + 1], dtype=torch.int64))
            ki.append(torch.arange(cuk[i], cuk[i + 1], dtype=torch.int64))
        cat = lambda xs: torch.cat(xs) if xs else torch.empty(0, dtype=torch.int64)
        plans.append(VarlenShard(
            rank=r, seq_ids=list(ids),
            cu_seqlens_q=torch.tensor(lcq, dtype=torch.int32, device=device),
            cu_seqlens_k=torch.tensor(lck, dtype=torch.int32, device=device),
            max_seqlen_q=max([lq[i] for i in ids], default=0), max_seqlen_k=max([lk[i] for i in ids], default=0),
            q_index=cat(qi), k_index=cat(ki), cost=sum(costs[i] for i in ids)))
    return plans


def flash_attn_varlen_sharded(q, k, v, cu_seqlens_q, cu_seqlens_k, rank: int, world: int, dropout_p=0.0,
                              softmax_scale=None, causal=False, attn_fn=None):
    """Rank ``rank``'s share of ``flash_attn_varlen_func`` over the GLOBAL packed tensors ``q, k, v`` (every rank
    holds, or can address, the whole batch -- the data-parallel loader hands each rank its sequences the same way).
    Returns ``(out_local, shard)``; ``out_local`` is differentiable w.r.t. ``q, k, v`` through the index_select, so a
    plain ``backward()`` leaves this rank's gradient rows in ``q.grad`` (zero rows elsewhere).  No collective."""
    if attn_fn is None:
        from .flash_attn_interface import flash_attn_varlen_func as attn_fn
    shard = shard_varlen(cu_seqlens_q, world, cu_seqlens_k, causal=causal, device=q.device)[rank]
    ql, kl, vl = shard.take_q(q), shard.take_k(k), shard.take_k(v)
    if ql.shape[0] == 0:
        return ql.new_zeros(ql.shape), shard
    out = attn_fn(ql, kl, vl, shard.cu_seqlens_q, shard.cu_seqlens_k, shard.max_seqlen_q, shard.max_seqlen_k,
                  dropout_p, softmax_scale, causal)
    return out, shard
