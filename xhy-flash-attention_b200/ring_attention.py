"""Ring attention: the one place where the hot path needs an exchange step (BASELINE config 5a).

A long sequence is sharded over the ``P`` ranks of a process group (one process per GPU).  K/V blocks travel round
the ring with point-to-point ``send``/``recv`` (NCCL over NVLink on the GPU box, posted on a side stream through
``torch.distributed.batch_isend_irecv`` so that the transfer of the next block overlaps the attention of the
current one); each rank runs the single-GPU kernels (``fa_b200_fwd`` / ``fa_b200_bwd``) on
``(Q_local, KV_current)`` and folds the partial result into its running ``(O, LSE)`` with the LSE-merge kernel
(``fa_b200_lse_merge``: the same algebra as the split-KV combine, csrc/flash_attn_v3/flash_fwd_combine_kernel.h:340-376).
The reference has no ring attention; parity is pinned against the single-device result (tests/test_ring_*.py).

Causal sharding is zig-zag: the sequence is cut into ``2P`` chunks and rank ``r`` owns chunks ``r`` and
``2P-1-r``, which balances the causal work.  With that layout every ring step is one of three dense cases:

    step 0 (own K/V)            : causal attention over the local pair of chunks (top-left == bottom-right, sq == sk)
    K/V from a rank j <  r      : all local queries see the FIRST chunk of that K/V, nothing of the second
    K/V from a rank j >  r      : only the SECOND local query chunk sees that K/V -- all of it

so no step needs a mask other than "causal" or "none", and half of each step's tile work is skipped structurally.

Backward rotates ``(K, V)`` the same way and lets the fp32 ``(dK, dV)`` accumulators travel with their K/V block;
``dQ`` stays local.  After ``P`` hops every accumulator is back on the rank that owns the block.

Host logic (schedule, peers, merge order) is independent of the device: ``block_fwd`` / ``block_bwd`` / ``merge``
can be replaced (the CPU ``gloo`` tests plug the oracle in); the defaults are the CUDA kernels and raise if the
library is missing.
"""
from __future__ import annotations

import os
from typing import Callable, Optional, Tuple

import torch
import torch.distributed as dist

from . import _capi
from .flash_attn_interface import _flash_attn_backward, _flash_attn_forward


# ----------------------------------------------------------------------------------------------------
# sharding helpers
# ----------------------------------------------------------------------------------------------------
def zigzag_shard(x: torch.Tensor, world: int, rank: int, dim: int = 1) -> torch.Tensor:
    """Return rank ``rank``'s zig-zag shard (chunks ``rank`` and ``2*world-1-rank`` of ``2*world``) along ``dim``."""
    n = x.shape[dim]
    if n % (2 * world) != 0:
        raise ValueError(f"sequence length {n} must be a multiple of 2 * world_size = {2 * world}")
    chunks = x.chunk(2 * world, dim=dim)
    return torch.cat([chunks[rank], chunks[2 * world - 1 - rank]], dim=dim).contiguous()


def zigzag_unshard(shards, dim: int = 1) -> torch.Tensor:
    """Inverse of :func:`zigzag_shard` given the list of all ranks' shards (rank order)."""
    world = len(shards)
    out = [None] * (2 * world)
    for r, s in enumerate(shards):
        a, b = s.chunk(2, dim=dim)
        out[r], out[2 * world - 1 - r] = a, b
    return torch.cat(out, dim=dim)


def contiguous_shard(x: torch.Tensor, world: int, rank: int, dim: int = 1) -> torch.Tensor:
    return x.chunk(world, dim=dim)[rank].contiguous()


# ----------------------------------------------------------------------------------------------------
# default block functions: the CUDA kernels
# ----------------------------------------------------------------------------------------------------
def _cuda_block_fwd(q, k, v, softmax_scale, causal):
    out, _, _, _, _, lse, _, _ = _flash_attn_forward(q, k, v, 0.0, softmax_scale, causal, False)
    return out, lse                                            # lse: (b, h, sq) fp32, +inf for rows that see nothing


def _cuda_block_bwd(dout, q, k, v, out, lse, softmax_scale, causal):
    # Tried in round 1 and removed: cutting a short-key step (8 ranks: 512 CTAs = 3.46 waves on 148 SMs) into two query
    # halves on two streams to fill the last wave.  Measured on 8 GPUs: 72.6 ms per step against 69.8 ms unsplit -- the
    # second set of dPsum / dQ-convert launches and the fp32 sum of the dK/dV partials cost more than the wave gains.
    dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
    _flash_attn_backward(dout, q, k, v, out, lse, dq, dk, dv, 0.0, softmax_scale, causal)
    return dq, dk, dv


def _cuda_merge(o_acc, lse_acc, o_step, lse_step):
    """(o_acc fp32 (b, s, h, d), lse_acc fp32 (b, s, h)) <- merge with (o_step 16-bit (b, s, h, d), lse_step (b, h, s)).
    One launch for the whole batch; ``o_acc`` / ``lse_acc`` may be views along the sequence (the second half of the
    running output): only their batch stride differs from a contiguous tensor's."""
    lib = _capi.load()
    b, s, h, d = o_acc.shape
    assert o_acc.stride()[1:] == (h * d, d, 1) and lse_acc.stride()[1:] == (h, 1)
    assert o_step.stride()[1:] == (h * d, d, 1) and lse_step.stride()[1:] == (s, 1)
    is_bf16 = 1 if o_step.dtype == torch.bfloat16 else 0
    with torch.cuda.device(o_acc.device):
        _capi.check(lib.fa_b200_lse_merge_batched(
            o_acc.data_ptr(), lse_acc.data_ptr(), o_step.data_ptr(), lse_step.data_ptr(), b, s, h, d, is_bf16, 1,
            o_acc.stride(0), o_step.stride(0), lse_acc.stride(0), h, 1, lse_step.stride(0), 1, s,
            _capi.stream_ptr(o_acc.device)), "fa_b200_lse_merge_batched")


class RingOps:
    """The three device operations the ring schedule is built from."""

    def __init__(self, block_fwd: Callable = _cuda_block_fwd, block_bwd: Callable = _cuda_block_bwd,
                 merge: Callable = _cuda_merge, exchange: bool = True):
        self.block_fwd, self.block_bwd, self.merge = block_fwd, block_bwd, merge
        # exchange=False replays the identical schedule (same block shapes, masks, merges) WITHOUT send/recv: every step
        # attends the rank's own K/V.  Results are meaningless; the run time is the compute-only time that
        # bench.py subtracts from the real step to report the exposed communication.
        self.exchange = exchange


_CUDA_OPS = RingOps()


# ----------------------------------------------------------------------------------------------------
# communication: one rotation of a list of tensors to rank+1, received from rank-1
# ----------------------------------------------------------------------------------------------------
RESERVE_SMS = int(os.environ.get("FA_B200_RING_RESERVE_SMS", "8"))   # SMs left to NCCL during the ring forward (0 = none)
STALL_TRACE = None     # set to a list to collect (tag, event before, event after) around every wait for a transfer


class _Rotation:
    """Posts send(to next) / recv(from prev) for a list of tensors; ``wait()`` returns the received tensors.

    On CUDA the operations are posted from a side stream so that they overlap the attention kernels running on the
    caller's stream; ``wait()`` makes the caller's stream wait for the transfer.
    """

    def __init__(self, group, tensors, comm_stream=None, exchange=True, tag=""):
        self.tag = tag
        if not exchange:                      # compute-only replay (RingOps.exchange)
            self.recv, self.reqs, self.staged, self.stream, self.keep = list(tensors), [], None, None, []
            return
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        nxt = dist.get_global_rank(group, (rank + 1) % world) if group is not None else (rank + 1) % world
        prv = dist.get_global_rank(group, (rank - 1) % world) if group is not None else (rank - 1) % world
        self.recv = [torch.empty_like(t) for t in tensors]
        self.stream = comm_stream
        self.keep = tensors
        self.staged = None
        if tensors and tensors[0].is_cuda and dist.get_backend(group) == "gloo":
            # debug transport for boxes with fewer GPUs than ranks (several ranks share one device, which NCCL refuses):
            # the blocks go through host memory.  Never taken with the nccl backend.
            self.staged = self.recv
            tensors = [t.cpu() for t in tensors]
            self.recv = [torch.empty_like(t) for t in tensors]
            self.stream = comm_stream = None
        ops = []
        for t, r in zip(tensors, self.recv):
            ops.append(dist.P2POp(dist.isend, t, nxt, group))
            ops.append(dist.P2POp(dist.irecv, r, prv, group))
        if comm_stream is not None:
            comm_stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(comm_stream):
                self.reqs = dist.batch_isend_irecv(ops)
        else:
            self.reqs = dist.batch_isend_irecv(ops)

    def wait(self):
        for r in self.reqs:
            r.wait()
        if self.staged is not None:
            for dst, src in zip(self.staged, self.recv):
                dst.copy_(src)
            return self.staged
        if self.stream is not None:
            cur = torch.cuda.current_stream()
            if STALL_TRACE is not None:          # diagnostics: how long the compute stream sits in this wait
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(cur)
                cur.wait_stream(self.stream)
                e1.record(cur)
                STALL_TRACE.append((self.tag, e0, e1))
            else:
                cur.wait_stream(self.stream)
            for t in self.keep + self.recv:      # allocated on the compute stream, used by the comm stream
                t.record_stream(self.stream)
        return self.recv


def _comm_stream(t: torch.Tensor):
    return torch.cuda.Stream(device=t.device) if t.is_cuda else None


def _step_plan(causal: bool, rank: int, src: int):
    """Which (query rows, key rows, causal flag) a ring step computes.

    Returns ``(q_part, k_part, causal_flag)`` with parts in {"all", "first", "second"}, or ``None`` if the step has no
    visible (query, key) pair.  ``src`` is the rank whose K/V block is being visited.
    """
    if not causal:
        return "all", "all", False
    if src == rank:
        return "all", "all", True
    if src < rank:
        return "all", "first", False
    return "second", "all", False


def _part(x: torch.Tensor, part: str, dim: int = 1) -> torch.Tensor:
    if part == "all":
        return x
    half = x.shape[dim] // 2
    return x.narrow(dim, 0 if part == "first" else half, half)


def ring_forward(q, k, v, softmax_scale, causal, group=None, ops: RingOps = _CUDA_OPS):
    """Forward over local shards ``q, k, v`` (batch, s_local, heads, d).  Returns ``(out, lse)`` with
    ``out`` in the input dtype and ``lse`` fp32 (batch, heads, s_local)."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    b, s, h, d = q.shape
    o_acc = torch.zeros((b, s, h, d), dtype=torch.float32, device=q.device)
    lse_acc = torch.full((b, s, h), float("inf"), dtype=torch.float32, device=q.device)   # +inf = nothing merged yet
    comm = _comm_stream(q)
    k_cur, v_cur = k.contiguous(), v.contiguous()
    # While a K/V block travels, NCCL's send / recv kernels hold a few SMs.  The forward kernel is persistent (one CTA per
    # SM, tile list cut statically): the CTAs that find their SM taken start late and the launch ends late by as much --
    # measured on 8 GPUs as 9 % of the step with the compute stream never waiting for data (tools/ring_trace.py).  Leaving
    # those SMs free costs the forward RESERVE_SMS / 148 of its throughput and removes the delay (70.2 -> 67.3 ms).
    # Worth it while the lost throughput (RESERVE_SMS / 148 of a step's forward, ~1 PFLOP/s) is below the delay it avoids
    # (one K/V block at ~250 GB/s): blocks up to ~32k tokens per rank at 32 / 8 heads of 128.
    reserve = 0
    if ops.exchange and world > 1 and q.is_cuda and ops.block_fwd is _cuda_block_fwd and RESERVE_SMS > 0:
        t_fwd = 4.0 * b * s * (s / 2) * h * d / 1.0e15
        t_xfer = 2.0 * k.numel() * k.element_size() / 2.5e11
        if RESERVE_SMS / 148.0 * t_fwd < t_xfer:
            reserve = RESERVE_SMS
    if reserve:
        _capi.load().fa_b200_set_forward_sm_reserve(reserve)
    try:
        return _ring_forward_steps(q, k_cur, v_cur, o_acc, lse_acc, comm, softmax_scale, causal, group, ops, world, rank)
    finally:
        if reserve:
            _capi.load().fa_b200_set_forward_sm_reserve(0)


def _ring_forward_steps(q, k_cur, v_cur, o_acc, lse_acc, comm, softmax_scale, causal, group, ops, world, rank):
    for t in range(world):
        rot = _Rotation(group, [k_cur, v_cur], comm, ops.exchange, f"fwd kv {t}") if t + 1 < world else None
        src = (rank - t) % world
        plan = _step_plan(causal, rank, src)
        if plan is not None:
            qp, kp, cz = plan
            # halves are strided views (batch stride of the whole shard): the kernels take them as they are
            o_step, lse_step = ops.block_fwd(_part(q, qp), _part(k_cur, kp), _part(v_cur, kp), softmax_scale, cz)
            ops.merge(_part(o_acc, qp), _part(lse_acc, qp), o_step, lse_step)
        if rot is not None:
            k_cur, v_cur = rot.wait()
    return o_acc.to(q.dtype), lse_acc.transpose(1, 2).contiguous()


def ring_backward(dout, q, k, v, out, lse, softmax_scale, causal, group=None, ops: RingOps = _CUDA_OPS):
    """Backward over local shards.  ``out``/``lse`` are the FINAL forward results (global softmax statistics), so every
    block's ``P = exp(S - LSE)`` and ``dPsum = rowsum(dO o O)`` are already the global ones and block gradients add up
    (csrc/flash_attn/src/flash_bwd_kernel.h:847,897-906).  Returns ``(dq, dk, dv)`` in the input dtype."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    comm = _comm_stream(q)
    dq = torch.zeros(q.shape, dtype=torch.float32, device=q.device)
    k_cur, v_cur = k.contiguous(), v.contiguous()
    dk_acc = torch.zeros(k.shape, dtype=torch.float32, device=q.device)   # accumulators of the block being visited
    dv_acc = torch.zeros(v.shape, dtype=torch.float32, device=q.device)
    acc_rot = None
    half = q.shape[1] // 2
    for t in range(world):
        kv_rot = _Rotation(group, [k_cur, v_cur], comm, ops.exchange, f"fwd kv {t}") if t + 1 < world else None
        src = (rank - t) % world
        plan = _step_plan(causal, rank, src)
        dk_blk = dv_blk = None
        if plan is not None:
            qp, kp, cz = plan
            lse_p = _part(lse, qp, dim=2).contiguous()      # (b, h, rows): the kernels index it densely
            dq_p, dk_p, dv_p = ops.block_bwd(_part(dout, qp), _part(q, qp), _part(k_cur, kp), _part(v_cur, kp),
                                             _part(out, qp), lse_p, softmax_scale, cz)
            _part(dq, qp).add_(dq_p)
            dk_blk, dv_blk = (kp, dk_p), (kp, dv_p)
        if acc_rot is not None:                      # accumulators of this block, as left by the previous rank
            dk_acc, dv_acc = acc_rot.wait()
        if dk_blk is not None:
            _part(dk_acc, dk_blk[0]).add_(dk_blk[1])
            _part(dv_acc, dv_blk[0]).add_(dv_blk[1])
        acc_rot = _Rotation(group, [dk_acc, dv_acc], comm, ops.exchange, f"bwd acc {t}")   # travels with the block; after `world` hops it is home
        if kv_rot is not None:
            k_cur, v_cur = kv_rot.wait()
    dk_acc, dv_acc = acc_rot.wait()
    return dq.to(q.dtype), dk_acc.to(k.dtype), dv_acc.to(v.dtype)


class RingFlashAttnFunc(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, softmax_scale, causal, group, ops):
        if softmax_scale is None:
            softmax_scale = q.shape[-1] ** (-0.5)
        out, lse = ring_forward(q, k, v, softmax_scale, causal, group, ops)
        ctx.save_for_backward(q, k, v, out, lse)
        ctx.softmax_scale, ctx.causal, ctx.group, ctx.ops = softmax_scale, causal, group, ops
        return out, lse

    @staticmethod
    def backward(ctx, dout, *args):
        q, k, v, out, lse = ctx.saved_tensors
        dq, dk, dv = ring_backward(dout.contiguous(), q, k, v, out, lse, ctx.softmax_scale, ctx.causal, ctx.group,
                                   ctx.ops)
        return dq, dk, dv, None, None, None, None


def ring_flash_attn_func(q, k, v, dropout_p=0.0, softmax_scale=None, causal=False, return_attn_probs=False, *,
                         group=None, ops: Optional[RingOps] = None):
    """``flash_attn_func`` over a sequence sharded across ``group`` (signature follows the reference's
    ``flash_attn_func``, flash_attn/flash_attn_interface.py:345-373, plus ``group``).

    ``q, k, v`` are this rank's shards (batch, s_local, heads, d): zig-zag shards (:func:`zigzag_shard`) when
    ``causal`` -- s_local must then be even -- and contiguous shards otherwise.  Returns the local shard of the output
    (and of the LSE with ``return_attn_probs``).
    """
    if dropout_p != 0.0:
        raise NotImplementedError("ring attention: dropout is not implemented")
    if causal and q.shape[1] % 2 != 0:
        raise ValueError("ring attention (causal, zig-zag): local sequence length must be even")
    if not dist.is_initialized():
        raise RuntimeError("ring attention needs an initialised torch.distributed process group")
    out, lse = RingFlashAttnFunc.apply(q, k, v, softmax_scale, causal, group, ops or _CUDA_OPS)
    return (out, lse, None) if return_attn_probs else out
