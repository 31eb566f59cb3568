"""End-to-end path for callers whose tensors live in HOST memory.

``flash_attn_func`` takes device tensors, like the reference.  When q, k, v (and dout) start in pinned host memory and
the results must end there -- the way ``bench.py`` measures its ``e2e`` number -- the PCIe copies (about 1.3 GB each
way for BASELINE config 3) cost three times as long as the kernels if they are done one after the other.
:class:`HostPipeline` cuts the batch into chunks and runs three streams: chunk *i+1* is copied host->device while
chunk *i* runs forward + backward and the results of chunk *i-1* are copied device->host (PCIe is full duplex and the
two directions use separate copy engines).  Inside a chunk the order is finer still: q/k/v travel before dout so the
forward starts while dout is on the wire, and ``out`` leaves for the host while the backward is running.  The batch dimension is independent for attention, so
chunking changes nothing in the results.  The kernels are the same ``fa_b200_fwd`` / ``fa_b200_bwd``; only the
orchestration differs.
"""
from __future__ import annotations

from typing import Optional

import torch

from .flash_attn_interface import _NO_MASK, _flash_attn_backward, _flash_attn_forward


class HostPipeline:
    """Forward + backward of ``flash_attn_func`` from pinned host inputs to pinned host outputs.

    ``run(q, k, v, dout, out, dq, dk, dv)``: all arguments are pinned host tensors, ``q/dout/out/dq`` of shape
    (batch, seqlen_q, heads, d), ``k/v/dk/dv`` of shape (batch, seqlen_k, heads_k, d).  Device staging buffers are
    allocated once and double-buffered.  The call returns after enqueueing; ``synchronize()`` waits for the results.
    """

    def __init__(self, device, chunk: int = 1, causal: bool = False, softmax_scale: Optional[float] = None):
        self.device = torch.device(device)
        self.chunk, self.causal, self.softmax_scale = chunk, causal, softmax_scale
        self.s_in = torch.cuda.Stream(self.device)
        self.s_run = torch.cuda.Stream(self.device)
        self.s_out = torch.cuda.Stream(self.device)
        self._bufs = None
        self._key = None
        self._in_free = self._out_free = None

    def _ensure(self, q, k):
        key = (tuple(q.shape[1:]), tuple(k.shape[1:]), q.dtype)
        if self._key == key:
            return
        c, dev, dt = self.chunk, self.device, q.dtype
        mk_q = lambda: torch.empty((c,) + tuple(q.shape[1:]), device=dev, dtype=dt)
        mk_k = lambda: torch.empty((c,) + tuple(k.shape[1:]), device=dev, dtype=dt)
        # two sets of {q, k, v, dout} inputs and {out, dq, dk, dv} outputs
        self._bufs = [dict(q=mk_q(), k=mk_k(), v=mk_k(), do=mk_q(), out=None, dq=mk_q(), dk=mk_k(), dv=mk_k())
                      for _ in range(2)]
        self._in_free = [None, None]    # event: the compute that read input set s has finished
        self._out_free = [None, None]   # event: the device->host copy of output set s has finished
        self._key = key

    def run(self, q, k, v, dout, out, dq, dk, dv):
        for t in (q, k, v, dout, out, dq, dk, dv):
            if t.is_cuda or not t.is_pinned():
                raise ValueError("HostPipeline.run takes pinned host tensors")
        self._ensure(q, k)
        b, c = q.shape[0], self.chunk
        scale = self.softmax_scale if self.softmax_scale is not None else q.shape[-1] ** (-0.5)
        cur = torch.cuda.current_stream(self.device)   # inputs are host tensors: nothing on `cur` has to finish first
        for i, lo in enumerate(range(0, b, c)):
            hi = min(lo + c, b)
            n = hi - lo
            B = self._bufs[i % 2]
            # host->device: q, k, v first (all the forward needs), dout behind them
            with torch.cuda.stream(self.s_in):
                if self._in_free[i % 2] is not None:
                    self.s_in.wait_event(self._in_free[i % 2])
                for name, src in (("k", k), ("v", v), ("q", q)):
                    B[name][:n].copy_(src[lo:hi], non_blocking=True)
                qkv_ready = torch.cuda.Event()
                qkv_ready.record(self.s_in)
                B["do"][:n].copy_(dout[lo:hi], non_blocking=True)
                do_ready = torch.cuda.Event()
                do_ready.record(self.s_in)
            # compute: the forward starts while dout is still on the wire
            with torch.cuda.stream(self.s_run):
                self.s_run.wait_event(qkv_ready)
                if self._out_free[i % 2] is not None:
                    self.s_run.wait_event(self._out_free[i % 2])
                o, qp, kp, vp, op, lse, _, _ = _flash_attn_forward(B["q"][:n], B["k"][:n], B["v"][:n], 0.0, scale,
                                                                   self.causal, False, _NO_MASK)
                B["out"] = o
                fwd_done = torch.cuda.Event()
                fwd_done.record(self.s_run)
                self.s_run.wait_event(do_ready)
                _flash_attn_backward(B["do"][:n], qp, kp, vp, op, lse, B["dq"][:n], B["dk"][:n], B["dv"][:n], 0.0, scale,
                                     self.causal, mask=_NO_MASK)
                done = torch.cuda.Event()
                done.record(self.s_run)
                self._in_free[i % 2] = done
            # device->host: out leaves while the backward runs, the gradients after it
            with torch.cuda.stream(self.s_out):
                self.s_out.wait_event(fwd_done)
                out[lo:hi].copy_(B["out"][:n], non_blocking=True)
                self.s_out.wait_event(done)
                for name, dst in (("dq", dq), ("dk", dk), ("dv", dv)):
                    dst[lo:hi].copy_(B[name][:n], non_blocking=True)
                B["out"].record_stream(self.s_out)
                freed = torch.cuda.Event()
                freed.record(self.s_out)
                self._out_free[i % 2] = freed
        for s_ in (self.s_in, self.s_run, self.s_out):
            cur.wait_stream(s_)

    def synchronize(self):
        torch.cuda.current_stream(self.device).synchronize()
