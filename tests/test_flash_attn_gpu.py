"""GPU parity tests of the operator API against the oracle (mirrors the reference's tests/test_flash_attn.py:
seeded randn inputs, mha/mqa/gqa, odd/unequal sequence lengths, head dims incl. non-multiples of 8, fp16/bf16,
causal, packed variants, varlen via padding masks; outputs, LSE and gradients under the 2x rule)."""
import importlib
import math

import pytest
import torch

from oracle import attention_oracle as orc
from tests._util import check_2x, check_lse, oracle_all

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _inputs(b, sq, sk, h, hk, d, dtype, seed=0):
    torch.manual_seed(seed)
    q = torch.randn(b, sq, h, d, device=DEV, dtype=dtype, requires_grad=True)
    k = torch.randn(b, sk, hk, d, device=DEV, dtype=dtype, requires_grad=True)
    v = torch.randn(b, sk, hk, d, device=DEV, dtype=dtype, requires_grad=True)
    g = torch.randn(b, sq, h, d, device=DEV, dtype=dtype)
    return q, k, v, g


def _run_and_check(fa, q, k, v, g, causal, generic, mask=None, **kw):
    out, lse, _ = fa.flash_attn_func(q, k, v, 0.0, None, causal, True, _force_generic=generic, **kw)
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), g)
    sq, sk = q.shape[1], k.shape[1]
    if mask is None and causal:
        mask = orc.causal_mask_top_left(sq, sk).to(DEV)
    ref = oracle_all(q.detach(), k.detach(), v.detach(), g, mask)
    check_2x("out", out, *ref["out"])
    check_lse(lse, ref["lse"])
    check_2x("dq", dq, *ref["dq"])
    check_2x("dk", dk, *ref["dk"])
    check_2x("dv", dv, *ref["dv"])


# the reference's grids: head dims tests/test_flash_attn.py:336,527; (seqlen_q, seqlen_k) :531; seqlen :341
REF_HEAD_DIMS = [32, 40, 59, 64, 80, 96, 111, 128, 160, 192, 224, 256]
REF_SEQ_PAIRS = [(113, 203), (128, 217), (113, 211), (108, 256), (256, 512), (512, 256), (1024, 1024), (1023, 1024),
                 (1024, 1023), (2048, 2048)]
REF_SEQLENS = [97, 128, 200, 256, 257, 384, 512, 768, 1024, 1025, 2048]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "fp16"])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", REF_HEAD_DIMS)
@pytest.mark.parametrize("sq,sk", REF_SEQ_PAIRS + [(384, 384)])
def test_flash_attn_output(fa, capi, sq, sk, d, causal, dtype):
    # the reference's whole (shape, head dim) grid; gqa with 6 q-heads, 2 kv-heads
    q, k, v, g = _inputs(2, sq, sk, 6, 2, d, dtype)
    capi.fallback_count(reset=True)
    _run_and_check(fa, q, k, v, g, causal, False)
    # every head dim of the reference's grid runs on the tcgen05 kernels, forward AND backward (the backward runs on
    # autograd's thread: the process-wide fallback counter sees it)
    assert capi.fallback_count() == 0, "a pass left the tcgen05 kernels"
    fa.flash_attn_func(q.detach(), k.detach(), v.detach(), 0.0, None, causal)
    assert capi.last_kernel_kind() == 2, f"forward left the tcgen05 kernels: {capi.last_fallback_reason()}"


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "fp16"])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", [32, 59, 64, 96, 128, 160, 256])
@pytest.mark.parametrize("sq,sk", [(113, 203), (128, 217), (256, 512), (108, 256), (1023, 1024), (512, 256), (384, 384)])
def test_flash_attn_output_generic(fa, sq, sk, d, causal, dtype):
    # the SIMT family as an independent second CUDA implementation
    q, k, v, g = _inputs(2, sq, sk, 6, 2, d, dtype)
    _run_and_check(fa, q, k, v, g, causal, True)


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", REF_HEAD_DIMS)
@pytest.mark.parametrize("seqlen", REF_SEQLENS)
def test_flash_attn_qkvpacked_grid(fa, seqlen, d, causal):
    # tests/test_flash_attn.py:333-347 (dropout_p = 0 half; the dropout half lives in tests/test_dropout_gpu.py)
    torch.manual_seed(0)
    b, h = 4 if seqlen < 2048 else 2, 4
    qkv = torch.randn(b, seqlen, 3, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    g = torch.randn(b, seqlen, h, d, device=DEV, dtype=torch.bfloat16)
    out, lse, _ = fa.flash_attn_qkvpacked_func(qkv, 0.0, None, causal, True)
    (dqkv,) = torch.autograd.grad(out, qkv, g)
    mask = orc.causal_mask_top_left(seqlen, seqlen).to(DEV) if causal else None
    q, k, v = qkv.detach().unbind(2)
    ref = oracle_all(q, k, v, g, mask)
    check_2x("out", out, *ref["out"])
    check_lse(lse, ref["lse"])
    for i, n in enumerate(("dq", "dk", "dv")):
        check_2x(n, dqkv[:, :, i], *ref[n])


@pytest.mark.parametrize("mha_type", ["mha", "mqa", "gqa"])
@pytest.mark.parametrize("causal", [False, True])
def test_flash_attn_head_types(fa, mha_type, causal):
    h = 9
    hk = {"mha": 9, "mqa": 1, "gqa": 3}[mha_type]  # tests/test_flash_attn.py:521,543
    q, k, v, g = _inputs(4, 257, 257, h, hk, 128, torch.bfloat16)
    _run_and_check(fa, q, k, v, g, causal, False)


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", [64, 128])
def test_flash_attn_qkvpacked(fa, d, causal):
    torch.manual_seed(0)
    b, s, h = 3, 300, 4
    qkv = torch.randn(b, s, 3, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    g = torch.randn(b, s, h, d, device=DEV, dtype=torch.bfloat16)
    out, lse, _ = fa.flash_attn_qkvpacked_func(qkv, 0.0, None, causal, True)
    (dqkv,) = torch.autograd.grad(out, qkv, g)
    mask = orc.causal_mask_top_left(s, s).to(DEV) if causal else None
    q, k, v = qkv.detach().unbind(2)
    ref = oracle_all(q, k, v, g, mask)
    check_2x("out", out, *ref["out"])
    check_lse(lse, ref["lse"])
    for i, n in enumerate(("dq", "dk", "dv")):
        check_2x(n, dqkv[:, :, i], *ref[n])


@pytest.mark.parametrize("causal", [False, True])
def test_flash_attn_kvpacked(fa, causal):
    torch.manual_seed(0)
    b, sq, sk, h, hk, d = 2, 200, 333, 8, 2, 128
    q = torch.randn(b, sq, h, d, device=DEV, dtype=torch.float16, requires_grad=True)
    kv = torch.randn(b, sk, 2, hk, d, device=DEV, dtype=torch.float16, requires_grad=True)
    g = torch.randn(b, sq, h, d, device=DEV, dtype=torch.float16)
    out = fa.flash_attn_kvpacked_func(q, kv, 0.0, None, causal)
    dq, dkv = torch.autograd.grad(out, (q, kv), g)
    mask = orc.causal_mask_top_left(sq, sk).to(DEV) if causal else None
    ref = oracle_all(q.detach(), kv.detach()[:, :, 0], kv.detach()[:, :, 1], g, mask)
    check_2x("out", out, *ref["out"])
    check_2x("dq", dq, *ref["dq"])
    check_2x("dk", dkv[:, :, 0], *ref["dk"])
    check_2x("dv", dkv[:, :, 1], *ref["dv"])


@pytest.mark.parametrize("seqlen", [1, 2, 5, 17, 128])
@pytest.mark.parametrize("causal", [False, True])
def test_flash_attn_bwd_overflow(fa, seqlen, causal):
    """tests/test_flash_attn.py:835-878: tiny sequences with scaled-up inputs must not produce NaN."""
    torch.manual_seed(0)
    d, h = 128, 5
    q = (torch.randn(2, seqlen, h, d, device=DEV, dtype=torch.bfloat16) * 5).requires_grad_()
    k, v = [(torch.randn(2, seqlen, h, d, device=DEV, dtype=torch.bfloat16) * 3).requires_grad_() for _ in range(2)]
    g = torch.randn(2, seqlen, h, d, device=DEV, dtype=torch.bfloat16)
    out = fa.flash_attn_func(q, k, v, causal=causal)
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), g)
    mask = orc.causal_mask_top_left(seqlen, seqlen).to(DEV) if causal else None
    ref = oracle_all(q.detach(), k.detach(), v.detach(), g, mask)
    for n, got in (("out", out), ("dq", dq), ("dk", dk), ("dv", dv)):
        assert not got.isnan().any()
        check_2x(n, got, *ref[n], factor=5.0, atol=1e-3)  # the reference uses 5x + 1e-3 here (:875-878)


def test_flash_attn_bwd_transpose(fa):
    """tests/test_flash_attn.py:881-926: non-contiguous dout (a transposed view) is accepted."""
    torch.manual_seed(0)
    b, s, h, d = 5, 512, 2, 128
    q, k, v = [torch.randn(b, s, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True) for _ in range(3)]
    out = fa.flash_attn_func(q, k, v, causal=True)
    g = torch.randn(b, h, s, d, device=DEV, dtype=torch.bfloat16).transpose(1, 2)  # strides != contiguous
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), g)
    ref = oracle_all(q.detach(), k.detach(), v.detach(), g, orc.causal_mask_top_left(s, s).to(DEV))
    for n, got in (("dq", dq), ("dk", dk), ("dv", dv)):
        check_2x(n, got, *ref[n])


def test_flash_attn_race_condition(fa):
    """tests/test_flash_attn.py:786-832: out/lse/dk/dv bit-identical across reruns; dq within a tight bound."""
    torch.manual_seed(0)
    q, k, v, g = _inputs(4, 512, 512, 4, 4, 128, torch.bfloat16)
    out0, lse0, _ = fa.flash_attn_func(q, k, v, 0.0, None, True, True)
    dq0, dk0, dv0 = torch.autograd.grad(out0, (q, k, v), g)
    for _ in range(20):
        out, lse, _ = fa.flash_attn_func(q, k, v, 0.0, None, True, True)
        dq, dk, dv = torch.autograd.grad(out, (q, k, v), g)
        assert torch.equal(out, out0) and torch.equal(lse, lse0)
        assert torch.equal(dk, dk0) and torch.equal(dv, dv0)
        assert (dq - dq0).abs().max().item() <= 2 * (dq0.abs().max().item() * 2 ** -7)


@pytest.mark.parametrize("causal", [False, True])
def test_host_pipeline_matches_device_path(fa, causal):
    """HostPipeline (pinned host in / out, chunked over three streams) == flash_attn_func on the device, bit for bit."""
    torch.manual_seed(0)
    b, s, h, hk, d = 5, 640, 4, 2, 128
    mk = lambda *shape: torch.randn(*shape, dtype=torch.bfloat16).pin_memory()
    q, k, v, g = mk(b, s, h, d), mk(b, s, hk, d), mk(b, s, hk, d), mk(b, s, h, d)
    outs = [torch.empty_like(t).pin_memory() for t in (q, q, k, v)]
    pipe = fa.HostPipeline("cuda", chunk=2, causal=causal)
    for _ in range(2):   # second run reuses the staging buffers
        pipe.run(q, k, v, g, *outs)
    pipe.synchronize()
    qd, kd, vd = [t.cuda().requires_grad_(True) for t in (q, k, v)]
    o = fa.flash_attn_func(qd, kd, vd, causal=causal)
    dq, dk, dv = torch.autograd.grad(o, (qd, kd, vd), g.cuda())
    assert torch.equal(outs[0], o.detach().cpu())
    assert torch.equal(outs[2], dk.cpu()) and torch.equal(outs[3], dv.cpu())
    torch.testing.assert_close(outs[1].float(), dq.cpu().float(), atol=2e-2, rtol=2e-2)   # dQ: fp32 atomics order
    with pytest.raises(ValueError):
        pipe.run(q.cuda(), k, v, g, *outs)


def test_training_step_in_cuda_graph(fa):
    """Forward + backward captured into a CUDA graph and replayed on new data in the same static buffers."""
    torch.manual_seed(0)
    b, s, h, hk, d = 2, 512, 8, 2, 128
    q = torch.randn(b, s, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(b, s, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(b, s, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    g = torch.randn(b, s, h, d, device=DEV, dtype=torch.bfloat16)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        torch.autograd.grad(fa.flash_attn_func(q, k, v, causal=True), (q, k, v), g)
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        out = fa.flash_attn_func(q, k, v, causal=True)
        dq, dk, dv = torch.autograd.grad(out, (q, k, v), g)
    for _ in range(2):
        with torch.no_grad():
            for t in (q, k, v, g):
                t.copy_(torch.randn_like(t))
        graph.replay()
        ref = oracle_all(q.detach(), k.detach(), v.detach(), g, orc.causal_mask_top_left(s, s).to(DEV))
        check_2x("out", out, *ref["out"])
        check_2x("dq", dq, *ref["dq"])
        check_2x("dk", dk, *ref["dk"])
        check_2x("dv", dv, *ref["dv"])


@pytest.mark.parametrize("case", ["causal_gqa", "full", "window", "varlen"])
def test_backward_deterministic(fa, case):
    """``deterministic`` (flat C ABI: num_splits == 1, csrc/flash_attn/src/flash_bwd_kernel.h:3203-3208; v3 handle:
    dq_semaphore): the fp32 adds into every dQ tile happen in key-block order, so repeated runs give the same BITS for
    dq as well (dk / dv always do); the values agree with the default, unordered accumulation."""
    itf = importlib.import_module("xhy-flash-attention_b200.flash_attn_interface")
    torch.manual_seed(0)
    b, s, h, hk, d = 3, 1000, 8, 2, 128
    kw = {}
    mask = itf._NO_MASK
    causal = case in ("causal_gqa", "varlen")
    if case == "varlen":
        lens = [300, 1000, 77]
        cu = torch.tensor([0, 300, 1300, 1377], device=DEV, dtype=torch.int32)
        shape_q, shape_k = (sum(lens), h, d), (sum(lens), hk, d)
        kw = dict(cu_seqlens_q=cu, cu_seqlens_k=cu, max_seqlen_q=max(lens), max_seqlen_k=max(lens))
    else:
        shape_q, shape_k = (b, s, h, d), (b, s, hk, d)
    if case == "window":
        mask = itf._Mask(window=(200, 50))
    q = torch.randn(shape_q, device=DEV, dtype=torch.bfloat16)
    k = torch.randn(shape_k, device=DEV, dtype=torch.bfloat16)
    v = torch.randn(shape_k, device=DEV, dtype=torch.bfloat16)
    g = torch.randn(shape_q, device=DEV, dtype=torch.bfloat16)
    scale = d ** -0.5
    out, qp, kp, vp, op, lse, _, _ = itf._flash_attn_forward(q, k, v, 0.0, scale, causal, False, mask,
                                                              *([kw[n] for n in ("cu_seqlens_q", "cu_seqlens_k", "max_seqlen_q", "max_seqlen_k")] if kw else []))

    def bwd(det):
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        itf._flash_attn_backward(g, qp, kp, vp, op, lse, dq, dk, dv, 0.0, scale, causal, mask=mask, deterministic=det, **kw)
        return dq, dk, dv

    ref = bwd(False)
    runs = [bwd(True) for _ in range(4)]
    for r in runs[1:]:
        for a, b_ in zip(runs[0], r):
            assert torch.equal(a, b_), "deterministic backward changed bits between runs"
    for a, b_ in zip(runs[0], ref):
        torch.testing.assert_close(a.float(), b_.float(), atol=2e-2, rtol=2e-2)
    assert torch.equal(runs[0][1], ref[1]) and torch.equal(runs[0][2], ref[2])   # dk, dv never depended on the order


@pytest.mark.parametrize("packed", [False, True])
@pytest.mark.parametrize("causal", [False, True])
def test_backward_expanded_head_grid(fa, causal, packed, monkeypatch):
    """The shallow-grid mode of the backward (one CTA per (key block, q-head), dK/dV summed over the GQA group
    afterwards) must give what the grouped mode gives; forced on here, chosen by the wave model in production."""
    iface = importlib.import_module("xhy-flash-attention_b200.flash_attn_interface")
    torch.manual_seed(0)
    b, sq, sk, h, hk, d = 2, 300, 520, 8, 2, 128
    q = torch.randn(b, sq, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    kv = torch.randn(b, sk, 2, hk, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    g = torch.randn(b, sq, h, d, device=DEV, dtype=torch.bfloat16)
    res = {}
    for mode in (False, True):
        monkeypatch.setattr(iface, "_FORCE_EXPANDED_HEADS", mode)
        if packed:      # dk/dv are strided slices of one packed gradient
            out = fa.flash_attn_kvpacked_func(q, kv, 0.0, None, causal)
            dq, dkv = torch.autograd.grad(out, (q, kv), g)
            res[mode] = (dq, dkv[:, :, 0], dkv[:, :, 1])
        else:
            k, v = kv[:, :, 0], kv[:, :, 1]
            out = fa.flash_attn_func(q, k, v, 0.0, None, causal)
            dq, dkv = torch.autograd.grad(out, (q, kv), g)
            res[mode] = (dq, dkv[:, :, 0], dkv[:, :, 1])
    mask = orc.causal_mask_top_left(sq, sk).to(DEV) if causal else None
    ref = oracle_all(q.detach(), kv.detach()[:, :, 0], kv.detach()[:, :, 1], g, mask)
    for i, n in enumerate(("dq", "dk", "dv")):
        check_2x(n, res[True][i], *ref[n], atol=2e-4)
        torch.testing.assert_close(res[True][i].float(), res[False][i].float(), atol=3e-2, rtol=3e-2)


def test_prefer_expanded_heads_wave_model():
    iface = importlib.import_module("xhy-flash-attention_b200.flash_attn_interface")
    dev = torch.device("cuda", 0)
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    if sms != 148:
        pytest.skip("wave model numbers below are for 148 SMs")
    assert iface._prefer_expanded_heads(1, 8192, 32, 8, dev)        # 512 CTAs = 3.46 waves -> 2048 CTAs = 13.8 waves
    assert not iface._prefer_expanded_heads(8, 8192, 32, 8, dev)    # BASELINE config 3: 4096 CTAs, deep enough
    assert not iface._prefer_expanded_heads(2, 512, 8, 2, dev)      # under one wave
