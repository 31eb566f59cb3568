"""GPU parity: dense additive attn_mask on the two varlen flat symbols (flash_attn_varlen_fwd / flash_attn_varlen_bwd,
include/fa_capi.h; reference csrc/capi/flash_attn.cu:690-790,960-1100): packed tensors + cu_seqlens, the mask is
(b, 1|h, 1|max_seqlen_q, max_seqlen_k) and each sequence reads its own top-left corner of it.  Also the first
flat-symbol varlen run on the tensor-core kernels: the symbols carry no totals, so the library patches the row extent of
its TMA tensor maps on the device from cu_seqlens[b] (no host readback, no stream synchronisation: the call only
enqueues, also under CUDA-graph capture), and the fp32 dq_accum workspace has exactly the reference's size."""
import ctypes as C

import pytest
import torch

from oracle import attention_oracle as orc
from tests._util import check_2x, check_lse

pytestmark = pytest.mark.gpu
DEV = "cuda"
vp = lambda t: None if t is None else C.c_void_p(t.data_ptr())
i32 = C.c_int
r128 = lambda x: (x + 127) // 128 * 128


def _seq(q, k, v, g, add, scale, upcast):
    q_, k_, v_ = (t.detach().clone().unsqueeze(0).requires_grad_(True) for t in (q, k, v))
    out, lse = orc.attention_masked_ref(q_, k_, v_, None, scale, upcast, not upcast,
                                        bias=add.float() if upcast else add)
    dq, dk, dv = torch.autograd.grad(out, (q_, k_, v_), g.unsqueeze(0).to(out.dtype))
    return out[0].detach(), lse[0], dq[0], dk[0], dv[0]


@pytest.mark.parametrize("d,expect_kind", [(128, 2), (64, 2), (160, 1)])
@pytest.mark.parametrize("mask_hs", [(1, True), (4, False), None])
def test_varlen_dense_mask(capi, d, expect_kind, mask_hs):
    lib = capi.load()
    torch.manual_seed(0)
    dtype = torch.bfloat16
    lens_q, lens_k = [300, 128, 77], [300, 256, 150]
    b, h, hk = 3, 4, 2
    msq, msk = max(lens_q), max(lens_k)
    tq, tk = sum(lens_q), sum(lens_k)
    q = torch.randn(tq, h, d, device=DEV, dtype=dtype)
    k = torch.randn(tk, hk, d, device=DEV, dtype=dtype)
    v = torch.randn(tk, hk, d, device=DEV, dtype=dtype)
    g = torch.randn(tq, h, d, device=DEV, dtype=dtype)
    cu_q = torch.tensor([0] + list(torch.tensor(lens_q).cumsum(0)), device=DEV, dtype=torch.int32)
    cu_k = torch.tensor([0] + list(torch.tensor(lens_k).cumsum(0)), device=DEV, dtype=torch.int32)
    mask = dims = None
    if mask_hs is not None:
        mh, per_row = mask_hs
        keep = torch.rand(b, mh, msq if per_row else 1, msk, device=DEV) > 0.3
        keep[..., 0] = True
        mask = (torch.where(keep, 0.0, -1e4) + torch.randn(keep.shape, device=DEV)).to(dtype).contiguous()
        dims = (C.c_int64 * 4)(*mask.shape)
    scale = d ** -0.5
    out = torch.empty_like(q)
    lse = torch.empty(b, h, msq, device=DEV, dtype=torch.float32)
    st = lambda t: (i32(t.stride(0)), i32(t.stride(1)), i32(0))      # (row, head, batch) strides of a packed tensor
    (qr, qh, qb), (kr, kh, kb), (vr, vh, vb), (or_, oh, ob) = st(q), st(k), st(v), st(out)
    lib.flash_attn_varlen_fwd.restype = C.c_bool
    ok = lib.flash_attn_varlen_fwd(
        vp(q), vp(k), vp(v), vp(cu_q), vp(cu_k), None, vp(out), None, vp(lse), i32(b), i32(msq), i32(msk),
        i32(r128(msq)), i32(r128(msk)), i32(h), i32(hk), i32(d), i32((d + 31) // 32 * 32), C.c_float(0.0),
        C.c_float(scale), C.c_float(1.0 / scale), C.c_bool(False), C.c_bool(False), C.c_bool(True), None,
        C.c_uint64(0), C.c_uint64(0), vp(mask), dims, qr, kr, vr, qh, kh, vh, or_, oh, qb, kb, vb, ob, C.c_bool(False))
    assert ok, capi.last_error()
    if d > 128 and mask_hs is None:
        expect_kind = 2               # plain calls with head dims 129..256 run on the tcgen05 kernels too
    assert capi.last_kernel_kind() == expect_kind
    dq = torch.empty_like(q)
    dk = torch.empty(tk, h, d, device=DEV, dtype=dtype)              # expanded: one dK/dV head per query head
    dv = torch.empty(tk, h, d, device=DEV, dtype=dtype)
    softmax_d = torch.empty(b, h, r128(msq), device=DEV, dtype=torch.float32)
    # the reference's size for this workspace: (b, h, seqlen_q_rounded, head_size_rounded) fp32 (flash_api.cpp:836-838)
    dq_accum = torch.empty(b * h * r128(msq) * ((d + 31) // 32 * 32) + 1, device=DEV, dtype=torch.float32)
    dq_accum[-1] = 12345.0                                            # canary right behind the reference-sized buffer
    (dqr, dqh, dqb), (dkr, dkh, dkb), (dvr, dvh, dvb), (gr, gh, gb) = st(dq), st(dk), st(dv), st(g)
    lib.flash_attn_varlen_bwd.restype = C.c_bool
    ok = lib.flash_attn_varlen_bwd(
        vp(g), vp(q), vp(k), vp(v), vp(out), vp(softmax_d), vp(lse), vp(cu_q), vp(cu_k), None, vp(dq), vp(dk), vp(dv),
        vp(dq_accum), i32(b), i32(msq), i32(msk), i32(r128(msq)), i32(r128(msk)), i32(h), i32(hk), i32(d),
        i32((d + 31) // 32 * 32), C.c_float(0.0), C.c_float(scale), C.c_float(1.0 / scale), C.c_bool(False),
        C.c_bool(True), i32(0), None, C.c_uint64(0), C.c_uint64(0), vp(mask), dims,
        qr, kr, vr, qh, kh, vh, or_, oh, qb, kb, vb, ob, dqr, dkr, dvr, dqh, dkh, dvh, gr, gh, dqb, dkb, dvb, gb,
        C.c_bool(False))
    assert ok, capi.last_error()
    assert capi.last_kernel_kind() == expect_kind
    torch.cuda.synchronize()
    assert dq_accum[-1].item() == 12345.0
    grp = h // hk
    for i in range(b):
        sq, sk = lens_q[i], lens_k[i]
        rq, rk = slice(int(cu_q[i]), int(cu_q[i + 1])), slice(int(cu_k[i]), int(cu_k[i + 1]))
        # the kernel adds mask * unscale to q.k before scaling: the additive term on the scaled logits is the mask itself
        add = (torch.zeros(1, 1, 1, 1, device=DEV, dtype=dtype) if mask is None else
               mask[i:i + 1, :, :sq if per_row else 1, :sk].expand(1, mh, sq, sk))
        ref = _seq(q[rq], k[rk], v[rk], g[rq], add, scale, True)
        pt = _seq(q[rq], k[rk], v[rk], g[rq], add, scale, False)
        check_2x(f"out[{i}]", out[rq], ref[0], pt[0])
        check_lse(lse[i, :, :sq], ref[1])
        check_2x(f"dq[{i}]", dq[rq], ref[2], pt[2], atol=1e-3)
        check_2x(f"dk[{i}]", dk[rk].view(sk, hk, grp, d).float().sum(2), ref[3], pt[3], atol=2e-2)
        check_2x(f"dv[{i}]", dv[rk].view(sk, hk, grp, d).float().sum(2), ref[4], pt[4], atol=2e-2)


def test_flat_varlen_under_graph_capture(capi):
    """The reference's contract for the flat symbols is "enqueue and return" (csrc/capi/flash_attn.cu:576-684).  Captured
    into a CUDA graph the call must stay on the tcgen05 kernels (no cu_seqlens readback) and replay correctly after the
    inputs -- and the sequence lengths -- have been rewritten in place."""
    lib = capi.load()
    torch.manual_seed(0)
    dtype, d, b, h, hk = torch.bfloat16, 128, 3, 4, 2
    msq = msk = 300
    cap = 3 * 300
    q = torch.randn(cap, h, d, device=DEV, dtype=dtype)
    k = torch.randn(cap, hk, d, device=DEV, dtype=dtype)
    v = torch.randn(cap, hk, d, device=DEV, dtype=dtype)
    cu = torch.tensor([0, 300, 428, 505], device=DEV, dtype=torch.int32)
    out = torch.empty_like(q)
    lse = torch.empty(b, h, msq, device=DEV, dtype=torch.float32)
    scale = d ** -0.5
    st = lambda t: (i32(t.stride(0)), i32(t.stride(1)), i32(0))
    (qr, qh, qb), (kr, kh, kb), (vr, vh, vb), (or_, oh, ob) = st(q), st(k), st(v), st(out)
    lib.flash_attn_varlen_fwd.restype = C.c_bool

    def call(stream):
        return lib.flash_attn_varlen_fwd(
            vp(q), vp(k), vp(v), vp(cu), vp(cu), None, vp(out), None, vp(lse), i32(b), i32(msq), i32(msk),
            i32(r128(msq)), i32(r128(msk)), i32(h), i32(hk), i32(d), i32(d), C.c_float(0.0), C.c_float(scale),
            C.c_float(1.0 / scale), C.c_bool(True), C.c_bool(False), C.c_bool(True), C.c_void_p(stream.cuda_stream),
            C.c_uint64(0), C.c_uint64(0), None, None, qr, kr, vr, qh, kh, vh, or_, oh, qb, kb, vb, ob, C.c_bool(False))

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        assert call(side), capi.last_error()          # warm-up outside capture (sets the kernel attributes)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    capi.fallback_count(reset=True)
    with torch.cuda.graph(graph, stream=side):
        assert call(side), capi.last_error()
    assert capi.last_kernel_kind() == 2 and capi.fallback_count() == 0, capi.last_fallback_reason()
    for lens in ([300, 128, 77], [100, 300, 250]):     # second replay: other lengths, written into the same cu_seqlens
        cu.copy_(torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32))
        q.normal_(); k.normal_(); v.normal_()
        out.zero_()
        graph.replay()
        torch.cuda.synchronize()
        ref, _ = orc.attention_varlen_ref(q, k, v, cu.tolist(), cu.tolist(), causal=True)
        pt, _ = orc.attention_varlen_ref(q, k, v, cu.tolist(), cu.tolist(), causal=True, upcast=False, reorder_ops=True)
        n = sum(lens)
        check_2x(f"out {lens}", out[:n], ref, pt)
