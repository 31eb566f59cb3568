"""GPU parity: the decode extras of flash_attn_with_kvcache -- paged KV (block_table),
cache_batch_idx, cache_leftpad, rotary on append, softcap, ALiBi -- against ``oracle.kvcache_ref``, on the TMA-fed
decode kernel (kind 3) and on the generic kernels (kind 1), plus softcap through the backward."""
import pytest
import torch

from oracle import attention_oracle as orc
from tests._util import check_2x, check_lse

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _ref_pair(q, kc, vc, **kw):
    o, l, kc2, vc2 = orc.kvcache_ref(q, kc, vc, **kw)
    op, _, _, _ = orc.kvcache_ref(q, kc, vc, upcast=False, **kw)
    return o, op, l, kc2, vc2


def _paged(kc, vc, page_size, extra_pages=3, seed=0):
    """Scatter a contiguous (b, sk, hk, d) cache into a shuffled page pool; sk must be a multiple of page_size."""
    b, sk, hk, d = kc.shape
    per = sk // page_size
    n = b * per + extra_pages
    g = torch.Generator().manual_seed(seed)
    perm = torch.randperm(n, generator=g)[: b * per].to(kc.device)
    pool_k = torch.full((n, page_size, hk, d), float("nan"), device=kc.device, dtype=kc.dtype)
    pool_v = torch.full((n, page_size, hk, d), float("nan"), device=kc.device, dtype=kc.dtype)
    pool_k[perm] = kc.reshape(b * per, page_size, hk, d)
    pool_v[perm] = vc.reshape(b * per, page_size, hk, d)
    return pool_k, pool_v, perm.view(b, per).to(torch.int32)


@pytest.mark.parametrize("num_splits", [0, 1, 3])
@pytest.mark.parametrize("page_size,kind", [(64, 3), (256, 3), (16, 1)])
@pytest.mark.parametrize("sq", [1, 3])
def test_paged_kv(fa, capi, sq, page_size, kind, num_splits):
    torch.manual_seed(0)
    b, sk, h, hk, d = 3, 1024, 8, 2, 128
    q = torch.randn(b, sq, h, d, device=DEV, dtype=torch.bfloat16)
    kc = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    vc = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    lens = torch.tensor([1024, 333, 65], device=DEV, dtype=torch.int32)
    for i in range(b):
        kc[i, int(lens[i]):] = float("nan")
        vc[i, int(lens[i]):] = float("nan")
    pool_k, pool_v, table = _paged(kc, vc, page_size)
    out, lse = fa.flash_attn_with_kvcache(q, pool_k, pool_v, cache_seqlens=lens, block_table=table, causal=True,
                                          num_splits=num_splits, return_softmax_lse=True)
    if num_splits == 1:
        assert capi.last_kernel_kind() == kind
    ref, pt, lse_ref, _, _ = _ref_pair(q, pool_k, pool_v, cache_seqlens=lens, block_table=table, causal=True)
    check_2x("out", out, ref, pt)
    check_lse(lse, lse_ref)


def test_paged_append_crosses_pages(fa):
    torch.manual_seed(1)
    b, h, hk, d, page_size, per = 2, 8, 2, 64, 64, 4
    q = torch.randn(b, 1, h, d, device=DEV, dtype=torch.float16)
    kc = torch.randn(b, per * page_size, hk, d, device=DEV, dtype=torch.float16)
    vc = torch.randn_like(kc)
    pool_k, pool_v, table = _paged(kc, vc, page_size)
    pool_k, pool_v = torch.nan_to_num(pool_k), torch.nan_to_num(pool_v)
    lens = torch.tensor([62, 127], device=DEV, dtype=torch.int32)     # 5 new rows straddle a page boundary
    kn = torch.randn(b, 5, hk, d, device=DEV, dtype=torch.float16)
    vn = torch.randn_like(kn)
    ref, pt, lse_ref, kref, vref = _ref_pair(q, pool_k, pool_v, k=kn, v=vn, cache_seqlens=lens, block_table=table,
                                              causal=True)
    out, lse = fa.flash_attn_with_kvcache(q, pool_k, pool_v, k=kn, v=vn, cache_seqlens=lens, block_table=table,
                                          causal=True, return_softmax_lse=True)
    assert torch.equal(pool_k, kref) and torch.equal(pool_v, vref), "paged append must be byte-exact"
    check_2x("out", out, ref, pt)
    check_lse(lse, lse_ref)


@pytest.mark.parametrize("sq,kind", [(1, 3), (20, 3), (40, 1)])
def test_cache_batch_idx(fa, capi, sq, kind):
    torch.manual_seed(2)
    b, bk, sk, h, hk, d = 3, 7, 512, 16, 4, 128
    q = torch.randn(b, sq, h, d, device=DEV, dtype=torch.bfloat16)
    kc = torch.randn(bk, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    vc = torch.randn_like(kc)
    idx = torch.tensor([5, 0, 6], device=DEV, dtype=torch.int32)
    lens = torch.tensor([512, 100, 257], device=DEV, dtype=torch.int32)
    kn = torch.randn(b, 2, hk, d, device=DEV, dtype=torch.bfloat16)
    vn = torch.randn_like(kn)
    lens_a = lens - 2
    ref, pt, lse_ref, kref, vref = _ref_pair(q, kc, vc, k=kn, v=vn, cache_seqlens=lens_a, cache_batch_idx=idx, causal=True)
    out, lse = fa.flash_attn_with_kvcache(q, kc, vc, k=kn, v=vn, cache_seqlens=lens_a, cache_batch_idx=idx, causal=True,
                                          num_splits=1, return_softmax_lse=True)
    assert capi.last_kernel_kind() == kind
    assert torch.equal(kc, kref) and torch.equal(vc, vref), "append through cache_batch_idx must be byte-exact"
    check_2x("out", out, ref, pt)
    check_lse(lse, lse_ref)


@pytest.mark.parametrize("num_splits", [1, 4])
@pytest.mark.parametrize("sq,kind", [(1, 3), (2, 3), (20, 3), (70, 1)])
@pytest.mark.parametrize("window", [(-1, -1), (100, 0)])
def test_cache_leftpad(fa, capi, sq, kind, window, num_splits):
    torch.manual_seed(3)
    b, sk, h, hk, d = 4, 700, 8, 2, 128
    q = torch.randn(b, sq, h, d, device=DEV, dtype=torch.bfloat16)
    kc = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    vc = torch.randn_like(kc)
    lp = torch.tensor([0, 1, 130, 64], device=DEV, dtype=torch.int32)
    lens = torch.tensor([700, 300, 131 + sq, 640], device=DEV, dtype=torch.int32)
    for i in range(b):   # NaN outside the valid window on both sides
        kc[i, :int(lp[i])] = float("nan"); vc[i, :int(lp[i])] = float("nan")
        kc[i, int(lens[i]):] = float("nan"); vc[i, int(lens[i]):] = float("nan")
    out, lse = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, cache_leftpad=lp, causal=True,
                                          window_size=window, num_splits=num_splits, return_softmax_lse=True)
    if num_splits == 1:
        assert capi.last_kernel_kind() == kind
    ref, pt, lse_ref, _, _ = _ref_pair(q, kc, vc, cache_seqlens=lens, cache_leftpad=lp, causal=True, window_size=window)
    check_2x("out", out, ref, pt)
    check_lse(lse, lse_ref)


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("interleaved", [False, True])
@pytest.mark.parametrize("rdim", [32, 128])
def test_rotary_append(fa, interleaved, causal, rdim):
    torch.manual_seed(4)
    b, sq, sk, h, hk, d = 3, 2, 384, 8, 2, 128
    dt = torch.bfloat16
    q = torch.randn(b, sq, h, d, device=DEV, dtype=dt)
    kc = torch.randn(b, sk, hk, d, device=DEV, dtype=dt)
    vc = torch.randn_like(kc)
    kn = torch.randn(b, sq, hk, d, device=DEV, dtype=dt)
    vn = torch.randn_like(kn)
    lens = torch.tensor([0, 100, 382], device=DEV, dtype=torch.int32)
    angle = torch.rand(sk, rdim // 2, device=DEV) * 6.28
    cos, sin = torch.cos(angle).to(dt), torch.sin(angle).to(dt)
    kw = dict(k=kn, v=vn, rotary_cos=cos, rotary_sin=sin, cache_seqlens=lens, causal=causal,
              rotary_interleaved=interleaved)
    ref, pt, lse_ref, kref, vref = _ref_pair(q, kc, vc, **kw)
    out, lse = fa.flash_attn_with_kvcache(q, kc, vc, return_softmax_lse=True, **kw)
    assert torch.equal(vc, vref)
    # rotated keys: fp32 arithmetic rounded once on both sides; an FMA contraction may move one bf16 ulp
    torch.testing.assert_close(kc.float(), kref.float(), atol=2e-2, rtol=8e-3)
    assert (kc != kref).float().mean().item() < 0.01
    check_2x("out", out, ref, pt, atol=2e-3)
    check_lse(lse, lse_ref, atol=5e-3)


def test_rotary_kernel_matches_oracle(capi):
    """fa_b200_rotary called directly: q-style fixed position, k-style running positions, rotary_dim < d tail copy."""
    torch.manual_seed(5)
    lib = capi.load()
    for dt, bf in ((torch.float16, 0), (torch.bfloat16, 1)):
        for interleaved in (0, 1):
            for fixed in (0, 1):
                b, s, h, d, rdim, ro = 2, 5, 3, 64, 32, 40
                x = torch.randn(b, s, h, d, device=DEV, dtype=dt)
                ang = torch.rand(ro, rdim // 2, device=DEV) * 6.28
                cos, sin = torch.cos(ang).to(dt), torch.sin(ang).to(dt)
                off = torch.tensor([3, 30], device=DEV, dtype=torch.int32)
                y = torch.empty_like(x)
                ok = lib.fa_b200_rotary(x.data_ptr(), y.data_ptr(), cos.data_ptr(), sin.data_ptr(), off.data_ptr(), b, s, h,
                                        d, rdim, ro, interleaved, fixed, bf, *x.stride()[:3], *y.stride()[:3],
                                        capi.stream_ptr(x.device))
                assert ok, capi.last_error()
                pos = off.long().view(b, 1) + (0 if fixed else 1) * torch.arange(s, device=DEV).view(1, s)
                ref = orc.rotary_ref(x, cos, sin, pos.expand(b, s), bool(interleaved))
                torch.testing.assert_close(y.float(), ref.float(), atol=2e-2, rtol=8e-3)
                assert torch.equal(y[..., rdim:], x[..., rdim:])


@pytest.mark.parametrize("sq,kind", [(1, 3), (20, 3), (70, 1)])
@pytest.mark.parametrize("softcap", [0.0, 15.0])
@pytest.mark.parametrize("alibi", [None, "h", "bh"])
def test_softcap_alibi(fa, capi, sq, kind, softcap, alibi):
    if softcap == 0.0 and alibi is None:
        pytest.skip("covered elsewhere")
    torch.manual_seed(6)
    b, sk, h, hk, d = 2, 777, 8, 2, 128
    q = torch.randn(b, sq, h, d, device=DEV, dtype=torch.bfloat16) * 3
    kc = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    vc = torch.randn_like(kc)
    lens = torch.tensor([777, 300], device=DEV, dtype=torch.int32)
    slopes = None
    if alibi == "h":
        slopes = torch.rand(h, device=DEV) * 0.3
    elif alibi == "bh":
        slopes = torch.rand(b, h, device=DEV) * 0.3
    kw = dict(cache_seqlens=lens, causal=True, softcap=softcap, alibi_slopes=slopes)
    out, lse = fa.flash_attn_with_kvcache(q, kc, vc, num_splits=1, return_softmax_lse=True, **kw)
    # many query rows: softcap alone has a tcgen05 instantiation, ALiBi needs the generic kernels
    assert capi.last_kernel_kind() == (2 if (kind == 1 and alibi is None) else kind)
    ref, pt, lse_ref, _, _ = _ref_pair(q, kc, vc, **kw)
    # tanh.approx (decode kernel) is good to ~1e-3 relative; the 2x rule still holds with a small absolute slack
    check_2x("out", out, ref, pt, atol=3e-3 if softcap > 0 else 1e-4)
    check_lse(lse, lse_ref, atol=2e-2 if softcap > 0 else 2e-3)
    out3, lse3 = fa.flash_attn_with_kvcache(q, kc, vc, num_splits=3, return_softmax_lse=True, **kw)
    check_2x("out(split)", out3, ref, pt, atol=3e-3 if softcap > 0 else 1e-4)


def test_decode_step_in_cuda_graph(fa):
    """A decode step (append + split-KV attention + combine) captured ONCE into a CUDA graph and replayed: all state
    that changes between steps (cache_seqlens, the new token's q/k/v) lives in static device buffers, the library
    makes no host-side decision that depends on them, and its only allocation is stream-ordered."""
    torch.manual_seed(8)
    b, sk, h, hk, d = 4, 1024, 16, 4, 128
    dt = torch.bfloat16
    kc = torch.randn(b, sk, hk, d, device=DEV, dtype=dt)
    vc = torch.randn_like(kc)
    kc_ref, vc_ref = kc.clone(), vc.clone()
    lens = torch.tensor([100, 1, 640, 1000], device=DEV, dtype=torch.int32)
    lens_ref = lens.clone()
    q_s = torch.zeros(b, 1, h, d, device=DEV, dtype=dt)
    kn_s, vn_s = torch.zeros(b, 1, hk, d, device=DEV, dtype=dt), torch.zeros(b, 1, hk, d, device=DEV, dtype=dt)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):   # warm-up outside the capture (library load, attribute setting)
        fa.flash_attn_with_kvcache(q_s, kc.clone(), vc.clone(), k=kn_s, v=vn_s, cache_seqlens=lens.clone(), causal=True,
                                   num_splits=4)
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        out_s = fa.flash_attn_with_kvcache(q_s, kc, vc, k=kn_s, v=vn_s, cache_seqlens=lens, causal=True, num_splits=4)
        lens.add_(1)
    for step in range(3):
        qn, kn, vn = torch.randn_like(q_s), torch.randn_like(kn_s), torch.randn_like(vn_s)
        q_s.copy_(qn); kn_s.copy_(kn); vn_s.copy_(vn)
        graph.replay()
        ref, _, kc_ref, vc_ref = orc.kvcache_ref(qn, kc_ref, vc_ref, k=kn, v=vn, cache_seqlens=lens_ref, causal=True)
        pt, _, _, _ = orc.kvcache_ref(qn, kc_ref, vc_ref, cache_seqlens=lens_ref + 1, causal=True, upcast=False)
        lens_ref = lens_ref + 1
        check_2x(f"out[step {step}]", out_s, ref, pt)
    assert torch.equal(lens, lens_ref) and torch.equal(kc, kc_ref) and torch.equal(vc, vc_ref)
