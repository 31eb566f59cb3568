"""GPU parity: flash_attn_with_kvcache (split-KV + LSE combine + append), the flat FA2 C symbols called directly
through ctypes (include/fa_capi.h), the FlashMask block scan and the ring LSE merge."""
import ctypes as C

import pytest
import torch

from oracle import attention_oracle as orc
from tests._util import check_2x, check_lse

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _kvcache_ref(q, kc, vc, seqlens, causal, window=(-1, -1)):
    outs, lses = [], []
    for i in range(q.shape[0]):
        n = int(seqlens[i])
        sq = q.shape[1]
        mask = None
        if causal:
            mask = orc.causal_mask_bottom_right(sq, n).to(DEV)
        if window != (-1, -1):
            m2 = orc.local_mask(sq, n, *window).to(DEV)
            mask = m2 if mask is None else (mask | m2)
        o, l = orc.attention_masked_ref(q[i:i + 1], kc[i:i + 1, :n], vc[i:i + 1, :n], mask, lse_empty=float("-inf"))
        op, _ = orc.attention_masked_ref(q[i:i + 1], kc[i:i + 1, :n], vc[i:i + 1, :n], mask, upcast=False,
                                         reorder_ops=True)
        outs.append((o, op)); lses.append(l)
    return (torch.cat([o[0] for o in outs]), torch.cat([o[1] for o in outs])), torch.cat(lses)


@pytest.mark.parametrize("num_splits", [1, 0, 3, 7])
@pytest.mark.parametrize("sq", [1, 4])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", [64, 128])
def test_kvcache_decode(fa, d, causal, sq, num_splits):
    torch.manual_seed(0)
    b, sk, h, hk = 3, 1500, 8, 2
    q = torch.randn(b, sq, h, d, device=DEV, dtype=torch.bfloat16)
    kc = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    vc = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    seqlens = torch.tensor([1500, 700, 33], device=DEV, dtype=torch.int32)
    out, lse = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=seqlens, causal=causal, num_splits=num_splits,
                                          return_softmax_lse=True)
    (ref, pt), lse_ref = _kvcache_ref(q, kc, vc, seqlens, causal)
    check_2x("out", out, ref, pt)
    check_lse(lse, lse_ref)


def test_kvcache_append(fa):
    torch.manual_seed(0)
    b, sk, h, hk, d = 2, 512, 4, 4, 128
    q = torch.randn(b, 1, h, d, device=DEV, dtype=torch.float16)
    kc = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.float16)
    vc = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.float16)
    kn = torch.randn(b, 1, hk, d, device=DEV, dtype=torch.float16)
    vn = torch.randn(b, 1, hk, d, device=DEV, dtype=torch.float16)
    seqlens = torch.tensor([100, 511], device=DEV, dtype=torch.int32)
    kc_ref, vc_ref = kc.clone(), vc.clone()
    for i in range(b):
        kc_ref[i, seqlens[i]] = kn[i, 0]; vc_ref[i, seqlens[i]] = vn[i, 0]
    out = fa.flash_attn_with_kvcache(q, kc, vc, k=kn, v=vn, cache_seqlens=seqlens, causal=True)
    assert torch.equal(kc, kc_ref) and torch.equal(vc, vc_ref)      # byte-exact append
    (ref, pt), _ = _kvcache_ref(q, kc_ref, vc_ref, seqlens + 1, True)
    check_2x("out", out, ref, pt)


def test_num_splits_heuristic(capi):
    lib = capi.load()
    # config C5b: b64, 8 kv-heads -> 512 m-blocks >= 0.8 * 148 SMs -> no split (csrc/flashmask_v2/heuristics.h:29-38)
    assert lib.fa_b200_num_splits(64, 8, 4, 32768, 128, 148) == 1
    # small batch, long context -> split
    assert lib.fa_b200_num_splits(1, 8, 4, 32768, 128, 148) > 1
    assert lib.fa_b200_num_splits(1, 8, 4, 512, 128, 148) == 1      # <= 4 n-blocks never split


def test_flat_capi_fwd_bwd(capi):
    """Call flash_attn_fwd / flash_attn_bwd exactly as a dlopen caller would (csrc/capi/flash_attn.h:11-55,100-161):
    expanded (per q-head) dk/dv, caller-summed; top-left causal; +inf lse convention."""
    lib = capi.load()
    torch.manual_seed(0)
    b, sq, sk, h, hk, d = 2, 200, 264, 4, 2, 128
    q = torch.randn(b, sq, h, d, device=DEV, dtype=torch.bfloat16)
    k = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    v = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    g = torch.randn(b, sq, h, d, device=DEV, dtype=torch.bfloat16)
    out = torch.empty_like(q)
    sq_r, sk_r = capi.round_up(sq, 128), capi.round_up(sk, 128)
    lse = torch.empty(b, h, sq, device=DEV, dtype=torch.float32)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    vp = lambda t: C.c_void_p(t.data_ptr() if t is not None else None)
    i32 = C.c_int
    st = lambda t: (i32(t.stride(1)), i32(t.stride(2)), i32(t.stride(0)))  # row, head, batch
    (qr, qh, qb), (kr, kh, kb), (vr, vh, vb), (or_, oh, ob) = st(q), st(k), st(v), st(out)
    lib.flash_attn_fwd.restype = C.c_bool
    ok = lib.flash_attn_fwd(
        vp(q), vp(k), vp(v), None, vp(out), None, vp(lse), i32(b), i32(sq), i32(sk), i32(sq_r), i32(sk_r), i32(h),
        i32(hk), i32(d), i32(d), C.c_float(0.0), C.c_float(d ** -0.5), C.c_float(1.0), C.c_bool(True), C.c_bool(False),
        C.c_bool(True), stream, C.c_uint64(0), C.c_uint64(0), None, None, None, None, None, None, None, None,
        qr, kr, vr, qh, kh, vh, or_, oh, qb, kb, vb, ob)
    assert ok, capi.last_error()
    mask = orc.causal_mask_top_left(sq, sk).to(DEV)
    from tests._util import oracle_all
    ref = oracle_all(q, k, v, g, mask)
    check_2x("out", out, *ref["out"])
    check_lse(lse, ref["lse"])
    dq = torch.empty_like(q)
    dk = torch.empty(b, sk, h, d, device=DEV, dtype=torch.bfloat16)   # expanded heads
    dv = torch.empty(b, sk, h, d, device=DEV, dtype=torch.bfloat16)
    softmax_d = torch.empty(b, h, sq_r, device=DEV, dtype=torch.float32)
    dq_accum = torch.empty(b, h, sq_r, d, device=DEV, dtype=torch.float32)
    (dqr, dqh, dqb), (dkr, dkh, dkb), (dvr, dvh, dvb), (gr, gh, gb) = st(dq), st(dk), st(dv), st(g)
    lib.flash_attn_bwd.restype = C.c_bool
    ok = lib.flash_attn_bwd(
        vp(g), vp(q), vp(k), vp(v), vp(out), vp(softmax_d), vp(lse), None, vp(dq), vp(dk), vp(dv), vp(dq_accum),
        i32(b), i32(sq), i32(sk), i32(sq_r), i32(sk_r), i32(h), i32(hk), i32(d), i32(d), C.c_float(0.0),
        C.c_float(d ** -0.5), C.c_float(1.0), C.c_bool(True), C.c_bool(True), i32(0), stream, C.c_uint64(0),
        C.c_uint64(0), None, None, None, None, None, None, None, None,
        qr, kr, vr, qh, kh, vh, or_, oh, qb, kb, vb, ob, dqr, dkr, dvr, dqh, dkh, dvh, gr, gh, dqb, dkb, dvb, gb)
    assert ok, capi.last_error()
    check_2x("dq", dq, *ref["dq"])
    g_ = h // hk
    check_2x("dk", dk.view(b, sk, hk, g_, d).float().sum(3), *ref["dk"], atol=2e-2)
    check_2x("dv", dv.view(b, sk, hk, g_, d).float().sum(3), *ref["dv"], atol=2e-2)


def test_flat_capi_errors(capi):
    lib = capi.load()
    lib.flash_attn_fwd_with_bias_and_mask.restype = C.c_bool
    lib.flash_attn_set_error(b"")
    q = torch.zeros(1, 8, 1, 12, device=DEV, dtype=torch.bfloat16)   # head_size 12 is not a multiple of 8
    lse = torch.zeros(1, 1, 8, device=DEV)
    vp = lambda t: C.c_void_p(t.data_ptr())
    i32 = C.c_int
    lib.flash_attn_fwd.restype = C.c_bool
    ok = lib.flash_attn_fwd(
        vp(q), vp(q), vp(q), None, vp(q), None, vp(lse), i32(1), i32(8), i32(8), i32(128), i32(128), i32(1), i32(1),
        i32(12), i32(32), C.c_float(0.0), C.c_float(1.0), C.c_float(1.0), C.c_bool(False), C.c_bool(False),
        C.c_bool(True), None, C.c_uint64(0), C.c_uint64(0), None, None, None, None, None, None, None, None,
        *[i32(12)] * 3, *[i32(12)] * 3, i32(12), i32(12), *[i32(96)] * 4)
    assert not ok and "head_size" in capi.last_error()


def test_flashmask_block_scan(capi):
    """Block min/max scan vs torch (csrc/flashmask_v2/flash_mask.hpp:162-245 ordering and padding)."""
    lib = capi.load()
    gen = torch.Generator().manual_seed(1)
    b, hm, sk, bn = 2, 3, 1000, 128
    vecs = [torch.randint(0, 1000, (b, hm, sk), generator=gen, dtype=torch.int32).to(DEV) for _ in range(4)]
    lts, lte, uts, ute = vecs
    nblock = (sk + bn - 1) // bn
    npad = (nblock + 3) & ~3
    ws = torch.full((8, b * hm, npad), -7, dtype=torch.int32, device=DEV)
    p = capi.new_params(torch.bfloat16)
    p.b, p.h, p.h_mask, p.seqlen_k = b, hm, hm, sk
    p.lt_start, p.lt_end, p.ut_start, p.ut_end = [t.data_ptr() for t in (lts, lte, uts, ute)]
    p.flashmask_maxmin = ws.data_ptr()
    assert lib.fa_b200_flashmask_prepare(p, bn, torch.cuda.current_stream().cuda_stream), capi.last_error()
    torch.cuda.synchronize()
    for idx, t in enumerate((lts, ute, lte, uts)):       # order: lts, ute, lte, uts (max then min)
        padded = torch.nn.functional.pad(t.view(b * hm, sk), (0, nblock * bn - sk), value=-1)
        blk = padded.view(b * hm, nblock, bn)
        mx = torch.where(blk >= 0, blk, torch.zeros_like(blk)).amax(-1)
        mn = torch.where(blk >= 0, blk, torch.full_like(blk, 2 ** 31 - 1)).amin(-1)
        assert torch.equal(ws[2 * idx, :, :nblock], mx)
        assert torch.equal(ws[2 * idx + 1, :, :nblock], mn)
        assert (ws[2 * idx, :, nblock:] == 0).all() and (ws[2 * idx + 1, :, nblock:] == 2 ** 31 - 1).all()


def test_lse_merge(capi):
    lib = capi.load()
    torch.manual_seed(0)
    rows, h, d = 300, 4, 128
    q = torch.randn(1, rows, h, d, device=DEV); k = torch.randn(1, 400, h, d, device=DEV)
    v = torch.randn(1, 400, h, d, device=DEV)
    full, lse_full = orc.attention_masked_ref(q, k, v)
    o1, l1 = orc.attention_masked_ref(q, k[:, :150], v[:, :150])
    o2, l2 = orc.attention_masked_ref(q, k[:, 150:], v[:, 150:])
    l1 = l1[0].t().contiguous(); l2 = l2[0].t().contiguous()        # (rows, h)
    l1[5] = float("inf"); o1[0, 5] = 0                               # an "empty" partial on the FA2 surface
    acc = o1[0].clone().contiguous()
    step = o2[0].to(torch.bfloat16).contiguous()
    assert lib.fa_b200_lse_merge(acc.data_ptr(), l1.data_ptr(), step.data_ptr(), l2.data_ptr(), rows, h, d, 1, 1,
                                 h, 1, h, 1, torch.cuda.current_stream().cuda_stream), capi.last_error()
    torch.cuda.synchronize()
    exp = full[0].clone(); exp[5] = o2[0, 5]
    torch.testing.assert_close(acc, exp, atol=2e-2, rtol=2e-2)
    exp_l = lse_full[0].t().clone(); exp_l[5] = l2[5]
    torch.testing.assert_close(l1, exp_l, atol=1e-4, rtol=1e-4)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "fp16"])
@pytest.mark.parametrize("h,hk,sq", [(32, 8, 1), (16, 2, 2), (8, 8, 3), (16, 1, 1), (32, 2, 1)])
@pytest.mark.parametrize("num_splits", [0, 1, 5])
def test_decode_kernel_shapes(fa, capi, h, hk, sq, num_splits, dtype):
    """The TMA-fed decode kernel (csrc/decode.cu): pack-GQA group sizes 1..16, NaN garbage beyond cache_seqlens,
    lengths that are not multiples of the 64-key tile; h=32/hk=2 (group 16, sq 1) fills all 16 packed rows."""
    torch.manual_seed(1)
    b, sk, d = 4, 2111, 128
    q = torch.randn(b, sq, h, d, device=DEV, dtype=dtype)
    kc = torch.randn(b, sk, hk, d, device=DEV, dtype=dtype)
    vc = torch.randn(b, sk, hk, d, device=DEV, dtype=dtype)
    seqlens = torch.tensor([2111, 1, 64, 1000], device=DEV, dtype=torch.int32)
    for i in range(b):  # rows beyond the used length hold NaN: they must not leak into the result
        kc[i, int(seqlens[i]):] = float("nan")
        vc[i, int(seqlens[i]):] = float("nan")
    out, lse = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=seqlens, causal=True, num_splits=num_splits,
                                          return_softmax_lse=True)
    assert capi.last_kernel_kind() in (1, 3)
    (ref, pt), lse_ref = _kvcache_ref(q, kc, vc, seqlens, True)
    check_2x("out", out, ref, pt)
    check_lse(lse, lse_ref)


def test_decode_kernel_is_used(fa, capi):
    torch.manual_seed(0)
    q = torch.randn(2, 1, 32, 128, device=DEV, dtype=torch.bfloat16)
    kc = torch.randn(2, 4096, 8, 128, device=DEV, dtype=torch.bfloat16)
    vc = torch.randn(2, 4096, 8, 128, device=DEV, dtype=torch.bfloat16)
    capi.launch_count(reset=True)
    fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=4096, causal=True, num_splits=1)
    assert capi.launch_count() == 1
    # the split launch is recorded as kind 3 (TMA-fed decode kernel); the combine kernel follows it
    lib = capi.load()
    p_kind = []
    fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=4096, causal=True, num_splits=1)
    p_kind.append(capi.last_kernel_kind())
    assert p_kind == [3]
    # more than 16 packed rows: several 16-row tiles per (batch, kv-head), still the decode kernel
    q2 = torch.randn(2, 5, 32, 128, device=DEV, dtype=torch.bfloat16)
    fa.flash_attn_with_kvcache(q2, kc, vc, cache_seqlens=4096, causal=True, num_splits=1)
    assert capi.last_kernel_kind() == 3
    # more than 128 packed rows (chunked prefill over the cache): one pass of the tcgen05 forward
    q3 = torch.randn(2, 64, 32, 128, device=DEV, dtype=torch.bfloat16)
    fa.flash_attn_with_kvcache(q3, kc, vc, cache_seqlens=4096, causal=True)
    assert capi.last_kernel_kind() == 2
    # an explicit split of such a shape is honoured by the generic kernels
    fa.flash_attn_with_kvcache(q3, kc, vc, cache_seqlens=4096, causal=True, num_splits=2)
    assert capi.last_kernel_kind() == 1


@pytest.mark.parametrize("window", [(128, 0), (300, -1), (-1, 0)])
@pytest.mark.parametrize("sq", [1, 4])
def test_decode_window(fa, window, sq):
    torch.manual_seed(2)
    b, sk, h, hk, d = 2, 1111, 8, 2, 64
    q = torch.randn(b, sq, h, d, device=DEV, dtype=torch.bfloat16)
    kc = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    vc = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    seqlens = torch.tensor([1111, 500], device=DEV, dtype=torch.int32)
    out, lse = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=seqlens, window_size=window, num_splits=3,
                                          return_softmax_lse=True)
    (ref, pt), lse_ref = _kvcache_ref(q, kc, vc, seqlens, False, window)
    check_2x("out", out, ref, pt)
    check_lse(lse, lse_ref)


@pytest.mark.parametrize("sk", [264, 261], ids=["sk264", "sk261_unaligned_rows"])
@pytest.mark.parametrize("mh,ms", [(1, 1), (4, 1), (1, 200), (4, 200)])
@pytest.mark.parametrize("causal", [False, True])
def test_flat_capi_dense_attn_mask(capi, mh, ms, causal, sk):
    """flash_attn_fwd / flash_attn_bwd with the dense additive attn_mask of the flat ABI (csrc/capi/flash_attn.h:36-37):
    logits = scale * q.k + mask, mask (b, 1-or-h, 1-or-sq, sk) broadcast by h % mod / row % mod."""
    lib = capi.load()
    torch.manual_seed(0)
    b, sq, h, d = 2, 200, 4, 64
    q, g = [torch.randn(b, sq, h, d, device=DEV, dtype=torch.float16) for _ in range(2)]
    k, v = [torch.randn(b, sk, h, d, device=DEV, dtype=torch.float16) for _ in range(2)]
    bias = (torch.randn(b, mh, ms, sk, device=DEV) * 2).to(torch.float16)
    bias[:, :, :, 250:] = float("-inf")                       # a padded key tail expressed through the mask
    out, dq, dk, dv = torch.empty_like(q), torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
    sq_r, sk_r = capi.round_up(sq, 128), capi.round_up(sk, 128)
    lse = torch.empty(b, h, sq, device=DEV, dtype=torch.float32)
    softmax_d = torch.empty(b, h, sq_r, device=DEV, dtype=torch.float32)
    dq_accum = torch.empty(b, h, sq_r, d, device=DEV, dtype=torch.float32)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    vp = lambda t: C.c_void_p(t.data_ptr() if t is not None else None)
    i32 = C.c_int
    st = lambda t: (i32(t.stride(1)), i32(t.stride(2)), i32(t.stride(0)))  # row, head, batch
    (qr, qh, qb), (kr, kh, kb), (vr, vh, vb), (or_, oh, ob) = st(q), st(k), st(v), st(out)
    dims = (C.c_int64 * 4)(b, mh, ms, sk)
    scale = d ** -0.5
    lib.flash_attn_fwd.restype = C.c_bool
    ok = lib.flash_attn_fwd(
        vp(q), vp(k), vp(v), None, vp(out), None, vp(lse), i32(b), i32(sq), i32(sk), i32(sq_r), i32(sk_r), i32(h),
        i32(h), i32(d), i32(d), C.c_float(0.0), C.c_float(scale), C.c_float(1.0 / scale), C.c_bool(causal), C.c_bool(False),
        C.c_bool(False), stream, C.c_uint64(0), C.c_uint64(0), vp(bias), dims, None, None, None, None, None, None,
        qr, kr, vr, qh, kh, vh, or_, oh, qb, kb, vb, ob)
    assert ok, capi.last_error()
    assert capi.last_kernel_kind() == 2, "dense attn_mask forward did not take the tcgen05 kernel"
    (dqr, dqh, dqb), (dkr, dkh, dkb), (dvr, dvh, dvb), (gr, gh, gb) = st(dq), st(dk), st(dv), st(g)
    lib.flash_attn_bwd.restype = C.c_bool
    ok = lib.flash_attn_bwd(
        vp(g), vp(q), vp(k), vp(v), vp(out), vp(softmax_d), vp(lse), None, vp(dq), vp(dk), vp(dv), vp(dq_accum),
        i32(b), i32(sq), i32(sk), i32(sq_r), i32(sk_r), i32(h), i32(h), i32(d), i32(d), C.c_float(0.0),
        C.c_float(scale), C.c_float(1.0 / scale), C.c_bool(causal), C.c_bool(False), i32(0), stream, C.c_uint64(0),
        C.c_uint64(0), vp(bias), dims, None, None, None, None, None, None,
        qr, kr, vr, qh, kh, vh, or_, oh, qb, kb, vb, ob, dqr, dkr, dvr, dqh, dkh, dvh, gr, gh, dqb, dkb, dvb, gb)
    assert ok, capi.last_error()
    assert capi.last_kernel_kind() == 2, "dense attn_mask backward did not take the tcgen05 kernels"
    torch.cuda.synchronize()
    mask = orc.causal_mask_top_left(sq, sk).to(DEV) if causal else None
    bias_full = bias.float().expand(b, mh, sq if ms == 1 else ms, sk)
    ref = orc.attention_grads_ref(q, k, v, g, mask, bias=bias_full)
    pt = orc.attention_grads_ref(q, k, v, g, mask, upcast=False, reorder_ops=True, bias=bias_full)
    check_2x("out", out, ref[0], pt[0])
    check_lse(lse, ref[1])
    check_2x("dq", dq, ref[2], pt[2])
    check_2x("dk", dk, ref[3], pt[3])
    check_2x("dv", dv, ref[4], pt[4])


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("sq,h,hk", [(5, 32, 8), (8, 16, 1), (33, 8, 8), (100, 4, 4), (300, 8, 2)])
def test_kvcache_many_query_rows(fa, capi, sq, h, hk, causal):
    """Speculative decoding / chunked prefill over a cache: 17..128 packed rows run as several 16-row tiles of the decode
    kernel (kind 3), more than that as one tcgen05 forward pass with seqused_k and the bottom-right mask (kind 2)."""
    torch.manual_seed(2)
    b, sk, d = 3, 1500, 128
    q = torch.randn(b, sq, h, d, device=DEV, dtype=torch.bfloat16)
    kc = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    vc = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    seqlens = torch.tensor([1500, 301, 777], device=DEV, dtype=torch.int32)
    for i in range(b):
        kc[i, int(seqlens[i]):] = float("nan"); vc[i, int(seqlens[i]):] = float("nan")
    (ref, pt), lse_ref = _kvcache_ref(q, kc, vc, seqlens, causal)
    for ns in (1, 0, 3):
        out, lse = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=seqlens, causal=causal, num_splits=ns,
                                              return_softmax_lse=True)
        if ns == 1:   # (with splits the last launch is the combine kernel)
            assert capi.last_kernel_kind() == (3 if sq * (h // hk) <= 128 else 2)
        check_2x("out", out, ref, pt)
        check_lse(lse, lse_ref)
