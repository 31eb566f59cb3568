"""fa_b200_fwd_splitkv_noalloc (include/fa_b200.h): split-KV decode + combine with the partials in the library's static
__device__ array -- an entry point with no counterpart in the reference, for callers that cannot allocate device memory.

Checked through the C ABI against the existing pair fa_b200_fwd_splitkv + fa_b200_fwd_combine (caller-owned partials) on
the same inputs, on TWO DIFFERENT problems back to back (the second, smaller call finds the first call's partials in the
array and must not be affected by them), and against the fp32 oracle with the decode tests' tolerance.  Collected last
(file name) so that a failure here cannot hide other tests under ``pytest -x``; passes alone in a fresh process too."""
import pytest
import torch

from oracle import attention_oracle as orc
from tests._util import check_2x, check_lse

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _params(capi, q, kc, vc, out, lse, lens, num_splits, causal):
    b, sq, h, d = q.shape
    sk, hk = kc.shape[1], kc.shape[2]
    p = capi.new_params(q.dtype)
    p.q, p.k, p.v, p.o = [capi.tdesc(t) for t in (q, kc, vc, out)]
    p.softmax_lse = lse.data_ptr()
    p.b, p.seqlen_q, p.seqlen_k, p.h, p.h_k, p.d = b, sq, sk, h, hk, d
    p.seqlen_q_rounded, p.seqlen_k_rounded, p.d_rounded = capi.round_up(sq, 128), capi.round_up(sk, 128), capi.round_up(d, 32)
    p.softmax_scale = d ** -0.5
    p.is_causal = 1 if causal else 0
    p.causal_bottom_right = 1
    p.lse_empty_is_neg_inf = 1
    p.seqused_k = lens.data_ptr()
    p.num_splits = num_splits
    return p


def _run_both(capi, b, sq, h, hk, d, sk, num_splits, causal, seed):
    lib = capi.load()
    torch.manual_seed(seed)
    q = torch.randn(b, sq, h, d, device=DEV, dtype=torch.bfloat16)
    kc = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    vc = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    lens = torch.randint(sk // 3, sk + 1, (b,), device=DEV, dtype=torch.int32)
    lens[0] = sk
    stream = capi.stream_ptr(q.device)
    # existing entry points, partials owned by the caller
    out_a = torch.full_like(q, float("nan"))
    lse_a = torch.full((b, h, sq), float("nan"), device=DEV, dtype=torch.float32)
    oacc = torch.empty(num_splits, b, sq, h, d, device=DEV, dtype=torch.float32)
    lacc = torch.empty(num_splits, b, h, sq, device=DEV, dtype=torch.float32)
    pa = _params(capi, q, kc, vc, out_a, lse_a, lens, num_splits, causal)
    pa.oaccum, pa.lseaccum = oacc.data_ptr(), lacc.data_ptr()
    capi.check(lib.fa_b200_fwd_splitkv(pa, stream), "fa_b200_fwd_splitkv")
    capi.check(lib.fa_b200_fwd_combine(pa, stream), "fa_b200_fwd_combine")
    # the new entry point: no partials passed
    out_b = torch.full_like(q, float("nan"))
    lse_b = torch.full((b, h, sq), float("nan"), device=DEV, dtype=torch.float32)
    pb = _params(capi, q, kc, vc, out_b, lse_b, lens, num_splits, causal)
    capi.check(lib.fa_b200_fwd_splitkv_noalloc(pb, stream), "fa_b200_fwd_splitkv_noalloc")
    torch.cuda.synchronize()
    assert not torch.isnan(out_b).any() and not torch.isnan(lse_b).any()
    # against the existing entry point: the tolerance the suite uses between two CUDA paths (tests/test_decode_capi_gpu.py)
    torch.testing.assert_close(out_b.float(), out_a.float(), atol=2e-2, rtol=2e-2)
    torch.testing.assert_close(lse_b, lse_a, atol=2e-3, rtol=1e-3)
    # and against the fp32 oracle under the suite's 2x rule (per batch element: its own cache length, bottom-right causal)
    for i in range(b):
        n = int(lens[i])
        mask = orc.causal_mask_bottom_right(sq, n).to(DEV) if causal else None
        ref, lse_ref = orc.attention_masked_ref(q[i:i + 1], kc[i:i + 1, :n], vc[i:i + 1, :n], mask, lse_empty=float("-inf"))
        pt, _ = orc.attention_masked_ref(q[i:i + 1], kc[i:i + 1, :n], vc[i:i + 1, :n], mask, upcast=False, reorder_ops=True)
        check_2x("out", out_b[i:i + 1], ref, pt)
        check_lse(lse_b[i:i + 1], lse_ref)


def test_noalloc_matches_allocating_path_twice_in_a_row(fa, capi):
    # call 1: decode shape of BASELINE config 5b scaled down (pack-GQA, TMA decode kernel), 4 splits
    _run_both(capi, b=4, sq=1, h=32, hk=8, d=128, sk=4096, num_splits=4, causal=True, seed=0)
    # call 2, right behind it on the same stream: a different, smaller problem over the same static array
    _run_both(capi, b=3, sq=2, h=8, hk=2, d=64, sk=1000, num_splits=3, causal=True, seed=1)
    # call 3: the first shape again with other data
    _run_both(capi, b=4, sq=1, h=32, hk=8, d=128, sk=4096, num_splits=4, causal=False, seed=2)


def test_noalloc_rejects_what_does_not_fit(fa, capi):
    lib = capi.load()
    b, sq, h, hk, d, sk = 64, 1, 32, 8, 128, 256
    q = torch.zeros(b, sq, h, d, device=DEV, dtype=torch.bfloat16)
    kc = torch.zeros(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    out = torch.empty_like(q)
    lse = torch.empty(b, h, sq, device=DEV, dtype=torch.float32)
    lens = torch.full((b,), sk, device=DEV, dtype=torch.int32)
    p = _params(capi, q, kc, kc, out, lse, lens, 64, True)      # 64 * 64 * 32 * 129 * 4 B = 67.6 MB > 16 MiB
    assert not lib.fa_b200_fwd_splitkv_noalloc(p, capi.stream_ptr(q.device))
    lib.flash_attn_error.restype = __import__("ctypes").c_char_p
    msg = lib.flash_attn_error().decode()
    assert "FA_B200_NOALLOC_SCRATCH_BYTES" in msg and "16777216" in msg, msg
    scratch = torch.empty(8, device=DEV, dtype=torch.float32)
    p = _params(capi, q, kc, kc, out, lse, lens, 2, True)
    p.oaccum = scratch.data_ptr()
    assert not lib.fa_b200_fwd_splitkv_noalloc(p, capi.stream_ptr(q.device))
