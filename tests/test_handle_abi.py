"""The setter-style params-handle ABI (include/fa_handle.h; reference csrc/flashmask_v2/flash_api.h:14-266 and
csrc/flash_attn_v3/flash_api.h).  CPU: every declared symbol is exported and getters return what setters stored.
GPU: forward / backward / split-KV driven exactly the way Paddle drives the reference (create, set fields, run)."""
import ctypes as C
import importlib
import math
import os
import re
import subprocess

import pytest
import torch

from oracle import attention_oracle as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
capi = importlib.import_module("xhy-flash-attention_b200._capi")


def _declared():
    """Expand the X-macros of fa_handle.h with the C preprocessor and collect the function names."""
    src = subprocess.run(["gcc", "-E", "-P", "-I", os.path.join(ROOT, "include"), "-I", "/usr/local/cuda/include",
                          os.path.join(ROOT, "include", "fa_handle.h")], capture_output=True, text=True, check=True).stdout
    return sorted(set(re.findall(r"\b((?:flashmaskv2|fa3)_\w+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(capi.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return C.CDLL(capi.LIB_PATH)


def test_every_handle_symbol_is_exported(lib):
    names = _declared()
    # 13 lifecycle/run functions per prefix; 4 accessors per shared field, 2 per backward-only field
    assert len(names) > 900 and "flashmaskv2_run_mha_fwd" in names and "fa3_bwd_params_set_dq_accum_ptr" in names
    assert "flashmaskv2_fwd_params_set_lt_start_ptr" in names and "fa3_fwd_params_set_lt_start_ptr" not in names
    out = subprocess.run(["nm", "-D", "--defined-only", capi.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = {line.split()[-1] for line in out.splitlines() if line.strip()}
    missing = [n for n in names if n not in exported]
    assert not missing, missing[:10]


def test_setters_and_getters_roundtrip(lib):
    for prefix in ("flashmaskv2", "fa3"):
        create = getattr(lib, f"{prefix}_create_bwd_params_handle"); create.restype = C.c_void_p
        cast = getattr(lib, f"{prefix}_cast_to_fwd_params_handle"); cast.restype = C.c_void_p; cast.argtypes = [C.c_void_p]
        h = create()
        assert h
        for member, ctype, val in (("seqlen_q", C.c_int, 777), ("q_row_stride", C.c_int64, 1 << 40),
                                   ("scale_softmax", C.c_float, 0.125), ("is_causal", C.c_bool, True),
                                   ("do_head_stride", C.c_int64, 128)):
            s = getattr(lib, f"{prefix}_bwd_params_set_{member}"); s.argtypes = [C.c_void_p, ctype]; s.restype = None
            g = getattr(lib, f"{prefix}_bwd_params_get_{member}"); g.argtypes = [C.c_void_p]; g.restype = ctype
            s(h, val)
            assert g(h) == val
        # the forward view of a backward handle sees the same fields
        g = getattr(lib, f"{prefix}_fwd_params_get_seqlen_q"); g.argtypes = [C.c_void_p]; g.restype = C.c_int
        assert g(cast(h)) == 777
        clear = getattr(lib, f"{prefix}_clear_bwd_params_handle"); clear.argtypes = [C.c_void_p]
        clear(h)
        assert g(cast(h)) == 0
        destroy = getattr(lib, f"{prefix}_destroy_bwd_params_handle"); destroy.argtypes = [C.c_void_p]
        destroy(h)


# ---------------------------------------------------------------------------------------------------- GPU
_CT = {int: C.c_int, float: C.c_float, bool: C.c_bool}


class Handle:
    def __init__(self, lib, prefix, kind):
        self.lib, self.prefix, self.kind = lib, prefix, kind
        f = getattr(lib, f"{prefix}_create_{kind}_params_handle"); f.restype = C.c_void_p
        self.h = f()

    def set(self, **kw):
        for k, v in kw.items():
            f = getattr(self.lib, f"{self.prefix}_{self.kind}_params_set_{k}")
            if isinstance(v, torch.Tensor):
                f.argtypes = [C.c_void_p, C.c_void_p]; v = v.data_ptr()
            elif k.endswith("_stride"):
                f.argtypes = [C.c_void_p, C.c_int64]
            else:
                f.argtypes = [C.c_void_p, _CT[type(v)]]
            f.restype = None
            f(self.h, v)

    def call(self, name, *args, restype=None):
        f = getattr(self.lib, f"{self.prefix}_{name}")
        f.argtypes = [C.c_void_p] + [C.c_void_p if not isinstance(a, bool) else C.c_bool for a in args]
        f.restype = restype
        return f(self.h, *args)

    def close(self):
        f = getattr(self.lib, f"{self.prefix}_destroy_{self.kind}_params_handle"); f.argtypes = [C.c_void_p]
        f(self.h)


def _set_qkvo(hd, q, k, v, o, packed=False):
    for name, t in (("q", q), ("k", k), ("v", v), ("o", o)):
        strides = (0, t.stride(0), t.stride(1)) if packed else (t.stride(0), t.stride(1), t.stride(2))
        hd.set(**{f"{name}_ptr": t, f"{name}_batch_stride": strides[0], f"{name}_row_stride": strides[1],
                  f"{name}_head_stride": strides[2]})


@pytest.mark.gpu
@pytest.mark.parametrize("prefix,mask", [("flashmaskv2", "doc"), ("flashmaskv2", None), ("fa3", None)])
def test_handle_fwd_bwd(fa, prefix, mask):
    from tests._util import check_2x, check_lse, oracle_all
    lib = capi.load()
    torch.manual_seed(0)
    b, s, h, hk, d = 2, 640, 4, 2, 128
    dev = "cuda"
    q, g = [torch.randn(b, s, h, d, device=dev, dtype=torch.bfloat16) for _ in range(2)]
    k, v = [torch.randn(b, s, hk, d, device=dev, dtype=torch.bfloat16) for _ in range(2)]
    o, dq = torch.empty_like(q), torch.empty_like(q)
    dk, dv = torch.empty_like(k), torch.empty_like(v)
    lse = torch.empty(b, h, s, device=dev, dtype=torch.float32)
    stream = torch.cuda.current_stream().cuda_stream
    common = dict(b=b, seqlen_q=s, seqlen_k=s, h=h, h_k=hk, d=d, dv=d, scale_softmax=1.0 / math.sqrt(d), is_bf16=True,
                  is_causal=True, num_splits=1, arch=100, num_sm=148, softmax_lse_ptr=lse)
    dense = orc.causal_mask_bottom_right(s, s)
    keep = []
    if mask == "doc":
        lts = orc.document_lt_start([200, 300, 140]).view(1, 1, s).expand(b, 1, s).contiguous().to(dev)
        ws = torch.empty(8 * b * ((s + 63) // 64 + 3), dtype=torch.int32, device=dev)
        common.update(lt_start_ptr=lts, h_flashmask=1, h_h_flashmask_ratio=h, flashmask_maxmin_ptr=ws)
        keep += [lts, ws]
        dense = dense | orc.flashmask_dense(s, s, lts.cpu())
    hf = Handle(lib, prefix, "fwd")
    _set_qkvo(hf, q, k, v, o)
    hf.set(**common)
    hf.call("run_mha_fwd", stream)
    hb = Handle(lib, prefix, "bwd")
    _set_qkvo(hb, q, k, v, o)
    hb.set(**common)
    for name, t in (("do", g), ("dq", dq), ("dk", dk), ("dv", dv)):
        hb.set(**{f"{name}_ptr": t, f"{name}_batch_stride": t.stride(0), f"{name}_row_stride": t.stride(1),
                  f"{name}_head_stride": t.stride(2)})
    hb.call("run_mha_bwd", stream)
    torch.cuda.synchronize()
    ref = oracle_all(q, k, v, g, dense.to(dev), lse_empty=float("-inf"))
    check_2x("out", o, *ref["out"])
    check_lse(lse, ref["lse"])
    check_2x("dq", dq, *ref["dq"])
    check_2x("dk", dk, *ref["dk"])
    check_2x("dv", dv, *ref["dv"])
    hf.close(); hb.close()


@pytest.mark.gpu
def test_handle_varlen_lse_is_packed(fa):
    from tests._util import check_2x
    lib = capi.load()
    torch.manual_seed(1)
    lens = [100, 333, 64]
    total, h, d = sum(lens), 4, 128
    dev = "cuda"
    q, k, v = [torch.randn(total, h, d, device=dev, dtype=torch.bfloat16) for _ in range(3)]
    o = torch.empty_like(q)
    cu = torch.tensor([0, 100, 433, 497], device=dev, dtype=torch.int32)
    lse = torch.full((h, total), float("nan"), device=dev, dtype=torch.float32)
    hf = Handle(lib, "flashmaskv2", "fwd")
    _set_qkvo(hf, q, k, v, o, packed=True)
    hf.set(b=3, seqlen_q=max(lens), seqlen_k=max(lens), total_q=total, total_k=total, h=h, h_k=h, d=d, dv=d,
           scale_softmax=1.0 / math.sqrt(d), is_bf16=True, is_causal=True, num_splits=1, softmax_lse_ptr=lse,
           cu_seqlens_q=cu, cu_seqlens_k=cu)
    hf.call("run_mha_fwd", torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ref, lses = orc.attention_varlen_ref(q, k, v, cu.tolist(), cu.tolist(), causal=True)
    pt, _ = orc.attention_varlen_ref(q, k, v, cu.tolist(), cu.tolist(), causal=True, upcast=False, reorder_ops=True)
    check_2x("out", o, ref, pt)
    torch.testing.assert_close(lse, torch.cat(lses, dim=1), atol=2e-3, rtol=1e-3)   # (h, total_q)
    hf.close()


@pytest.mark.gpu
def test_handle_decode_split_and_combine(fa):
    from tests._util import check_2x
    lib = capi.load()
    torch.manual_seed(2)
    b, sk, h, hk, d, splits = 3, 2000, 8, 2, 128, 4
    dev = "cuda"
    q = torch.randn(b, 1, h, d, device=dev, dtype=torch.bfloat16)
    kc, vc = [torch.randn(b, sk, hk, d, device=dev, dtype=torch.bfloat16) for _ in range(2)]
    o = torch.empty_like(q)
    lse = torch.empty(b, h, 1, device=dev, dtype=torch.float32)
    used = torch.tensor([2000, 1234, 77], device=dev, dtype=torch.int32)
    oacc = torch.empty(splits, b, 1, h, d, device=dev, dtype=torch.float32)
    lacc = torch.empty(splits, b, h, 1, device=dev, dtype=torch.float32)
    hf = Handle(lib, "fa3", "fwd")
    _set_qkvo(hf, q, kc, vc, o)
    hf.set(b=b, seqlen_q=1, seqlen_k=sk, h=h, h_k=hk, d=d, dv=d, scale_softmax=1.0 / math.sqrt(d), is_bf16=True,
           is_causal=True, seqused_k=used, softmax_lse_ptr=lse, num_sm=148)
    assert hf.call("get_pack_gqa", restype=C.c_bool)
    assert hf.call("get_num_splits", restype=C.c_int) >= 1
    hf.set(num_splits=splits, oaccum_ptr=oacc, softmax_lseaccum_ptr=lacc)
    stream = torch.cuda.current_stream().cuda_stream
    hf.call("run_mha_fwd", stream)
    hf.call("run_mha_fwd_combine", stream, False)
    torch.cuda.synchronize()
    for i in range(b):
        n = int(used[i])
        ref, _ = orc.attention_masked_ref(q[i:i + 1], kc[i:i + 1, :n], vc[i:i + 1, :n])
        pt, _ = orc.attention_masked_ref(q[i:i + 1], kc[i:i + 1, :n], vc[i:i + 1, :n], upcast=False, reorder_ops=True)
        check_2x("out", o[i:i + 1], ref, pt)
    hf.close()


@pytest.mark.gpu
@pytest.mark.parametrize("s,h,hk,d,cap", [(200, 4, 2, 64, 8.0), (1024, 8, 2, 128, 30.0), (333, 4, 4, 96, 15.0)])
def test_handle_softcap_fwd_bwd(fa, s, h, hk, d, cap):
    """softcap through the fa3 handle (csrc/flash_attn_v3/flash_api.cu:42,235): forward and backward, dS carries 1 - tanh^2."""
    from tests._util import check_2x
    lib = capi.load()
    torch.manual_seed(3)
    b = 2
    dev = "cuda"
    q, g = [torch.randn(b, s, h, d, device=dev, dtype=torch.bfloat16) * 2 for _ in range(2)]
    k, v = [torch.randn(b, s, hk, d, device=dev, dtype=torch.bfloat16) for _ in range(2)]
    o, dq = torch.empty_like(q), torch.empty_like(q)
    dk, dv = torch.empty_like(k), torch.empty_like(v)
    lse = torch.empty(b, h, s, device=dev, dtype=torch.float32)
    stream = torch.cuda.current_stream().cuda_stream
    common = dict(b=b, seqlen_q=s, seqlen_k=s, h=h, h_k=hk, d=d, dv=d, scale_softmax=1.0 / math.sqrt(d), is_bf16=True,
                  is_causal=True, num_splits=1, softcap=cap, softmax_lse_ptr=lse)
    hf = Handle(lib, "fa3", "fwd")
    _set_qkvo(hf, q, k, v, o)
    hf.set(**common)
    hf.call("run_mha_fwd", stream)
    hb = Handle(lib, "fa3", "bwd")
    _set_qkvo(hb, q, k, v, o)
    hb.set(**common)
    for name, t in (("do", g), ("dq", dq), ("dk", dk), ("dv", dv)):
        hb.set(**{f"{name}_ptr": t, f"{name}_batch_stride": t.stride(0), f"{name}_row_stride": t.stride(1),
                  f"{name}_head_stride": t.stride(2)})
    hb.call("run_mha_bwd", stream)
    torch.cuda.synchronize()
    assert capi.last_kernel_kind() == 2, "softcap did not take the tcgen05 kernels"
    mask = orc.causal_mask_bottom_right(s, s).to(dev)
    ref = orc.attention_grads_ref(q, k, v, g, mask, softcap=cap)
    pt = orc.attention_grads_ref(q, k, v, g, mask, upcast=False, reorder_ops=True, softcap=cap)
    for name, got, i in (("out", o, 0), ("dq", dq, 2), ("dk", dk, 3), ("dv", dv, 4)):
        check_2x(name, got, ref[i], pt[i], atol=3e-3)       # tanh.approx / lg2.approx: ~1e-3 relative
    torch.testing.assert_close(lse, ref[1], atol=2e-2, rtol=1e-3)
    hf.close(); hb.close()


@pytest.mark.gpu
@pytest.mark.parametrize("rotary", [False, True])
def test_handle_paged_append_decode(fa, rotary):
    """page_table + knew / vnew (+ rotary) through the fa3 handle: append into the pages, then decode over them."""
    from tests._util import check_2x
    lib = capi.load()
    torch.manual_seed(4)
    b, h, hk, d, page, per = 2, 8, 2, 128, 64, 5
    dev, dt = "cuda", torch.bfloat16
    n_pages = b * per + 2
    pool_k = torch.randn(n_pages, page, hk, d, device=dev, dtype=dt)
    pool_v = torch.randn_like(pool_k)
    table = torch.randperm(n_pages)[: b * per].view(b, per).to(torch.int32).to(dev)
    q = torch.randn(b, 1, h, d, device=dev, dtype=dt)
    kn, vn = torch.randn(b, 1, hk, d, device=dev, dtype=dt), torch.randn(b, 1, hk, d, device=dev, dtype=dt)
    used = torch.tensor([255, 64], device=dev, dtype=torch.int32)
    o = torch.empty_like(q)
    lse = torch.empty(b, h, 1, device=dev, dtype=torch.float32)
    kw = {}
    if rotary:
        ang = torch.rand(per * page, 32, device=dev) * 6.28
        kw = dict(rotary_cos=torch.cos(ang).to(dt), rotary_sin=torch.sin(ang).to(dt), rotary_interleaved=False)
    ref, _, kref, vref = orc.kvcache_ref(q, pool_k, pool_v, k=kn, v=vn, cache_seqlens=used, block_table=table,
                                         causal=True, **kw)
    pt, _, _, _ = orc.kvcache_ref(q, pool_k, pool_v, k=kn, v=vn, cache_seqlens=used, block_table=table, causal=True,
                                  upcast=False, **kw)
    hf = Handle(lib, "fa3", "fwd")
    _set_qkvo(hf, q, pool_k, pool_v, o)
    hf.set(b=b, seqlen_q=1, seqlen_k=per * page, h=h, h_k=hk, d=d, dv=d, scale_softmax=1.0 / math.sqrt(d), is_bf16=True,
           is_causal=True, seqused_k=used, softmax_lse_ptr=lse, num_splits=1, page_table=table,
           page_table_batch_stride=table.stride(0), page_size=page, num_pages=n_pages, knew_ptr=kn, vnew_ptr=vn,
           seqlen_knew=1, knew_batch_stride=kn.stride(0), knew_row_stride=kn.stride(1), knew_head_stride=kn.stride(2),
           vnew_batch_stride=vn.stride(0), vnew_row_stride=vn.stride(1), vnew_head_stride=vn.stride(2))
    if rotary:
        hf.set(rotary_cos_ptr=kw["rotary_cos"], rotary_sin_ptr=kw["rotary_sin"], rotary_dim=64, is_rotary_interleaved=False)
    hf.call("run_mha_fwd", torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert torch.equal(pool_v, vref)
    torch.testing.assert_close(pool_k.float(), kref.float(), atol=2e-2, rtol=8e-3)
    check_2x("out", o, ref, pt, atol=2e-3)
    hf.close()


@pytest.mark.gpu
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("with_k", [False, True])
def test_handle_seqused_q(fa, causal, with_k):
    """seqused_q (+ seqused_k) on a padded batch (csrc/flashmask_v2/seqlen.h:17-91): element i attends its first
    seqused_q[i] rows over its first seqused_k[i] keys, bottom-right aligned on those lengths; the rest stays untouched."""
    from tests._util import check_2x, check_lse, oracle_all
    lib = capi.load()
    torch.manual_seed(0)
    b, s, h, hk, d = 3, 512, 4, 2, 128
    dev = "cuda"
    q, g = [torch.randn(b, s, h, d, device=dev, dtype=torch.bfloat16) for _ in range(2)]
    k, v = [torch.randn(b, s, hk, d, device=dev, dtype=torch.bfloat16) for _ in range(2)]
    used_q, used_k = [512, 130, 77], ([512, 300, 77] if with_k else [s] * b)
    uq = torch.tensor(used_q, device=dev, dtype=torch.int32)
    uk = torch.tensor(used_k, device=dev, dtype=torch.int32)
    for i in range(b):                               # a cache tail holds arbitrary bits: make them the worst ones
        k[i, used_k[i]:] = float("nan")
        v[i, used_k[i]:] = float("nan")
    sentinel = 7.0
    o, dq = torch.full_like(q, sentinel), torch.full_like(q, sentinel)
    dk, dv = torch.full_like(k, sentinel), torch.full_like(v, sentinel)
    lse = torch.full((b, h, s), sentinel, device=dev, dtype=torch.float32)
    stream = torch.cuda.current_stream().cuda_stream
    common = dict(b=b, seqlen_q=s, seqlen_k=s, h=h, h_k=hk, d=d, dv=d, scale_softmax=1.0 / math.sqrt(d), is_bf16=True,
                  is_causal=causal, num_splits=1, arch=100, num_sm=148, softmax_lse_ptr=lse, seqused_q=uq)
    if with_k:
        common.update(seqused_k=uk)
    hf = Handle(lib, "fa3", "fwd")
    _set_qkvo(hf, q, k, v, o)
    hf.set(**common)
    hf.call("run_mha_fwd", stream)
    hb = Handle(lib, "fa3", "bwd")
    _set_qkvo(hb, q, k, v, o)
    hb.set(**common)
    for name, t in (("do", g), ("dq", dq), ("dk", dk), ("dv", dv)):
        hb.set(**{f"{name}_ptr": t, f"{name}_batch_stride": t.stride(0), f"{name}_row_stride": t.stride(1),
                  f"{name}_head_stride": t.stride(2)})
    hb.call("run_mha_bwd", stream)
    assert capi.last_kernel_kind() == 2              # tcgen05 kernels, with or without seqused_k
    torch.cuda.synchronize()
    for i in range(b):
        nq, nk = used_q[i], used_k[i]
        mask = orc.causal_mask_bottom_right(nq, nk).to(dev) if causal else None
        ref = oracle_all(q[i:i + 1, :nq], k[i:i + 1, :nk], v[i:i + 1, :nk], g[i:i + 1, :nq], mask, lse_empty=float("-inf"))
        check_2x(f"out[{i}]", o[i:i + 1, :nq], *ref["out"])
        check_lse(lse[i:i + 1, :, :nq], ref["lse"])
        check_2x(f"dq[{i}]", dq[i:i + 1, :nq], *ref["dq"])
        check_2x(f"dk[{i}]", dk[i:i + 1, :nk], *ref["dk"])
        check_2x(f"dv[{i}]", dv[i:i + 1, :nk], *ref["dv"])
        assert (o[i, nq:] == sentinel).all() and (dq[i, nq:] == sentinel).all() and (lse[i, :, nq:] == sentinel).all()
        for t_ in (dk[i, nk:], dv[i, nk:]):            # keys beyond seqused_k receive no gradient: untouched or zero
            assert ((t_ == sentinel) | (t_ == 0)).all()
    hf.close(); hb.close()


@pytest.mark.gpu
@pytest.mark.parametrize("dynamic", [False, True])
@pytest.mark.parametrize("varlen_q", [False, True])
def test_handle_varlen_split_decode(fa, varlen_q, dynamic):
    """Split-KV decode over PACKED K/V (cu_seqlens_k) -- optionally packed queries too -- with per-batch split counts:
    num_splits_dynamic_ptr is filled by the library (prepare_varlen_num_blocks, flash_prepare_scheduler.cu:15-124) unless
    skip_scheduler_metadata_computation is set; short sequences get fewer splits, the surplus partials are empty."""
    from tests._util import check_2x
    lib = capi.load()
    torch.manual_seed(3)
    dev = "cuda"
    lens_k = [4000, 150, 1290, 64]
    lens_q = [2, 1, 3, 1] if varlen_q else [1, 1, 1, 1]
    b, h, hk, d, splits = len(lens_k), 8, 2, 128, 5
    tk, tq, msq = sum(lens_k), sum(lens_q), max(lens_q)
    k, v = [torch.randn(tk, hk, d, device=dev, dtype=torch.bfloat16) for _ in range(2)]
    cu_k = torch.tensor([0] + torch.tensor(lens_k).cumsum(0).tolist(), device=dev, dtype=torch.int32)
    cu_q = torch.tensor([0] + torch.tensor(lens_q).cumsum(0).tolist(), device=dev, dtype=torch.int32)
    if varlen_q:
        q = torch.randn(tq, h, d, device=dev, dtype=torch.bfloat16)
        lse = torch.full((h, tq), float("nan"), device=dev, dtype=torch.float32)
    else:
        q = torch.randn(b, 1, h, d, device=dev, dtype=torch.bfloat16)
        lse = torch.full((b, h, 1), float("nan"), device=dev, dtype=torch.float32)
    o = torch.full_like(q, float("nan"))
    oacc = torch.empty(splits, b, msq, h, d, device=dev, dtype=torch.float32)
    lacc = torch.empty(splits, b, h, msq, device=dev, dtype=torch.float32)
    dyn = torch.full((b,), -1, device=dev, dtype=torch.int32)
    hf = Handle(lib, "fa3", "fwd")
    for name, t, packed in (("q", q, varlen_q), ("k", k, True), ("v", v, True), ("o", o, varlen_q)):
        st = (0, t.stride(0), t.stride(1)) if packed else (t.stride(0), t.stride(1), t.stride(2))
        hf.set(**{f"{name}_ptr": t, f"{name}_batch_stride": st[0], f"{name}_row_stride": st[1], f"{name}_head_stride": st[2]})
    hf.set(b=b, seqlen_q=msq, seqlen_k=max(lens_k), total_k=tk, h=h, h_k=hk, d=d, dv=d, scale_softmax=1.0 / math.sqrt(d),
           is_bf16=True, is_causal=True, cu_seqlens_k=cu_k, softmax_lse_ptr=lse, num_sm=148, num_splits=splits,
           oaccum_ptr=oacc, softmax_lseaccum_ptr=lacc)
    if varlen_q:
        hf.set(cu_seqlens_q=cu_q, total_q=tq)
    if dynamic:
        hf.set(num_splits_dynamic_ptr=dyn)
    stream = torch.cuda.current_stream().cuda_stream
    hf.call("run_mha_fwd", stream)
    assert capi.last_kernel_kind() == 3, "varlen split decode left the TMA decode kernel"
    hf.call("run_mha_fwd_combine", stream, False)
    torch.cuda.synchronize()
    if dynamic:
        got = dyn.tolist()
        assert all(1 <= x <= splits for x in got) and got[0] >= got[2] >= got[1], got     # longer sequences, more splits
        # surplus splits of short sequences are empty
        for i in range(b):
            assert torch.isinf(lacc[got[i]:, i, :, :lens_q[i]]).all()
    for i in range(b):
        rk = slice(int(cu_k[i]), int(cu_k[i + 1]))
        qi = q[int(cu_q[i]):int(cu_q[i + 1])].unsqueeze(0) if varlen_q else q[i:i + 1]
        oi = o[int(cu_q[i]):int(cu_q[i + 1])].unsqueeze(0) if varlen_q else o[i:i + 1]
        sq, sk = qi.shape[1], lens_k[i]
        mask = orc.causal_mask_bottom_right(sq, sk).to(dev)
        ref, lse_ref = orc.attention_masked_ref(qi, k[rk].unsqueeze(0), v[rk].unsqueeze(0), mask)
        pt, _ = orc.attention_masked_ref(qi, k[rk].unsqueeze(0), v[rk].unsqueeze(0), mask, upcast=False, reorder_ops=True)
        check_2x(f"out[{i}]", oi, ref, pt)
        got_lse = lse[:, int(cu_q[i]):int(cu_q[i + 1])] if varlen_q else lse[i]
        torch.testing.assert_close(got_lse.reshape(h, sq), lse_ref[0], atol=2e-3, rtol=1e-3)
    hf.close()


@pytest.mark.gpu
def test_prepare_num_splits_formula(fa):
    """fa_b200_prepare_num_splits restates the reference's formula: blocks_per_sm = ceil(1.1 * h_k * sum m_b n_b / num_sm),
    splits_b = clamp(ceil(n_b / blocks_per_sm), 1, num_splits)."""
    lib = capi.load()
    dev = "cuda"
    used = torch.tensor([32768, 100, 4096, 9000, 1], device=dev, dtype=torch.int32)
    p = capi.new_params(torch.bfloat16)
    p.b, p.seqlen_q, p.seqlen_k, p.h, p.h_k, p.d, p.num_splits = 5, 1, 32768, 32, 8, 128, 16
    p.seqused_k = used.data_ptr()
    out = torch.zeros(5, device=dev, dtype=torch.int32)
    capi.check(lib.fa_b200_prepare_num_splits(p, out.data_ptr(), 148, torch.cuda.current_stream().cuda_stream), "prepare")
    n = [math.ceil(x / 128) for x in used.tolist()]
    m = math.ceil(1 * (32 // 8) / 16)
    bps = max(math.ceil(sum(m * x for x in n) * 1.1 * 8 / 148), 1)
    want = [max(min(math.ceil(x / bps), 16), 1) for x in n]
    assert out.tolist() == want, (out.tolist(), want)
