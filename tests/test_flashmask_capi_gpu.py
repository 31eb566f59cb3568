"""GPU parity: FlashMask through the FLAT FA2 C ABI -- non-NULL flashmask_downstart/upend/downend/upstart/maxmin
passed to flash_attn_fwd AND flash_attn_bwd (csrc/capi/flash_attn.h:36-43, csrc/capi/flash_attn.cu:204-210,623), the
entry Paddle's FlashMask-v1 callers use.  All four flavours: {lts} causal, {lts, lte} causal,
{lts, ute} non-causal, {lts, lte, uts, ute}; mask heads broadcast over h / h_mask q-heads.  Checked against the dense
restatement of the element rule (oracle.flashmask_dense, csrc/flash_attn/src/softmax.h:180-298) under the 2x rule."""
import pytest
import torch

from oracle import attention_oracle as orc
from tests import _flat_capi as fc
from tests._util import check_2x, check_lse, oracle_all

pytestmark = pytest.mark.gpu
DEV = "cuda"


def make_flashmask(flavour, b, h_mask, sq, sk, seed):
    """Random but valid vectors: bands lie where the (causal) visible region is, every row keeps column 0 or itself."""
    g = torch.Generator().manual_seed(seed)
    col = torch.arange(sk).view(1, 1, sk).expand(b, h_mask, sk)
    r = lambda lo, hi: torch.randint(lo, hi, (b, h_mask, sk), generator=g)
    fm = {}
    if flavour == "lts":                     # causal document-style: rows >= lts[j] do not see column j
        fm["lts"] = torch.maximum(col + 1 + r(0, sq // 2), torch.tensor(1))
    elif flavour == "lts_lte":               # causal with a masked band [lts, lte)
        lts = col + 1 + r(0, sq // 3)
        fm["lts"], fm["lte"] = lts, lts + r(0, sq // 3)
    elif flavour == "lts_ute":               # non-causal: rows >= lts and rows < ute are masked
        fm["ute"] = torch.clamp(col - r(0, sq // 3), min=0)
        fm["lts"] = col + 1 + r(0, sq // 3)
    else:                                    # two bands
        uts = torch.clamp(col - r(sq // 4, sq // 2), min=0)
        fm["uts"], fm["ute"] = uts, torch.minimum(uts + r(0, sq // 4), col)
        lts = col + 1 + r(0, sq // 4)
        fm["lts"], fm["lte"] = lts, lts + r(0, sq // 3)
    return {k: v.clamp(0, sq).int().contiguous().to(DEV) for k, v in fm.items()}


@pytest.mark.parametrize("num_splits", [0, 1], ids=["atomic", "deterministic"])
@pytest.mark.parametrize("h,hk,h_mask", [(4, 4, 4), (4, 2, 1), (6, 2, 2)])
@pytest.mark.parametrize("flavour", ["lts", "lts_lte", "lts_ute", "lts_lte_uts_ute"])
@pytest.mark.parametrize("sq,sk", [(256, 256), (200, 264), (513, 385)])
def test_flat_capi_flashmask(capi, sq, sk, flavour, h, hk, h_mask, num_splits):
    lib = capi.load()
    torch.manual_seed(0)
    b, d = 2, 128
    causal = flavour in ("lts", "lts_lte")
    q = torch.randn(b, sq, h, d, device=DEV, dtype=torch.bfloat16)
    k = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    v = torch.randn(b, sk, hk, d, device=DEV, dtype=torch.bfloat16)
    g = torch.randn(b, sq, h, d, device=DEV, dtype=torch.bfloat16)
    fm = make_flashmask(flavour, b, h_mask, sq, sk, seed=sq + sk)
    out, lse = fc.flat_fwd(lib, q, k, v, causal, fm=fm)
    dense = orc.flashmask_dense(sq, sk, fm["lts"].cpu(), fm["lte"].cpu() if "lte" in fm else None,
                                fm["uts"].cpu() if "uts" in fm else None, fm["ute"].cpu() if "ute" in fm else None)
    dense = dense.repeat_interleave(h // h_mask, dim=1).to(DEV)          # (b, h, sq, sk)
    if causal:
        dense = dense | orc.causal_mask_top_left(sq, sk).to(DEV)
    ref = oracle_all(q, k, v, g, dense)
    check_2x("out", out, *ref["out"])
    check_lse(lse, ref["lse"])
    dq, dk, dv = fc.flat_bwd(lib, g, q, k, v, out, lse, causal, fm=fm, num_splits=num_splits)
    check_2x("dq", dq, *ref["dq"])
    check_2x("dk", dk, *ref["dk"], atol=2e-4)     # summed over the GQA group on the caller's side, as the ABI demands
    check_2x("dv", dv, *ref["dv"], atol=2e-4)


def test_flat_capi_flashmask_requires_workspace(capi):
    """csrc/capi/flash_attn.cu:204-210: vectors without the maxmin workspace are an error, reported through
    flash_attn_error()."""
    import ctypes as C
    lib = capi.load()
    q = torch.randn(1, 128, 2, 128, device=DEV, dtype=torch.bfloat16)
    fm = make_flashmask("lts", 1, 1, 128, 128, 0)
    (ds, dims, ue, de, us, ws), keep = fc.flashmask_args(fm, 1, 128, DEV)
    out = torch.empty_like(q)
    lse = torch.empty(1, 2, 128, device=DEV, dtype=torch.float32)
    st = fc._strides(q)
    lib.flash_attn_fwd.restype = C.c_bool
    ok = lib.flash_attn_fwd(
        fc._vp(q), fc._vp(q), fc._vp(q), None, fc._vp(out), None, fc._vp(lse), C.c_int(1), C.c_int(128), C.c_int(128),
        C.c_int(128), C.c_int(128), C.c_int(2), C.c_int(2), C.c_int(128), C.c_int(128), C.c_float(0.0),
        C.c_float(128 ** -0.5), C.c_float(1.0), C.c_bool(True), C.c_bool(False), C.c_bool(True),
        C.c_void_p(torch.cuda.current_stream().cuda_stream), C.c_uint64(0), C.c_uint64(0), None, None,
        ds, dims, ue, de, us, None, st[0], st[0], st[0], st[1], st[1], st[1], st[0], st[1], st[2], st[2], st[2], st[2])
    assert not ok
    assert "maxmin" in capi.last_error()
