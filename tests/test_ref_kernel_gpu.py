"""GPU: this library against the REFERENCE'S OWN FA2 KERNELS on the same box, same inputs, through the same flat C ABI.

``oracle/_ref/libflashattn_ref.so`` is the reference's csrc/capi/flash_attn.cu + its hdim128 bf16 kernels (plain, causal,
flashmask, causal_flashmask; fwd + bwd) + calc_reduced_attn_scores, compiled for sm_100 from the sources where they lie
under /root/reference by oracle/build_ref.sh (test infrastructure; nothing in the product loads it).  It is what
north_star's tolerance names: "max abs error vs an fp32 attention_ref no more than 2x the REFERENCE KERNEL'S OWN error".
It also pins FlashMask v1 (all four vector flavours) and calc_reduced_attn_scores against outputs the reference produced.
Skipped when the comparator was not built (no /root/reference at build time)."""
import ctypes as C
import os

import pytest
import torch

from oracle import attention_oracle as orc
from tests import _flat_capi as fc
from tests.test_flashmask_capi_gpu import make_flashmask

pytestmark = pytest.mark.gpu
DEV = "cuda"
REF_SO = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libflashattn_ref.so")


@pytest.fixture(scope="module")
def reflib():
    if not os.path.exists(REF_SO):
        pytest.skip("reference comparator not built (oracle/build_ref.sh)")
    return C.CDLL(REF_SO, mode=os.RTLD_LAZY)     # only the hdim128 bf16 specialisations are linked in


def _err(x, ref):
    return (x.float() - ref.float()).abs().max().item()


def _inputs(b, sq, sk, h, hk, d=128, seed=0):
    torch.manual_seed(seed)
    mk = lambda *s: torch.randn(*s, device=DEV, dtype=torch.bfloat16)
    return mk(b, sq, h, d), mk(b, sk, hk, d), mk(b, sk, hk, d), mk(b, sq, h, d)


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("h,hk", [(4, 4), (8, 2)], ids=["mha", "gqa"])
@pytest.mark.parametrize("sq,sk", [(512, 512), (384, 640), (1024, 1024), (113, 203)])
def test_error_vs_reference_kernel(capi, reflib, sq, sk, h, hk, causal):
    """north_star's criterion, literally: our max abs error against the fp32 oracle <= 2 x the reference kernel's."""
    lib = capi.load()
    q, k, v, g = _inputs(2, sq, sk, h, hk)
    mask = orc.causal_mask_top_left(sq, sk).to(DEV) if causal else None
    o_ref, lse_ref, dq_ref, dk_ref, dv_ref = orc.attention_grads_ref(q, k, v, g, mask)
    ours_o, ours_lse = fc.flat_fwd(lib, q, k, v, causal)
    ref_o, ref_lse = fc.flat_fwd(reflib, q, k, v, causal)
    ours_g = fc.flat_bwd(lib, g, q, k, v, ours_o, ours_lse, causal)
    ref_g = fc.flat_bwd(reflib, g, q, k, v, ref_o, ref_lse, causal)
    fin = torch.isfinite(lse_ref)
    for name, a, r, want in (("out", ours_o, ref_o, o_ref), ("dq", ours_g[0], ref_g[0], dq_ref),
                             ("dk", ours_g[1], ref_g[1], dk_ref), ("dv", ours_g[2], ref_g[2], dv_ref)):
        e_ours, e_ref = _err(a, want), _err(r, want)
        assert e_ours <= 2.0 * e_ref + 1e-4, f"{name}: ours {e_ours:.3e} > 2 x reference kernel {e_ref:.3e}"
    assert _err(ours_lse[fin], lse_ref[fin]) <= 2.0 * _err(ref_lse[fin], lse_ref[fin]) + 1e-3
    assert torch.equal(torch.isinf(ours_lse), torch.isinf(ref_lse))       # same +inf pattern for empty rows


@pytest.mark.parametrize("h,hk,h_mask", [(4, 4, 4), (4, 2, 1)])
@pytest.mark.parametrize("flavour", ["lts", "lts_lte", "lts_ute", "lts_lte_uts_ute"])
@pytest.mark.parametrize("sq,sk", [(256, 256), (512, 512), (384, 640)])
def test_flashmask_v1_vs_reference_kernel(capi, reflib, sq, sk, flavour, h, hk, h_mask):
    """FlashMask v1 through flash_attn_fwd / flash_attn_bwd: identical argument lists go to both libraries; outputs
    agree to 16-bit rounding and ours is within 2x of the reference kernel's error against the dense oracle."""
    lib = capi.load()
    causal = flavour in ("lts", "lts_lte")
    q, k, v, g = _inputs(2, sq, sk, h, hk, seed=1)
    fm = make_flashmask(flavour, 2, h_mask, sq, sk, seed=sq * 7 + sk)
    ours_o, ours_lse = fc.flat_fwd(lib, q, k, v, causal, fm=fm)
    ref_o, ref_lse = fc.flat_fwd(reflib, q, k, v, causal, fm=fm)
    ours_g = fc.flat_bwd(lib, g, q, k, v, ours_o, ours_lse, causal, fm=fm)
    ref_g = fc.flat_bwd(reflib, g, q, k, v, ref_o, ref_lse, causal, fm=fm)
    dense = orc.flashmask_dense(sq, sk, fm["lts"].cpu(), fm["lte"].cpu() if "lte" in fm else None,
                                fm["uts"].cpu() if "uts" in fm else None, fm["ute"].cpu() if "ute" in fm else None)
    dense = dense.repeat_interleave(h // h_mask, dim=1).to(DEV)
    if causal:
        dense = dense | orc.causal_mask_top_left(sq, sk).to(DEV)
    want = orc.attention_grads_ref(q, k, v, g, dense)
    # rows the mask empties completely: the reference kernel's output there is not a defined quantity
    # (csrc/flash_attn/src/flash_fwd_kernel.h:513-514 writes out = 0 * garbage-free accumulators, lse = +inf) -- compare
    # the pattern, then only the rows with at least one visible key
    live = ~dense.all(dim=-1)                                            # (b, h, sq)
    assert torch.equal(torch.isinf(ours_lse), ~live)
    rows = live.transpose(1, 2).unsqueeze(-1)                            # (b, sq, h, 1)
    for name, a, r, w in (("out", ours_o, ref_o, want[0]), ("dq", ours_g[0], ref_g[0], want[2])):
        a, r, w = a * rows, r * rows, w * rows
        assert _err(a, w) <= 2.0 * _err(r, w) + 1e-4, f"{name}: ours {_err(a, w):.3e} vs reference kernel {_err(r, w):.3e}"
        torch.testing.assert_close(a.float(), r.float(), atol=3e-2, rtol=3e-2)
    for name, a, r, w in (("dk", ours_g[1], ref_g[1], want[3]), ("dv", ours_g[2], ref_g[2], want[4])):
        assert _err(a, w) <= 2.0 * _err(r, w) + 2e-4, f"{name}: ours {_err(a, w):.3e} vs reference kernel {_err(r, w):.3e}"
    assert _err(ours_lse[live], want[1][live]) <= 2.0 * _err(ref_lse[live], want[1][live]) + 1e-3


@pytest.mark.parametrize("return_softmax", [False, True])
@pytest.mark.parametrize("h,hk", [(4, 4), (4, 2)])
def test_reduced_scores_vs_reference_kernel(capi, reflib, h, hk, return_softmax):
    lib = capi.load()
    q, k, v, _ = _inputs(2, 256, 384, h, hk, seed=2)
    _, lse = fc.flat_fwd(reflib, q, k, v, causal=False)
    red_ref, p_ref = fc.reduced_scores(reflib, q, k, lse, return_softmax=return_softmax)
    red, p = fc.reduced_scores(lib, q, k, lse, return_softmax=return_softmax)
    torch.testing.assert_close(red, red_ref, atol=2e-3, rtol=2e-3)
    if return_softmax:
        torch.testing.assert_close(p, p_ref, atol=1e-4, rtol=2e-3)
