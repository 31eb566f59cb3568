"""Shared helpers for the GPU parity tests: oracle calls + the reference's "2x" error criterion
(tests/test_flash_attn.py:418-427,628-639 of the reference): the kernel's max abs error against the fp32
oracle must not exceed twice the error of the same oracle evaluated in low precision (upcast=False,
reorder_ops=True), plus a tiny absolute slack for near-zero baselines."""
import torch

from oracle import attention_oracle as orc


def oracle_all(q, k, v, dout, mask=None, scale=None, lse_empty=float("inf")):
    """fp32 oracle + low-precision PyTorch baseline, both evaluated on the GPU copies of the inputs."""
    out_ref, lse_ref, dq_ref, dk_ref, dv_ref = orc.attention_grads_ref(q, k, v, dout, mask, scale, upcast=True)
    out_pt, _, dq_pt, dk_pt, dv_pt = orc.attention_grads_ref(q, k, v, dout, mask, scale, upcast=False, reorder_ops=True)
    if lse_empty != float("inf"):
        lse_ref = torch.where(torch.isinf(lse_ref), torch.full_like(lse_ref, lse_empty), lse_ref)
    return dict(out=(out_ref, out_pt), lse=lse_ref, dq=(dq_ref, dq_pt), dk=(dk_ref, dk_pt), dv=(dv_ref, dv_pt))


def check_2x(name, got, ref, pt, factor=2.0, atol=1e-4):
    err = (got.float() - ref.float()).abs().max().item()
    base = (pt.float() - ref.float()).abs().max().item()
    assert not torch.isnan(got.float()).any(), f"{name}: NaN in kernel result"
    assert err <= factor * base + atol, f"{name}: kernel err {err:.3e} > {factor} x baseline err {base:.3e}"
    return err, base


def check_lse(got, ref, atol=2e-3, rtol=1e-3):
    fin = torch.isfinite(ref)
    assert torch.equal(torch.isinf(got), ~fin), "lse: empty-row pattern differs"
    assert torch.equal(got[~fin], ref[~fin]), "lse: sign of infinity differs on empty rows"
    torch.testing.assert_close(got[fin], ref[fin], atol=atol, rtol=rtol)
