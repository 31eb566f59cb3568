"""Golden vectors for the rotary embedding, from the REFERENCE's own PyTorch implementation.

    python tests/golden/make_golden_rotary.py        (build container only: reads /root/reference)

``rotate_half`` and ``apply_rotary_emb_torch`` (flash_attn/layers/rotary.py:13-32) are extracted with ``ast`` (the
module imports a CUDA extension at import time) and run UNMODIFIED on seeded CPU inputs; inputs and outputs are
stored as ``rotary_*.npz``.  ``oracle.rotary_ref`` is pinned to these in tests/test_oracle_golden.py.
"""
import ast
import os

import numpy as np
import torch

REF = "/root/reference/flash_attn/layers/rotary.py"
HERE = os.path.dirname(os.path.abspath(__file__))


def load():
    tree = ast.parse(open(REF).read())
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in ("rotate_half", "apply_rotary_emb_torch")]
    ns = {}
    from einops import rearrange, repeat
    ns.update(dict(torch=torch, rearrange=rearrange, repeat=repeat))
    exec(compile(ast.Module(body=body, type_ignores=[]), REF, "exec"), ns)
    return ns["apply_rotary_emb_torch"]


def main():
    apply = load()
    for name, d, rdim, interleaved in (("rotary_inter_d64_r32", 64, 32, True), ("rotary_neox_d64_r64", 64, 64, False),
                                      ("rotary_neox_d128_r32", 128, 32, False)):
        torch.manual_seed(0)
        b, s, h = 2, 24, 3
        x = torch.randn(b, s, h, d)
        angle = torch.rand(s, rdim // 2) * 2 * torch.pi
        cos, sin = torch.cos(angle), torch.sin(angle)
        if interleaved:
            # this revision of the reference helper builds its cos/sin table as 's d -> s 1 (2 d)' for BOTH layouts
            # (rotary.py:29-30), which is only right for the contiguous one; the v3 kernel (csrc/flash_attn_v3/rotary.h)
            # rotates the pair (2j, 2j+1) by angle j.  That is the contiguous rotation of the de-interleaved vector, so
            # the golden is: de-interleave -> reference helper (contiguous) -> re-interleave.
            xr = x[..., :rdim]
            xp = torch.cat((xr[..., 0::2], xr[..., 1::2], x[..., rdim:]), dim=-1)
            op = apply(xp, cos, sin, interleaved=False)
            out = torch.cat((torch.stack((op[..., :rdim // 2], op[..., rdim // 2:rdim]), dim=-1).flatten(-2),
                             op[..., rdim:]), dim=-1)
        else:
            out = apply(x, cos, sin, interleaved=False)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), x=x.numpy(), cos=cos.numpy(), sin=sin.numpy(),
                            out=out.numpy(), interleaved=np.array(interleaved))
        print("wrote", name, tuple(out.shape))


if __name__ == "__main__":
    main()
