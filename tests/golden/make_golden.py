"""Generate golden vectors by running the REFERENCE's own oracle code.

Run once in the build container (where /root/reference exists):

    python tests/golden/make_golden.py

It extracts ``attention_ref`` (tests/test_flash_attn.py:111-159) and the
padding helpers (``generate_random_padding_mask`` :26-35) from the reference
tree with ``ast`` -- the module itself cannot be imported without a GPU
(module-level ``torch.cuda.get_device_capability``, :20-23) -- executes them
UNMODIFIED on seeded CPU inputs and stores inputs + outputs (+ autograd grads)
as small ``.npz`` fixtures.  Nothing at test/bench time reads /root/reference;
only these committed fixtures travel.
"""
import ast
import math
import os
import sys

import numpy as np
import torch

REF = "/root/reference/tests/test_flash_attn.py"
HERE = os.path.dirname(os.path.abspath(__file__))


def load_reference_functions():
    src = open(REF).read()
    tree = ast.parse(src)
    wanted = {"attention_ref", "generate_random_padding_mask"}
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in wanted]
    mod = ast.Module(body=body, type_ignores=[])
    ns = {}
    from einops import rearrange, repeat
    ns.update(dict(torch=torch, math=math, rearrange=rearrange, repeat=repeat))
    exec(compile(mod, REF, "exec"), ns)
    return ns["attention_ref"], ns["generate_random_padding_mask"]


CASES = [
    # name, b, sq, sk, h, hk, d, causal, padding
    ("mha_nc_d64", 2, 48, 48, 4, 4, 64, False, None),
    ("mha_causal_d64", 2, 64, 64, 4, 4, 64, True, None),
    ("gqa_causal_d128", 1, 80, 80, 6, 2, 128, True, None),
    ("mqa_nc_d32_sq_ne_sk", 2, 33, 71, 3, 1, 32, False, None),
    ("causal_sq_gt_sk", 1, 96, 40, 2, 2, 64, True, None),   # rows with every key masked? no: top-left keeps col 0
    ("causal_sq_lt_sk", 1, 40, 96, 2, 2, 64, True, None),
    ("padded_random", 3, 64, 64, 4, 2, 64, False, "random"),
    ("padded_causal", 3, 64, 64, 4, 4, 40, True, "random"),
]


def main():
    attention_ref, gen_pad = load_reference_functions()
    for name, b, sq, sk, h, hk, d, causal, padding in CASES:
        torch.manual_seed(0)
        q = torch.randn(b, sq, h, d, dtype=torch.float32, requires_grad=True)
        k = torch.randn(b, sk, hk, d, dtype=torch.float32, requires_grad=True)
        v = torch.randn(b, sk, hk, d, dtype=torch.float32, requires_grad=True)
        g = torch.randn(b, sq, h, d, dtype=torch.float32)
        qpm = kpm = None
        if padding is not None:
            qpm = gen_pad(sq, b, "cpu", mode=padding)
            kpm = gen_pad(sk, b, "cpu", mode=padding)
        out, attn = attention_ref(q, k, v, qpm, kpm, 0.0, None, causal=causal)
        dq, dk, dv = torch.autograd.grad(out, (q, k, v), g)
        # the reference's low-precision baseline (upcast=False, reorder_ops=True) in bf16
        out_pt, _ = attention_ref(q.detach().bfloat16(), k.detach().bfloat16(), v.detach().bfloat16(),
                                  qpm, kpm, 0.0, None, causal=causal, upcast=False, reorder_ops=True)
        arrs = dict(q=q.detach().numpy(), k=k.detach().numpy(), v=v.detach().numpy(), g=g.numpy(),
                    out=out.detach().numpy(), attn=attn.detach().numpy(),
                    dq=dq.numpy(), dk=dk.numpy(), dv=dv.numpy(),
                    out_pt_bf16=out_pt.float().numpy(),
                    causal=np.array(causal))
        if qpm is not None:
            arrs["query_padding_mask"] = qpm.numpy()
            arrs["key_padding_mask"] = kpm.numpy()
        path = os.path.join(HERE, f"{name}.npz")
        np.savez_compressed(path, **arrs)
        print("wrote", path, {k_: v_.shape for k_, v_ in arrs.items() if hasattr(v_, "shape")})


if __name__ == "__main__":
    if not os.path.exists(REF):
        sys.exit("reference tree not present; golden fixtures are generated in the build container only")
    main()
