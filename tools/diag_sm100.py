"""Development diagnostics for the tcgen05 kernels (not a test): runs small structured cases, each in its own
process (a device trap poisons the CUDA context), and prints where the tensor-core path deviates from the
generic SIMT kernels / a dense fp32 softmax.  Usage: python tools/diag_sm100.py [case ...]"""
import importlib
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CASES = ["p_identity", "fwd_small", "fwd_causal", "fwd_multi", "fwd_d64", "fwd_varlen", "bwd_small", "bwd_causal",
         "bwd_gqa", "time_fwd", "time_bwd"]


def run_case(name):
    import torch
    fa = importlib.import_module("xhy-flash-attention_b200")
    capi = importlib.import_module("xhy-flash-attention_b200._capi")
    dev = "cuda"
    torch.manual_seed(0)

    def cmp(tag, a, b):
        a, b = a.float(), b.float()
        err = (a - b).abs()
        print(f"  {tag}: max_err {err.max().item():.4e} mean_err {err.mean().item():.4e} ref_absmax {b.abs().max().item():.3e}"
              f" nan {int(torch.isnan(a).sum())}")
        if err.max().item() > 0.05 * max(1e-6, b.abs().max().item()):
            idx = (err > 0.05 * b.abs().max()).nonzero()
            print(f"    bad elements: {idx.shape[0]} / {err.numel()}; first {idx[:6].tolist()}")
            if a.dim() == 4:
                bad_rows = sorted(set(idx[:, 1].tolist()))
                bad_cols = sorted(set(idx[:, 3].tolist()))
                print(f"    bad rows (first 16) {bad_rows[:16]} n={len(bad_rows)}; bad cols (first 16) {bad_cols[:16]} n={len(bad_cols)}")

    def dense_ref(q, k, v):
        # fp32 softmax(q k^T / sqrt(d)) v on the device, one head per column of the layout (b, s, h, d)
        qf, kf, vf = (t.float().permute(0, 2, 1, 3) for t in (q, k, v))
        s = qf @ kf.transpose(-1, -2) / q.shape[-1] ** 0.5
        return (torch.softmax(s, -1) @ vf).permute(0, 2, 1, 3), torch.logsumexp(s, -1)

    def fwd_case(b, sq, sk, h, hk, d, causal, dtype=torch.bfloat16):
        q = torch.randn(b, sq, h, d, device=dev, dtype=dtype)
        k = torch.randn(b, sk, hk, d, device=dev, dtype=dtype)
        v = torch.randn(b, sk, hk, d, device=dev, dtype=dtype)
        o_tc, l_tc, _ = fa.flash_attn_func(q, k, v, 0.0, None, causal, True)
        print("  kernel kind:", capi.last_kernel_kind())
        o_g, l_g, _ = fa.flash_attn_func(q, k, v, 0.0, None, causal, True, _force_generic=True)
        torch.cuda.synchronize()
        cmp("out tc vs generic", o_tc, o_g)
        cmp("lse tc vs generic", l_tc, l_g)

    def bwd_case(b, sq, sk, h, hk, d, causal, dtype=torch.bfloat16):
        q = torch.randn(b, sq, h, d, device=dev, dtype=dtype, requires_grad=True)
        k = torch.randn(b, sk, hk, d, device=dev, dtype=dtype, requires_grad=True)
        v = torch.randn(b, sk, hk, d, device=dev, dtype=dtype, requires_grad=True)
        g = torch.randn(b, sq, h, d, device=dev, dtype=dtype)
        res = []
        for generic in (False, True):
            o = fa.flash_attn_func(q, k, v, 0.0, None, causal, False, _force_generic=generic)
            res.append(torch.autograd.grad(o, (q, k, v), g))
            if not generic:
                print("  bwd kernel kind:", capi.last_kernel_kind())
        torch.cuda.synchronize()
        for n, a, b_ in zip(("dq", "dk", "dv"), *res):
            cmp(f"{n} tc vs generic", a, b_)

    def timeit(fn, n=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    if name == "p_identity":
        # V = I (d = 128 = sk): out == softmax probabilities, exposes S / P layout errors directly
        sq = sk = d = 128
        q = torch.randn(1, sq, 1, d, device=dev, dtype=torch.bfloat16)
        k = torch.randn(1, sk, 1, d, device=dev, dtype=torch.bfloat16)
        v = torch.eye(sk, device=dev, dtype=torch.bfloat16).view(1, sk, 1, d)
        o_tc, l_tc, _ = fa.flash_attn_func(q, k, v, 0.0, None, False, True)
        print("  kernel kind:", capi.last_kernel_kind())
        ref, lse = dense_ref(q, k, v)
        cmp("P (out with V=I) vs dense fp32", o_tc, ref)
        cmp("lse vs dense fp32", l_tc, lse)
        # V = one-hot shifted: detects d-column permutations in the PV B operand
        v2 = torch.randn(1, sk, 1, d, device=dev, dtype=torch.bfloat16)
        o2 = fa.flash_attn_func(q, k, v2)
        ref2, _ = dense_ref(q, k, v2)
        cmp("out random V vs dense fp32", o2, ref2)
    elif name == "fwd_small":
        fwd_case(1, 128, 128, 1, 1, 128, False)
        fwd_case(1, 100, 77, 1, 1, 128, False)
    elif name == "fwd_causal":
        fwd_case(1, 256, 256, 2, 2, 128, True)
        fwd_case(2, 300, 517, 4, 2, 128, True)
    elif name == "fwd_multi":
        fwd_case(2, 1024, 1024, 8, 2, 128, False)
        fwd_case(3, 2048, 2048, 16, 4, 128, True, torch.float16)
    elif name == "fwd_d64":
        fwd_case(2, 512, 512, 4, 4, 64, False)
        fwd_case(2, 300, 300, 4, 2, 96, True)
        fwd_case(2, 300, 300, 4, 2, 40, True)
    elif name == "fwd_varlen":
        lens = [76, 34, 146, 1, 300, 129]
        cu = torch.tensor([0] + list(__import__("itertools").accumulate(lens)), device=dev, dtype=torch.int32)
        t = sum(lens)
        q = torch.randn(t, 4, 128, device=dev, dtype=torch.bfloat16)
        k = torch.randn(t, 2, 128, device=dev, dtype=torch.bfloat16)
        v = torch.randn(t, 2, 128, device=dev, dtype=torch.bfloat16)
        a = fa.flash_attn_varlen_func(q, k, v, cu, cu, max(lens), max(lens), causal=True)
        print("  kernel kind:", capi.last_kernel_kind())
        b_ = fa.flash_attn_varlen_func(q, k, v, cu, cu, max(lens), max(lens), causal=True, _force_generic=True)
        cmp("varlen out tc vs generic", a[None], b_[None])
    elif name == "bwd_small":
        bwd_case(1, 128, 128, 1, 1, 128, False)
        bwd_case(1, 256, 384, 1, 1, 128, False)
        bwd_case(1, 100, 77, 1, 1, 128, False)
    elif name == "bwd_causal":
        bwd_case(2, 512, 512, 2, 2, 128, True)
        bwd_case(2, 300, 517, 4, 4, 128, True)
    elif name == "bwd_gqa":
        bwd_case(2, 512, 512, 8, 2, 128, True)
        bwd_case(2, 333, 333, 4, 1, 64, False)
    elif name in ("time_fwd", "time_bwd"):
        for (b, s, h, hk, causal) in ((8, 2048, 16, 16, False), (2, 8192, 32, 8, True)):
            q = torch.randn(b, s, h, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
            k = torch.randn(b, s, hk, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
            v = torch.randn(b, s, hk, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
            flops = 4 * b * s * s * h * 128 / (2 if causal else 1)
            if name == "time_fwd":
                with torch.no_grad():
                    ms = timeit(lambda: fa.flash_attn_func(q, k, v, causal=causal))
                print(f"  fwd b{b} s{s} h{h}/{hk} causal={causal}: {ms:.3f} ms  {flops / ms / 1e9:.1f} TFLOPS (kind {capi.last_kernel_kind()})")
            else:
                o = fa.flash_attn_func(q, k, v, causal=causal)
                g = torch.randn_like(o)
                ms = timeit(lambda: torch.autograd.grad(o, (q, k, v), g, retain_graph=True))
                print(f"  bwd b{b} s{s} h{h}/{hk} causal={causal}: {ms:.3f} ms  {2.5 * flops / ms / 1e9:.1f} TFLOPS (kind {capi.last_kernel_kind()})")
    else:
        raise SystemExit(f"unknown case {name}")


if __name__ == "__main__":
    if len(sys.argv) >= 3 and sys.argv[1] == "--one":
        run_case(sys.argv[2])
        sys.exit(0)
    cases = sys.argv[1:] or CASES
    for c in cases:
        print(f"== {c}", flush=True)
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--one", c], capture_output=True, text=True,
                               timeout=240)
            print(r.stdout, end="")
            if r.returncode != 0:
                print(f"  EXIT {r.returncode}\n  " + "\n  ".join(r.stderr.strip().splitlines()[-12:]))
        except subprocess.TimeoutExpired:
            print("  TIMEOUT")
        sys.stdout.flush()
