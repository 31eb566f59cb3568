"""Quick parity probe of one kernel family in its own process (a trap poisons the CUDA context of the process):
    python tools/quick_check.py fwd|bwd d [causal]
prints max abs errors against the fp32 oracle next to the bf16-baseline errors."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import attention_oracle as orc
fa = importlib.import_module("xhy-flash-attention_b200")
capi = importlib.import_module("xhy-flash-attention_b200._capi")
what, d = sys.argv[1], int(sys.argv[2])
dev = "cuda"
ok = True
for (b, sq, sk, h, hk, causal) in [(2, 113, 203, 6, 2, False), (2, 256, 512, 6, 2, True), (1, 1023, 1024, 4, 4, True), (2, 384, 384, 8, 2, True),
                                   (1, 2048, 2048, 8, 2, True), (2, 512, 256, 6, 2, False), (1, 700, 1500, 4, 1, False)]:
    torch.manual_seed(0)
    q = torch.randn(b, sq, h, d, device=dev, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(b, sk, hk, d, device=dev, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(b, sk, hk, d, device=dev, dtype=torch.bfloat16, requires_grad=True)
    g = torch.randn(b, sq, h, d, device=dev, dtype=torch.bfloat16)
    capi.fallback_count(reset=True)
    out, lse, _ = fa.flash_attn_func(q, k, v, 0.0, None, causal, True)
    torch.cuda.synchronize()
    kind = capi.last_kernel_kind()
    mask = orc.causal_mask_top_left(sq, sk).to(dev) if causal else None
    ref = orc.attention_grads_ref(q.detach(), k.detach(), v.detach(), g, mask)
    pt = orc.attention_grads_ref(q.detach(), k.detach(), v.detach(), g, mask, upcast=False, reorder_ops=True)
    got = [out, lse]
    if what == "bwd":
        got += list(torch.autograd.grad(out, (q, k, v), g))
        torch.cuda.synchronize()
    line = f"{what} d{d} b{b} sq{sq} sk{sk} h{h}/{hk} causal={causal} kind={kind} fallbacks={capi.fallback_count()}:"
    for n, a, r_, p_ in zip(("out", "lse", "dq", "dk", "dv"), got, ref, pt):
        if n == "lse":
            fin = torch.isfinite(r_)
            e = (a[fin] - r_[fin]).abs().max().item()
            line += f" lse {e:.2e}"
            ok &= e < 5e-3
            continue
        e, base = (a.float() - r_.float()).abs().max().item(), (p_.float() - r_.float()).abs().max().item()
        bad = not (e <= 2 * base + 2e-4) or torch.isnan(a.float()).any().item()
        ok &= not bad
        line += f" {n} {e:.2e}/{base:.2e}{' !!' if bad else ''}"
    print(line, flush=True)
print("QUICK_OK" if ok else "QUICK_FAIL", what, d, flush=True)
