"""One small fwd+bwd per head dim for ncu captures: python tools/ncu_case.py [b s h hk causal reps [d,d,...]]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
fa = importlib.import_module("xhy-flash-attention_b200")
a = sys.argv[1:]
b, s, h, hk, causal, reps = (int(a[0]), int(a[1]), int(a[2]), int(a[3]), bool(int(a[4])), int(a[5])) if len(a) >= 6 else (2, 4096, 16, 4, True, 2)
dims = [int(x) for x in a[6].split(",")] if len(a) >= 7 else [128]
torch.manual_seed(0)
for d in dims:
    q = torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    g = torch.randn_like(q)
    for _ in range(reps):
        o = fa.flash_attn_func(q, k, v, causal=causal)
        torch.autograd.grad(o, (q, k, v), g)
    torch.cuda.synchronize()
print("done")
