"""Decode case for ncu captures: python tools/ncu_decode.py [b ctx reps]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
fa = importlib.import_module("xhy-flash-attention_b200")
a = sys.argv[1:]
b, ctx, reps = (int(a[0]), int(a[1]), int(a[2])) if len(a) >= 3 else (64, 32768, 3)
torch.manual_seed(0)
q = torch.randn(b, 1, 32, 128, device="cuda", dtype=torch.bfloat16)
kc = torch.randn(b, ctx, 8, 128, device="cuda", dtype=torch.bfloat16)
vc = torch.randn(b, ctx, 8, 128, device="cuda", dtype=torch.bfloat16)
sl = torch.full((b,), ctx, device="cuda", dtype=torch.int32)
for _ in range(reps):
    o = fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=sl, causal=True)
torch.cuda.synchronize()
print("done")
