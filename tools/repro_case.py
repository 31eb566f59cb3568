"""Repeat one forward case N times and print the max abs error against the fp32 oracle each time (flakiness probe).
    python tools/repro_case.py b sq sk h hk d causal dtype [reps]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import attention_oracle as orc
fa = importlib.import_module("xhy-flash-attention_b200")
b, sq, sk, h, hk, d = [int(x) for x in sys.argv[1:7]]
causal = sys.argv[7] == "1"
dt = torch.float16 if sys.argv[8] == "fp16" else torch.bfloat16
reps = int(sys.argv[9]) if len(sys.argv) > 9 else 10
torch.manual_seed(0)
q = torch.randn(b, sq, h, d, device="cuda", dtype=dt)
k = torch.randn(b, sk, hk, d, device="cuda", dtype=dt)
v = torch.randn(b, sk, hk, d, device="cuda", dtype=dt)
mask = orc.causal_mask_top_left(sq, sk).cuda() if causal else None
ref, lse_ref = orc.attention_masked_ref(q, k, v, mask)
errs = []
for i in range(reps):
    out, lse, _ = fa.flash_attn_func(q, k, v, 0.0, None, causal, True)
    torch.cuda.synchronize()
    e = (out.float() - ref.float()).abs()
    errs.append(e.max().item())
    if e.max().item() > 0.05:
        idx = (e > 0.05).nonzero()
        print("bad elements:", idx.shape[0], "first", idx[:4].tolist(), "rows", sorted(set(idx[:, 1].tolist()))[:20],
              "heads", sorted(set(idx[:, 2].tolist())), "cols", sorted(set(idx[:, 3].tolist()))[:12], "...", flush=True)
print(sys.argv[1:9], ["%.3g" % e for e in errs])
