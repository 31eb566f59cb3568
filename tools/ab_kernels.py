"""Kernel timing of one library build: python tools/ab_kernels.py  (FA_B200_LIB selects the .so).  Prints fwd/bwd kernel ms."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
fa = importlib.import_module("xhy-flash-attention_b200")
capi = importlib.import_module("xhy-flash-attention_b200._capi")
lib = capi.load()
torch.manual_seed(0)
tag = os.path.basename(capi.LIB_PATH)
for name, (b, s, h, hk, causal) in {"c3": (8, 8192, 32, 8, True), "c2": (8, 2048, 16, 16, False)}.items():
    q = torch.randn(b, s, h, 128, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(b, s, hk, 128, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(b, s, hk, 128, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    g = torch.randn_like(q)
    def step():
        o = fa.flash_attn_func(q, k, v, causal=causal)
        torch.autograd.grad(o, (q, k, v), g)
    for _ in range(5): step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): step()
    e1.record(); torch.cuda.synchronize()
    tot = e0.elapsed_time(e1) / 20
    lib.fa_b200_profile_enable(1)
    f, bw = [], []
    for _ in range(10):
        step(); torch.cuda.synchronize()
        f.append(lib.fa_b200_profile_last_ms(0)); bw.append(lib.fa_b200_profile_last_ms(1))
    lib.fa_b200_profile_enable(0)
    fl = 4.0 * b * s * s * h * 128 / (2 if causal else 1)
    fm, bm = sum(f) / 10, sum(bw) / 10
    print(f"{tag:28s} {name}: step {tot:7.3f} ms ({3.5*fl/tot/1e9:6.0f} TF) | fwd {fm:6.3f} ms ({fl/fm/1e9:5.0f} TF) | bwd {bm:6.3f} ms ({2.5*fl/bm/1e9:5.0f} TF)", flush=True)
