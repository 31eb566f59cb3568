import importlib, torch, sys
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
fa = importlib.import_module("xhy-flash-attention_b200")
capi = importlib.import_module("xhy-flash-attention_b200._capi"); lib = capi.load()
for (b, s, h, hk, d, causal) in [(8, 8192, 32, 8, 64, True), (8, 8192, 32, 32, 64, True), (8, 2048, 16, 16, 64, False), (8, 8192, 32, 8, 96, True), (8, 8192, 32, 8, 128, True)]:
    q = torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    g = torch.randn_like(q)
    fl = 4.0 * b * s * s * h * d / (2 if causal else 1)
    def step():
        o = fa.flash_attn_func(q, k, v, causal=causal)
        torch.autograd.grad(o, (q, k, v), g)
    step(); step()
    lib.fa_b200_profile_enable(1)
    f, bw = [], []
    for _ in range(5):
        step(); torch.cuda.synchronize(); f.append(lib.fa_b200_profile_last_ms(0)); bw.append(lib.fa_b200_profile_last_ms(1))
    lib.fa_b200_profile_enable(0)
    fm, bm = sorted(f)[2], sorted(bw)[2]
    print(f"b{b} s{s} h{h}/{hk} d{d} causal={causal}: fwd {fm:.3f} ms ({fl/fm/1e9:.0f} TF)  bwd {bm:.3f} ms ({2.5*fl/bm/1e9:.0f} TF) kind {capi.last_kernel_kind()}", flush=True)
