"""FP8 forward timing on the C3 shape (b8 s8192 hq32/hkv8 d128 causal) and d256, next to the bf16 forward."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
fa = importlib.import_module("xhy-flash-attention_b200")

def timeit(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
    ev[0].record()
    for i in range(n):
        fn(); ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(n))
    return ts[len(ts) // 2]

for (b, s, h, hk, d) in [(8, 8192, 32, 8, 128), (4, 8192, 32, 8, 256), (8, 8192, 32, 8, 64)]:
    q = torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16)
    k = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16)
    v = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16)
    q8, k8, v8 = [x.to(torch.float8_e4m3fn) for x in (q, k, v)]
    fl = 4 * b * s * s * h * d / 2
    t16 = timeit(lambda: fa.flash_attn_func(q, k, v, causal=True))
    t8 = timeit(lambda: fa.flash_attn_fp8_func(q8, k8, v8, causal=True))
    print(f"d{d}: bf16 {t16:.3f} ms {fl / t16 / 1e9:.0f} TFLOPS | e4m3 {t8:.3f} ms {fl / t8 / 1e9:.0f} TFLOPS", flush=True)
