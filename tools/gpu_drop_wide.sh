#!/bin/bash
timeout 900 python -m pytest tests/test_dropout_gpu.py tests/test_flash_attn_gpu.py -m gpu -q --timeout 300 -x 2>&1 | tail -4
timeout 300 python - <<'PY' 2>&1 | grep -v Warning | tail -4
import importlib, torch
fa = importlib.import_module("xhy-flash-attention_b200")
capi = importlib.import_module("xhy-flash-attention_b200._capi")
for d in (192, 256):
    b, s, h, hk = 4, 8192, 32, 8
    q = torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    g = torch.randn_like(q)
    fl = 4.0 * b * s * s * h * d / 2
    def step(p):
        o = fa.flash_attn_func(q, k, v, p, None, True); torch.autograd.grad(o, (q, k, v), g)
    for p in (0.0, 0.1):
        for _ in range(2): step(p)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3): step(p)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        print(f"d{d} p={p}: fwd+bwd {ms:.2f} ms {3.5 * fl / ms / 1e9:.0f} TFLOPS kind {capi.last_kernel_kind()} fallbacks {capi.fallback_count()}")
PY
