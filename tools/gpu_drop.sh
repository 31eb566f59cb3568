#!/bin/bash
mkdir -p gpurun_out
D=xhy-flash-attention_b200/lib_ab
timeout 900 python -m pytest tests/test_dropout_gpu.py -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/drop_pytest.log
timeout 900 python -m pytest tests/test_flash_attn_gpu.py tests/test_features_gpu.py tests/test_decode_capi_gpu.py -m gpu -x -q 2>&1 | tail -2 | tee -a gpurun_out/drop_pytest.log
AB_CASES="c3" timeout 600 python tools/ab_inproc.py $D/lib_head.so new 2>&1 | tee gpurun_out/ab_drop.log
python - <<'PY' 2>&1 | tee -a gpurun_out/ab_drop.log
import importlib, torch
fa = importlib.import_module("xhy-flash-attention_b200")
capi = importlib.import_module("xhy-flash-attention_b200._capi"); lib = capi.load()
b, s, h, hk = 8, 8192, 32, 8
q = torch.randn(b, s, h, 128, device="cuda", dtype=torch.bfloat16, requires_grad=True)
k = torch.randn(b, s, hk, 128, device="cuda", dtype=torch.bfloat16, requires_grad=True)
v = torch.randn(b, s, hk, 128, device="cuda", dtype=torch.bfloat16, requires_grad=True)
g = torch.randn_like(q)
fl = 4.0 * b * s * s * h * 128 / 2
for p in (0.0, 0.1):
    def step():
        o = fa.flash_attn_func(q, k, v, p, None, True)
        torch.autograd.grad(o, (q, k, v), g)
    step(); step()
    lib.fa_b200_profile_enable(1)
    f, bw = [], []
    for _ in range(5):
        step(); torch.cuda.synchronize(); f.append(lib.fa_b200_profile_last_ms(0)); bw.append(lib.fa_b200_profile_last_ms(1))
    lib.fa_b200_profile_enable(0)
    fm, bm = sorted(f)[2], sorted(bw)[2]
    print(f"C3 dropout_p={p}: fwd {fm:.3f} ms ({fl/fm/1e9:.0f} TF)  bwd {bm:.3f} ms ({2.5*fl/bm/1e9:.0f} TF)  kind {capi.last_kernel_kind()}")
PY
