"""Forward / backward kernel times of one library build on a few dense shapes (CUDA events inside the library).
    FA_B200_FWD2=0|1 python tools/fwd_time.py [tag]        -> one JSON line
Used for A/B runs of the two forward kernels (one-thread-per-row sm100_fwd.cu vs split-row sm100_fwd2.cu)."""
import importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
fa = importlib.import_module("xhy-flash-attention_b200")
capi = importlib.import_module("xhy-flash-attention_b200._capi")
lib = capi.load()
dev, dt = "cuda", torch.bfloat16
cases = {"c3 b8 s8192 h32/8 d128 causal": (8, 8192, 32, 8, 128, True), "c2 b8 s2048 h16 d128": (8, 2048, 16, 16, 128, False),
         "b64 s1024 h32/8 d128 causal": (64, 1024, 32, 8, 128, True), "b2 s32768 h32/8 d128 causal": (2, 32768, 32, 8, 128, True),
         "b8 s8192 h32/8 d64 causal": (8, 8192, 32, 8, 64, True), "b8 s8192 h32/8 d128 noncausal": (4, 8192, 32, 8, 128, False),
         "b4 s8192 h32/8 d192 causal": (4, 8192, 32, 8, 192, True), "b4 s8192 h32/8 d256 causal": (4, 8192, 32, 8, 256, True),
         "b4 s8192 h16/16 d256 noncausal": (2, 8192, 16, 16, 256, False)}
res = {"tag": sys.argv[1] if len(sys.argv) > 1 else "", "fwd2": os.environ.get("FA_B200_FWD2", "1")}
for name, (b, s, h, hk, d, causal) in cases.items():
    torch.manual_seed(0)
    q = torch.randn(b, s, h, d, device=dev, dtype=dt, requires_grad=True)
    k = torch.randn(b, s, hk, d, device=dev, dtype=dt, requires_grad=True)
    v = torch.randn(b, s, hk, d, device=dev, dtype=dt, requires_grad=True)
    g = torch.randn_like(q)
    fa.flash_attn_func(q[:1, :256], k[:1, :256], v[:1, :256], causal=causal)
    if capi.last_kernel_kind() != 2:
        res[name] = {"kind": capi.last_kernel_kind(), "why": capi.last_fallback_reason()}
        continue
    def fwd():
        return fa.flash_attn_func(q, k, v, causal=causal)
    for _ in range(3): fwd()
    lib.fa_b200_profile_enable(1)
    ts = []
    for _ in range(7):
        fwd(); torch.cuda.synchronize(); ts.append(lib.fa_b200_profile_last_ms(0))
    tb = []
    if True:
        for _ in range(3):
            o = fwd(); torch.autograd.grad(o, (q, k, v), g); torch.cuda.synchronize(); tb.append(lib.fa_b200_profile_last_ms(1))
    lib.fa_b200_profile_enable(0)
    ts.sort(); tb.sort()
    fl = 4.0 * b * s * s * h * d / (2 if causal else 1)
    res[name] = {"fwd_ms": ts[len(ts) // 2], "fwd_tflops": fl / ts[len(ts) // 2] / 1e9,
                 "bwd_tflops": 2.5 * fl / tb[len(tb) // 2] / 1e9 if tb else None}
    del q, k, v, g
    torch.cuda.empty_cache()
print(json.dumps(res))
