"""Forward/backward kernel time against sequence length at a fixed token count (65536), causal GQA 32/8 d128 bf16:
python tools/len_sweep.py   (FA_B200_LIB selects the build).  Shows the per-tile overhead of short sequences."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
fa = importlib.import_module("xhy-flash-attention_b200")
capi = importlib.import_module("xhy-flash-attention_b200._capi")
lib = capi.load()
tag = os.path.basename(capi.LIB_PATH)
torch.manual_seed(0)
out = []
for s in (512, 1024, 2048, 4096, 8192, 16384):
    b = 65536 // s
    q = torch.randn(b, s, 32, 128, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(b, s, 8, 128, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(b, s, 8, 128, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    g = torch.randn_like(q)
    def step():
        o = fa.flash_attn_func(q, k, v, causal=True)
        torch.autograd.grad(o, (q, k, v), g)
    for _ in range(3): step()
    lib.fa_b200_profile_enable(1)
    f, bw = [], []
    for _ in range(8):
        step(); torch.cuda.synchronize()
        f.append(lib.fa_b200_profile_last_ms(0)); bw.append(lib.fa_b200_profile_last_ms(1))
    lib.fa_b200_profile_enable(0)
    fl = 4.0 * b * s * s * 32 * 128 / 2
    fm, bm = min(f), min(bw)
    out.append(f"s={s}: fwd {fm:.3f} ms {fl/fm/1e9:.0f} TF, bwd {bm:.3f} ms {2.5*fl/bm/1e9:.0f} TF")
print(tag, "|", " | ".join(out), flush=True)
