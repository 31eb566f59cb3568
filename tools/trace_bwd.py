"""Timeline of the backward main kernel on C3 from the FA_TRACE variant build (see tools/trace_fwd.py).  CTA (0,0,0):
key block 0 of (batch 0, kv-head 0): it walks all 64 query blocks of its 4 q-heads."""
import ctypes as C, importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
fa = importlib.import_module("xhy-flash-attention_b200")
capi = importlib.import_module("xhy-flash-attention_b200._capi")
lib = capi.load()
b, s, h, hk, d = 8, 8192, 32, 8, 128
q = torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16, requires_grad=True)
k = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16, requires_grad=True)
v = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16, requires_grad=True)
g = torch.randn_like(q)
for _ in range(3):
    o = fa.flash_attn_func(q, k, v, causal=True)
    torch.autograd.grad(o, (q, k, v), g)
torch.cuda.synchronize()
buf = (C.c_ulonglong * (4 * 96 * 8))()
assert lib.fa_b200_trace_bwd_dump(buf) == 0
t = np.array(buf, dtype=np.int64).reshape(4, 96, 8)
print("P = exps done + (st wait, arrive); dS = (first dp chunk + math) + second chunk till st-wait + fences/arrive")
print("it | WG0: waitS P waitdP dS | WG1: waitS P waitdP dS | issuer: waitDO/DQE ->dP; waitP ; (dV,S) ; waitDS ; | reducer: waitDQ ld rest | period")
for it in range(4, 40):
    row = []
    for r in range(2):
        e = t[r, it]
        row.append((e[1] - e[0], f"{e[5] - e[1]}+{e[2] - e[5]}", e[3] - e[2], f"{e[6] - e[3]}+{e[7] - e[6]}+{e[4] - e[7]}"))
    i = t[2, it]
    rd = t[3, it]
    per = t[2, it, 0] - t[2, it - 1, 0]
    print(it, "|", *row[0], "|", *row[1], "|", i[1] - i[0], i[2] - i[1], i[3] - i[2], i[4] - i[3], "|", rd[1] - rd[0], rd[2] - rd[1], t[3, it + 1, 0] - rd[2], "|", per)
