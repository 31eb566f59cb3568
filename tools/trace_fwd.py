"""Timeline of the one-thread-per-row forward kernel on C3, from the FA_TRACE variant build:
    bash tools/build_variant.sh trace "-DFA_TRACE"
    FA_B200_LIB=$PWD/xhy-flash-attention_b200/lib_ab/lib_trace.so FA_B200_FWD2=0 python tools/trace_fwd.py
Prints, for CTA 0's first work tile, per key block and per softmax warpgroup: wait for S, TMEM load, max (+mask), exp
up to the first half of P, second half; and for the MMA issuer: wait for P (half / full) and issue times."""
import ctypes as C, importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
fa = importlib.import_module("xhy-flash-attention_b200")
capi = importlib.import_module("xhy-flash-attention_b200._capi")
lib = capi.load()
b, s, h, hk, d = 8, 8192, 32, 8, 128
q = torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16)
k = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16)
v = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16)
for _ in range(3):
    fa.flash_attn_func(q, k, v, causal=True)
torch.cuda.synchronize()
buf = (C.c_ulonglong * (3 * 96 * 8))()
assert lib.fa_b200_trace_dump(buf) == 0
import numpy as np
t = np.array(buf, dtype=np.int64).reshape(3, 96, 8)
t0 = t[0, 0, 0]
print("block | WG0: waitS ld max exp1 exp2 total | WG1: waitS ld max exp1 exp2 total | issuer t0: waitPh waitPf issueS | t1: waitPh waitPf issueS | period")
prev = None
for blk in range(2, 40):
    row = []
    for r in range(2):
        e = t[r, blk]
        row.append((e[1] - e[0], e[2] - e[1], e[3] - e[2], e[4] - e[3], e[5] - e[4], e[5] - e[1]))
    i = t[2, blk]
    per = t[0, blk, 1] - t[0, blk - 1, 1]
    print(blk, "|", *row[0], "|", *row[1], "|", i[1] - i[0], i[2] - i[1], i[3] - i[2], "|", i[5] - i[4], i[6] - i[5], i[7] - i[6], "|", per,
          "| abs S0", t[0, blk, 1] - t0, "S1", t[1, blk, 1] - t0)
