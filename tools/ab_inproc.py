"""In-process A/B of library builds: python tools/ab_inproc.py libA.so libB.so ...   ("" or "new" = the in-tree build).
The builds are loaded side by side and called alternately on the same tensors, so clock / power drift between runs
(+-5 % on this pool) hits all of them equally.  Reports the median forward / backward kernel time per case."""
import importlib, os, statistics, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
fa = importlib.import_module("xhy-flash-attention_b200")
capi = importlib.import_module("xhy-flash-attention_b200._capi")
default = capi.LIB_PATH
libs = {}
for a in sys.argv[1:] or ["new"]:
    path = default if a in ("", "new") else a
    capi._lib, capi.LIB_PATH = None, path
    libs[os.path.basename(path)[:24]] = capi.load()
torch.manual_seed(0)
dev, dt = "cuda", torch.bfloat16


def dense(b, s, causal=True, h=32, hk=8, d=128):
    q = torch.randn(b, s, h, d, device=dev, dtype=dt, requires_grad=True)
    k = torch.randn(b, s, hk, d, device=dev, dtype=dt, requires_grad=True)
    v = torch.randn(b, s, hk, d, device=dev, dtype=dt, requires_grad=True)
    g = torch.randn_like(q)
    def step():
        o = fa.flash_attn_func(q, k, v, causal=causal)
        torch.autograd.grad(o, (q, k, v), g)
    return step


def varlen(lens, h=32, hk=8):
    total = sum(lens)
    q = torch.randn(total, h, 128, device=dev, dtype=dt, requires_grad=True)
    k = torch.randn(total, hk, 128, device=dev, dtype=dt, requires_grad=True)
    v = torch.randn(total, hk, 128, device=dev, dtype=dt, requires_grad=True)
    g = torch.randn_like(q)
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), device=dev, dtype=torch.int32)
    def step():
        o = fa.flash_attn_varlen_func(q, k, v, cu, cu, max(lens), max(lens), 0.0, None, True)
        torch.autograd.grad(o, (q, k, v), g)
    return step


def fmdoc(lens, h=32, hk=8):
    total = sum(lens)
    q = torch.randn(1, total, h, 128, device=dev, dtype=dt, requires_grad=True)
    k = torch.randn(1, total, hk, 128, device=dev, dtype=dt, requires_grad=True)
    v = torch.randn(1, total, hk, 128, device=dev, dtype=dt, requires_grad=True)
    g = torch.randn_like(q)
    cu = torch.tensor(lens, device=dev).cumsum(0)
    ends = torch.repeat_interleave(cu, torch.tensor(lens, device=dev)).view(1, 1, total).int().contiguous()
    def step():
        o = fa.flash_attn_func(q, k, v, 0.0, None, True, flashmask=(ends, None, None, None))
        torch.autograd.grad(o, (q, k, v), g)
    return step


gen = torch.Generator().manual_seed(0)
rag = [int(x) for x in (128 + torch.rand(32, generator=gen) * 7800).round().tolist()]
cases = {"c3 b8 s8192": dense(8, 8192), "d64 b8 s8192 h32/8": dense(8, 8192, True, 32, 8, 64), "b64 s1024": dense(64, 1024), "b16 s4096": dense(16, 4096),
         "c2 b8 s2048 nc h16": dense(8, 2048, False, 16, 16), f"varlen 32 seqs {sum(rag)} tok": varlen(rag),
         f"fmdoc 32 docs {sum(rag)} tok": fmdoc(rag)}
only = os.environ.get("AB_CASES")
for name, step in cases.items():
    if only and not any(o in name for o in only.split(",")):
        continue
    res = {n: ([], []) for n in libs}
    for n, l in libs.items():
        capi._lib = l
        step(); step()
    for _ in range(int(os.environ.get("AB_ROUNDS", "12"))):
        for n, l in libs.items():
            capi._lib = l
            l.fa_b200_profile_enable(1)
            step(); torch.cuda.synchronize()
            res[n][0].append(l.fa_b200_profile_last_ms(0)); res[n][1].append(l.fa_b200_profile_last_ms(1))
            l.fa_b200_profile_enable(0)
    base = None
    for n, (f, b) in res.items():
        fm, bm = statistics.median(f), statistics.median(b)
        base = base or (fm, bm)
        print(f"{name:34s} {n:26s} fwd {fm:7.3f} ms ({100 * (fm / base[0] - 1):+5.1f} %)  bwd {bm:7.3f} ms ({100 * (bm / base[1] - 1):+5.1f} %)", flush=True)
