"""Timing probe for the FlashMask forward: python tools/fm_probe.py [case]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
fa = importlib.import_module("xhy-flash-attention_b200")
capi = importlib.import_module("xhy-flash-attention_b200._capi")
lib = capi.load()
dev = "cuda"
torch.manual_seed(0)

def timeit(fn, n=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    lib.fa_b200_profile_enable(1)
    ms = []
    for _ in range(n):
        fn(); torch.cuda.synchronize(); ms.append(lib.fa_b200_profile_last_ms(0))
    lib.fa_b200_profile_enable(0)
    return sum(ms) / len(ms)

def case(b, s, h, hk, doc_lens=None, label=""):
    q = torch.randn(b, s, h, 128, device=dev, dtype=torch.bfloat16)
    k = torch.randn(b, s, hk, 128, device=dev, dtype=torch.bfloat16)
    v = torch.randn(b, s, hk, 128, device=dev, dtype=torch.bfloat16)
    t0 = timeit(lambda: fa.flash_attn_func(q, k, v, causal=True))
    full = torch.full((b, 1, s), s, device=dev, dtype=torch.int32)
    t1 = timeit(lambda: fa.flash_attn_func(q, k, v, causal=True, flashmask=(full, None, None, None)))
    out = f"{label}: causal {t0:.3f} ms | FM all-clear {t1:.3f} ms"
    if doc_lens is not None:
        ends = torch.repeat_interleave(torch.tensor(doc_lens).cumsum(0), torch.tensor(doc_lens)).view(1, 1, s).int().to(dev).expand(b, 1, s).contiguous()
        t2 = timeit(lambda: fa.flash_attn_func(q, k, v, causal=True, flashmask=(ends, None, None, None)))
        pairs = sum(n * (n + 1) // 2 for n in doc_lens)
        out += f" | FM docs {t2:.3f} ms ({4 * b * h * 128 * pairs / t2 / 1e9:.0f} TFLOPS on unmasked pairs)"
    print(out, flush=True)

which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which in ("all", "c3"):
    case(8, 8192, 32, 8, [2048] * 4, "b8 s8192, docs 4x2048")
    case(8, 8192, 32, 8, [4096, 1000, 3096], "b8 s8192, docs 4096/1000/3096")
if which in ("all", "long"):
    case(1, 65536, 32, 8, [4096] * 16, "b1 s65536, docs 16x4096")
    case(1, 65536, 32, 8, [4000, 4192] * 8, "b1 s65536, docs unaligned")
if which in ("all", "ncu"):
    case(1, 65536, 32, 8, [4000, 4192] * 8, "ncu case")
