"""Small invocations of every kernel family, one after the other (a quick "does every path launch" run; also usable under a debugger or profiler):
    python tools/all_kernels_case.py [train|flashmask|decode|generic]
Shapes are tiny but hit: tcgen05 fwd/bwd (dense causal GQA, d64, ragged
lengths, varlen, FlashMask with skipped + partial blocks), decode kernel (contiguous, paged, leftpad, splits + combine),
generic kernels (d 256, dropout), rotary + append, LSE merge."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
fa = importlib.import_module("xhy-flash-attention_b200")
torch.manual_seed(0)
dev, dt = "cuda", torch.bfloat16


def fwd_bwd(fn, *tensors, **kw):
    ts = [t.requires_grad_(True) for t in tensors]
    o = fn(*ts, **kw)
    torch.autograd.grad(o, ts, torch.randn_like(o))


def r(*shape, dtype=dt):
    return torch.randn(*shape, device=dev, dtype=dtype)


which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which in ("all", "train"):
    fwd_bwd(fa.flash_attn_func, r(2, 384, 8, 128), r(2, 384, 2, 128), r(2, 384, 2, 128), causal=True)
    fwd_bwd(fa.flash_attn_func, r(1, 200, 4, 64), r(1, 333, 4, 64), r(1, 333, 4, 64), causal=False)
    fwd_bwd(fa.flash_attn_func, r(1, 130, 2, 96, dtype=torch.float16), r(1, 130, 2, 96, dtype=torch.float16),
            r(1, 130, 2, 96, dtype=torch.float16), causal=True)
    cu = torch.tensor([0, 100, 356, 640], device=dev, dtype=torch.int32)
    fwd_bwd(fa.flash_attn_varlen_func, r(640, 4, 128), r(640, 2, 128), r(640, 2, 128), cu_seqlens_q=cu, cu_seqlens_k=cu,
            max_seqlen_q=284, max_seqlen_k=284, causal=True)
    print("train ok")
if which in ("all", "flashmask"):
    s = 1024
    ends = torch.tensor([300, 400, 900, 1024]).cumsum(0) * 0 + torch.tensor([300, 400, 900, 1024])
    lts = torch.cat([torch.full((n,), e) for n, e in zip([300, 100, 500, 124], ends.tolist())]).view(1, 1, s, 1).to(dev).to(torch.int32)
    fwd_bwd(lambda q, k, v: fa.flashmask_attention(q, k, v, lts, causal=True), r(1, s, 4, 128), r(1, s, 2, 128), r(1, s, 2, 128))
    print("flashmask ok")
if which in ("all", "decode"):
    q, kc, vc = r(3, 1, 16, 128), r(3, 1000, 4, 128), r(3, 1000, 4, 128)
    lens = torch.tensor([1000, 1, 513], device=dev, dtype=torch.int32)
    for ns in (1, 4):
        fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, causal=True, num_splits=ns)
    fa.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, cache_leftpad=torch.tensor([3, 0, 70], device=dev, dtype=torch.int32),
                               causal=True, softcap=20.0, alibi_slopes=torch.rand(16, device=dev))
    pool_k, pool_v = r(20, 64, 4, 128), r(20, 64, 4, 128)
    table = torch.randperm(20, device=dev)[:18].view(3, 6).to(torch.int32)
    ang = torch.rand(384, 32, device=dev)
    fa.flash_attn_with_kvcache(q, pool_k, pool_v, k=r(3, 1, 4, 128), v=r(3, 1, 4, 128), rotary_cos=torch.cos(ang).to(dt),
                               rotary_sin=torch.sin(ang).to(dt), cache_seqlens=torch.tensor([383, 0, 64], device=dev, dtype=torch.int32),
                               block_table=table, causal=True, num_splits=2)
    print("decode ok")
if which in ("all", "generic"):
    fwd_bwd(fa.flash_attn_func, r(1, 70, 2, 256), r(1, 90, 1, 256), r(1, 90, 1, 256), causal=True)
    fwd_bwd(fa.flash_attn_func, r(1, 128, 2, 64), r(1, 128, 2, 64), r(1, 128, 2, 64), dropout_p=0.2, causal=True)
    fa.flash_attn_with_kvcache(r(2, 20, 8, 128), r(2, 300, 2, 128), r(2, 300, 2, 128), cache_seqlens=300, causal=True, num_splits=3)
    print("generic ok")
torch.cuda.synchronize()
print("done")
