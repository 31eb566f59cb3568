"""Same-box yardstick (SURVEY.md §8(d) "reference kernel timed beside it"): the stock upstream `flash_attn` wheel that
ships in the image (FA2 mma.sync kernels recompiled for sm_100 -- the kind of kernel this repo is meant to beat) timed
in the same process, on the same tensors, alternating with this repo's kernels.  A comparator only: nothing in the
product path imports it.
    python tools/yardstick_upstream.py [out.json]
Prints one line per workload (C3, C2 fwd+bwd; C5b decode; C4 varlen fwd+bwd) and the max abs difference between the two outputs."""
import importlib, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
ours = importlib.import_module("xhy-flash-attention_b200")
import flash_attn as up  # upstream wheel (site-packages)

assert "site-packages" in up.__file__, up.__file__
torch.manual_seed(0)
dev, dt = "cuda", torch.bfloat16


def timed(fn, warm=3, reps=10):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


res = {"upstream_version": up.__version__, "device": torch.cuda.get_device_name(0), "workloads": {}}
for name, (b, s, h, hk, causal) in {"c3": (8, 8192, 32, 8, True), "c2": (8, 2048, 16, 16, False)}.items():
    q = torch.randn(b, s, h, 128, device=dev, dtype=dt, requires_grad=True)
    k = torch.randn(b, s, hk, 128, device=dev, dtype=dt, requires_grad=True)
    v = torch.randn(b, s, hk, 128, device=dev, dtype=dt, requires_grad=True)
    g = torch.randn_like(q)
    fl = 4.0 * b * s * s * h * 128 / (2 if causal else 1)
    row = {}
    outs = {}
    for tag, mod in (("ours", ours), ("upstream", up), ("ours_again", ours)):
        def fwd():
            return mod.flash_attn_func(q, k, v, causal=causal)
        def step():
            o = fwd()
            return torch.autograd.grad(o, (q, k, v), g)
        tf = timed(lambda: fwd())
        ts = timed(step)
        outs[tag] = (fwd().detach().float(), *[x.float() for x in step()])
        row[tag] = {"fwd_ms": tf, "fwd_tflops": fl / tf / 1e9, "fwd_bwd_ms": ts, "fwd_bwd_tflops": 3.5 * fl / ts / 1e9}
    row["max_abs_diff_out_dq_dk_dv"] = [float((a - c).abs().max()) for a, c in zip(outs["ours"], outs["upstream"])]
    row["speedup_fwd_bwd"] = row["upstream"]["fwd_bwd_ms"] / row["ours"]["fwd_bwd_ms"]
    res["workloads"][name] = row
    print(name, json.dumps(row), flush=True)
    del q, k, v, g, outs
    torch.cuda.empty_cache()

# C5b decode: b64, cache 32768, hq32 / hkv8, one query token
b, ctx, h, hk = 64, 32768, 32, 8
q = torch.randn(b, 1, h, 128, device=dev, dtype=dt)
kc = torch.randn(b, ctx, hk, 128, device=dev, dtype=dt)
vc = torch.randn(b, ctx, hk, 128, device=dev, dtype=dt)
lens = torch.full((b,), ctx, device=dev, dtype=torch.int32)
byts = 2.0 * b * ctx * hk * 128 * 2
row = {}
outs = {}
for tag, mod in (("ours", ours), ("upstream", up), ("ours_again", ours)):
    fn = lambda: mod.flash_attn_with_kvcache(q, kc, vc, cache_seqlens=lens, causal=True)
    t = timed(fn, warm=3, reps=20)
    outs[tag] = fn().float()
    row[tag] = {"ms": t, "gbs": byts / t / 1e6, "tok_s": b / t * 1e3}
row["max_abs_diff_out"] = float((outs["ours"] - outs["upstream"]).abs().max())
row["speedup"] = row["upstream"]["ms"] / row["ours"]["ms"]
res["workloads"]["c5b_decode"] = row
print("c5b", json.dumps(row), flush=True)
del q, kc, vc, outs
torch.cuda.empty_cache()

# C4 varlen: 64 sequences (128..16k tokens, 262144 in total), cu_seqlens encoding, causal (lengths from bench.c4_lengths)
from bench import c4_lengths
lens_l = c4_lengths()
total, h, hk = sum(lens_l), 32, 8
fl = 3.5 * 4.0 * h * 128 * sum(n * (n + 1) // 2 for n in lens_l)
q = torch.randn(total, h, 128, device=dev, dtype=dt, requires_grad=True)
k = torch.randn(total, hk, 128, device=dev, dtype=dt, requires_grad=True)
v = torch.randn(total, hk, 128, device=dev, dtype=dt, requires_grad=True)
g = torch.randn_like(q)
cu = torch.tensor([0] + list(torch.tensor(lens_l).cumsum(0)), device=dev, dtype=torch.int32)
row, outs = {}, {}
for tag, mod in (("ours", ours), ("upstream", up), ("ours_again", ours)):
    def step():
        o = mod.flash_attn_varlen_func(q, k, v, cu, cu, max(lens_l), max(lens_l), 0.0, None, True)
        return (o.detach(), *torch.autograd.grad(o, (q, k, v), g))
    t = timed(step, warm=3, reps=5)
    outs[tag] = [x.float() for x in step()]
    row[tag] = {"fwd_bwd_ms": t, "fwd_bwd_tflops": fl / t / 1e9}
row["max_abs_diff_out_dq_dk_dv"] = [float((a - c).abs().max()) for a, c in zip(outs["ours"], outs["upstream"])]
row["speedup_fwd_bwd"] = row["upstream"]["fwd_bwd_ms"] / row["ours"]["fwd_bwd_ms"]
res["workloads"]["c4_varlen"] = row
print("c4", json.dumps(row), flush=True)
if len(sys.argv) > 1:
    json.dump(res, open(sys.argv[1], "w"), indent=1)
