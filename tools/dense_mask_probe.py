import importlib, sys, os, ctypes as C, torch
sys.path.insert(0, os.getcwd())
capi = importlib.import_module("xhy-flash-attention_b200._capi"); lib = capi.load()
itf = importlib.import_module("xhy-flash-attention_b200.flash_attn_interface")
b, s, h, hk, d = 8, 8192, 32, 8, 128
q = torch.randn(b, s, h, d, device="cuda", dtype=torch.bfloat16)
k = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16)
v = torch.randn(b, s, hk, d, device="cuda", dtype=torch.bfloat16)
g = torch.randn_like(q)
mask = torch.zeros(b, 1, s, s, device="cuda", dtype=torch.bfloat16)
mask.masked_fill_(torch.triu(torch.ones(s, s, device="cuda", dtype=torch.bool), 1), float("-inf"))   # causal as a dense mask
p = capi.new_params(q.dtype)
out = torch.empty_like(q); lse = torch.empty(b, h, s, device="cuda", dtype=torch.float32)
dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
sd = torch.empty(b, h, s, device="cuda", dtype=torch.float32); acc = torch.empty(b * h * s * d, device="cuda", dtype=torch.float32)
p.q, p.k, p.v, p.o, p.dout, p.dq, p.dk, p.dv = [capi.tdesc(t) for t in (q, k, v, out, g, dq, dk, dv)]
p.softmax_lse, p.softmax_d, p.dq_accum = lse.data_ptr(), sd.data_ptr(), acc.data_ptr()
p.b, p.seqlen_q, p.seqlen_k, p.h, p.h_k, p.d = b, s, s, h, hk, d
p.seqlen_q_rounded, p.seqlen_k_rounded, p.d_rounded = s, s, d
p.softmax_scale = d ** -0.5
p.attn_mask, p.mask_head_mod, p.mask_seq_mod, p.unscale_softmax = mask.data_ptr(), 1, s, d ** 0.5
st = capi.stream_ptr(q.device)
fl = 4.0 * b * s * s * h * d / 2
for _ in range(2):
    assert lib.fa_b200_fwd(p, st), capi.last_error(); assert lib.fa_b200_bwd(p, st), capi.last_error()
lib.fa_b200_profile_enable(1)
lib.fa_b200_fwd(p, st); lib.fa_b200_bwd(p, st); torch.cuda.synchronize()
f, bw = lib.fa_b200_profile_last_ms(0), lib.fa_b200_profile_last_ms(1)
print(f"C3 causal as a DENSE bf16 mask (b,1,s,s): fwd {f:.2f} ms ({fl/f/1e9:.0f} TF on the causal half), bwd {bw:.2f} ms ({2.5*fl/bw/1e9:.0f} TF), kind {capi.last_kernel_kind()}")
