"""Where does a ring step wait?  torchrun ... tools/ring_trace.py [tokens]: per rank, the time the compute stream spends in
every wait for a transfer (CUDA events around the wait), for one traced step after warm-up."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
ring = importlib.import_module("xhy-flash-attention_b200.ring_attention")
total = int(sys.argv[1]) if len(sys.argv) > 1 else 16384 * world
s_local, h, hk, d = total // world, 32, 8, 128
torch.manual_seed(rank)
q = torch.randn(1, s_local, h, d, device=dev, dtype=torch.bfloat16, requires_grad=True)
k = torch.randn(1, s_local, hk, d, device=dev, dtype=torch.bfloat16, requires_grad=True)
v = torch.randn(1, s_local, hk, d, device=dev, dtype=torch.bfloat16, requires_grad=True)
do = torch.randn(1, s_local, h, d, device=dev, dtype=torch.bfloat16)

def step():
    out = ring.ring_flash_attn_func(q, k, v, causal=True)
    return torch.autograd.grad(out, (q, k, v), do)

for _ in range(3):
    step()
for rep in range(2):
    dist.barrier(); torch.cuda.synchronize()
    ring.STALL_TRACE = []
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record(); step(); t1.record(); torch.cuda.synchronize()
    tr, ring.STALL_TRACE = ring.STALL_TRACE, None
    stalls = [(tag, a.elapsed_time(b)) for tag, a, b in tr]
    tot = sum(x for _, x in stalls)
    line = f"rank {rank} rep {rep}: step {t0.elapsed_time(t1):.2f} ms, waits {tot:.2f} ms: " + " ".join(f"[{t} {x:.2f}]" for t, x in stalls if x > 0.05)
    for r in range(world):
        dist.barrier()
        if r == rank:
            print(line, flush=True)
dist.destroy_process_group()
