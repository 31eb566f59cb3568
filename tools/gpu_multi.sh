#!/bin/bash
# N-GPU pass (gpurun --gpus N): ring + sharded tests over NCCL, then the default bench line at N ranks (c3_strong, c4_sharded, ring extras)
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi_multi.txt 2>&1
timeout 600 python -m pytest tests/test_ring_gpu.py tests/test_varlen_shard_gpu.py -m gpu -q --timeout 300 > gpurun_out/pytest_ring_n$N.log 2>&1; echo "pytest ring exit $?"
tail -3 gpurun_out/pytest_ring_n$N.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_c3_n$N.json 2> gpurun_out/bench_c3_n$N.err; echo "bench exit $?"
tail -3 gpurun_out/bench_c3_n$N.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err
python - <<PY
import json
d=json.load(open('gpurun_out/bench_c3_n$N.json'))
print('c3 n$N', round(d['value'],1), round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), d['e2e'].get('host_bw'), 'frac_floor', d['e2e'].get('frac_of_copy_floor'))
for k in ('c3_strong','c4_sharded','ring'):
    print(k, json.dumps(d.get(k))[:900])
r=json.load(open('gpurun_out/bench_ref_n$N.json'))
print('ref', r['value'], r['cpu_baseline']['cores'])
PY
