#!/bin/bash
# ring exchange experiments at N ranks: same per-rank block sizes as the 8-rank C5a run (16384 tokens per rank)
N=${1:-2}; TOK=$((16384 * N))
mkdir -p gpurun_out
run() {
  tag=$1; shift
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 \
      bench.py --gpus $N --workload c5a --ring-tokens $TOK --steps 6 --warmup 3 > gpurun_out/ring_$tag.json 2> gpurun_out/ring_$tag.err
  python - <<PY
import json
try:
    r = json.load(open('gpurun_out/ring_$tag.json'))['ring']
    print('$tag', 'ms', round(r['ms_per_step'], 3), 'compute-only', round(r['compute_only_ms_per_step'], 3), 'exposed %', round(r['exposed_comm_pct'], 2), 'TF/GPU', round(r['per_gpu_tflops'], 1))
except Exception as e:
    print('$tag', 'failed', e); print(open('gpurun_out/ring_$tag.err').read()[-1500:])
PY
}
shift
if [ $# -eq 0 ]; then set -- FOO=1 NCCL_P2P_USE_CUDA_MEMCPY=1; fi
for kv in "$@"; do run "$kv" $kv; done
