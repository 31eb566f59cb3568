#!/bin/bash
# multi-GPU pass: ring test over NCCL, ring bench, batch x kv-head partition bench.  usage: gpu_multi.sh N
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi_multi.txt
timeout 600 python -m pytest tests/test_ring_gpu.py -m gpu -x -q > gpurun_out/pytest_ring_n$N.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_ring_n$N.log
tail -5 gpurun_out/pytest_ring_n$N.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus $N --workload c5a --steps 5 --warmup 3 > gpurun_out/bench_ring_n$N.json 2> gpurun_out/bench_ring_n$N.err; echo "ring exit $?"; tail -3 gpurun_out/bench_ring_n$N.err; cat gpurun_out/bench_ring_n$N.json
timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_c3_n$N.json 2> gpurun_out/bench_c3_n$N.err; echo "c3 exit $?"; tail -3 gpurun_out/bench_c3_n$N.err; cat gpurun_out/bench_c3_n$N.json | cut -c1-400
