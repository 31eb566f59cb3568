#!/bin/bash
# compute-sanitizer over tiny invocations of every kernel family (memcheck on all; synccheck and racecheck on the
# tcgen05 training kernels and the decode kernel).  gpurun -- 'bash tools/sanitize.sh'
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
python tools/all_kernels_case.py > gpurun_out/san_plain.log 2>&1 || { echo "plain run failed"; tail gpurun_out/san_plain.log; exit 1; }
for part in train flashmask decode generic; do
  timeout 900 $CS --tool memcheck --print-limit 20 python tools/all_kernels_case.py $part > gpurun_out/san_memcheck_$part.log 2>&1
  echo "memcheck $part exit $? : $(grep -E 'ERROR SUMMARY' gpurun_out/san_memcheck_$part.log)"
done
for tool in synccheck racecheck; do
  for part in train decode; do
    timeout 900 $CS --tool $tool --print-limit 20 python tools/all_kernels_case.py $part > gpurun_out/san_${tool}_$part.log 2>&1
    echo "$tool $part exit $? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' gpurun_out/san_${tool}_$part.log)"
  done
done
