#!/bin/bash
# oracle/build_ref.sh -- compiles the REFERENCE's own FA2 kernel path for sm_100 into oracle/_ref/libflashattn_ref.so.
#
# Test infrastructure only (same-box comparator of tests/ and tools/): nothing under xhy-flash-attention_b200/ loads it.
# The sources stay where they lie under /root/reference; only outputs go to oracle/_ref/ (git-ignored, NOT
# gpurun-ignored, so the built library travels to the GPU box).  No reference build system is run: this is the list
# of nvcc commands csrc/CMakeLists.txt:37-66,108-122 boils down to for the translation units the comparator needs --
#   csrc/capi/flash_attn.cu                      the C ABI (flash_attn_fwd / flash_attn_bwd / calc_reduced_attn_scores ...)
#   csrc/flash_attn/src/cuda_utils.cu            its device-property cache
#   flash_{fwd,bwd}_hdim128_bf16{,_causal}{,_flashmask}_sm80.cu   the instantiation stubs the reference's own
#                                                generator (csrc/flash_attn/src/generate_kernels.py) writes
#   calc_reduced_attn_scores_dispatch/hdim128_bf16_sm80.cu
# Every other run_mha_* specialisation stays an undefined symbol of the shared library (lazy binding: only the
# specialisations above may be called; the comparator sticks to head dim 128, bf16).
# ~3 min of nvcc per kernel TU; 8 in parallel.
set -e
REF=${REF:-/root/reference}
HERE=$(cd "$(dirname "$0")" && pwd)
OUT=$HERE/_ref
[ -d "$REF/csrc/capi" ] || { echo "no reference tree at $REF: keeping whatever is in $OUT"; exit 0; }
mkdir -p "$OUT/gen" "$OUT/obj"
# the generator rewrites every stub on each run: only run it when the stubs are missing, so objects stay cached
[ -f "$OUT/gen/flash_fwd_hdim128_bf16_causal_sm80.cu" ] || python "$REF/csrc/flash_attn/src/generate_kernels.py" -o "$OUT/gen"
S=$REF/csrc
FLAGS="-w -Xcompiler -fPIC -Xcompiler -O3 -std=c++17 -U__CUDA_NO_HALF_OPERATORS__ -U__CUDA_NO_HALF_CONVERSIONS__ \
 -U__CUDA_NO_HALF2_OPERATORS__ -U__CUDA_NO_BFLOAT16_CONVERSIONS__ --expt-relaxed-constexpr --expt-extended-lambda \
 --use_fast_math -gencode arch=compute_100,code=sm_100 -I $S/flash_attn/src -I $S/flash_attn -I $S/cutlass/include \
 -I $S/capi -I $S/flash_attn_with_bias_and_mask -I $S/flash_attn_with_bias_and_mask/src -I $S"
TUS="$S/capi/flash_attn.cu $S/flash_attn/src/cuda_utils.cu $S/flash_attn/src/calc_reduced_attn_scores_dispatch/hdim128_bf16_sm80.cu"
for dir in fwd bwd; do
  for sw in "" _causal _flashmask _causal_flashmask; do
    TUS="$TUS $OUT/gen/flash_${dir}_hdim128_bf16${sw}_sm80.cu"
  done
done
pids=""
for f in $TUS; do
  o=$OUT/obj/$(basename "$f" .cu).o
  if [ ! -f "$o" ] || [ "$f" -nt "$o" ]; then
    ( nvcc $FLAGS -c "$f" -o "$o" 2> "$o.log" || { echo "FAILED $f"; tail -5 "$o.log"; rm -f "$o"; } ) &
    pids="$pids $!"
  fi
done
for p in $pids; do wait $p; done
nvcc -shared -o "$OUT/libflashattn_ref.so" $OUT/obj/*.o -lcudart
ls -la "$OUT/libflashattn_ref.so"
