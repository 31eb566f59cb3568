// noalloc.cu -- fa_b200_fwd_splitkv_noalloc: split-KV decode + LSE combine with the split partials in a file-scope
// __device__ array instead of caller- or allocator-provided scratch.
//
// No counterpart in the reference; added for callers that cannot allocate device
// memory (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The existing pair
// fa_b200_fwd_splitkv + fa_b200_fwd_combine wants two fp32 scratch tensors from its caller -- oaccum (splits, b, sq, h, d)
// and lseaccum (splits, b, h, sq); this entry point takes the same params struct with those two pointers NULL and puts
// the partials in `g_splitkv_scratch` below.  It calls no cudaMalloc / cudaMallocAsync / cudaFree and no allocating
// library routine: the kernels it launches (decode.cu or the generic split kernel, aux_kernels.cu's combine) take
// their tensor maps as kernel arguments and need nothing else.
//
// Limits and sharing:
//   * the array is FA_B200_NOALLOC_SCRATCH_BYTES (16 MiB) per GPU that loads the library, held until the process exits;
//     a call whose partials do not fit is rejected with an error naming both sizes (BASELINE config 5b -- b64, 32 q heads,
//     d128, 4 splits -- needs 4.2 MiB);
//   * every call on a device shares the one array: calls that overlap on the GPU (two streams, two threads) RACE; calls
//     on one stream are ordered and safe;
//   * the array keeps the last call's contents.  Nothing in it has to start at zero: the split kernels write every
//     (split, batch, row, head) partial of the call -- an empty split writes lse = -inf, which the combine skips -- before
//     the combine reads them, so stale values are never read (tests/test_zz_no_alloc.py runs two different problems back
//     to back and compares each with the allocating path).
#include <mutex>

#include "common.h"
#include "../../include/fa_capi.h"

namespace fa {
namespace {

// partials of one call: [oaccum: splits * b * sq * h * d floats][lseaccum: splits * b * h * sq floats]
static __device__ __align__(256) float g_splitkv_scratch[FA_B200_NOALLOC_SCRATCH_BYTES / sizeof(float)];

// In host code `g_splitkv_scratch` is a host-side stand-in for the symbol, not a device address: the device address is
// asked for once per device, with that device current, and cached per device ordinal.
float *scratch_device_address() {
  static std::mutex mu;
  static float *addr[64] = {};
  int dev = 0;
  FA_CUDA(cudaGetDevice(&dev));
  FA_CHECK(dev >= 0 && dev < 64, "device ordinal out of range");
  std::lock_guard<std::mutex> g(mu);
  if (addr[dev] == nullptr) {
    void *ptr = nullptr;
    FA_CUDA(cudaGetSymbolAddress(&ptr, g_splitkv_scratch));
    FA_CHECK(ptr != nullptr, "cudaGetSymbolAddress returned a null address for the split-KV scratch array");
    addr[dev] = static_cast<float *>(ptr);
  }
  return addr[dev];
}

}  // namespace

void run_splitkv_noalloc(const Params &caller, cudaStream_t s) {
  FA_CHECK(caller.oaccum == nullptr && caller.lseaccum == nullptr,
           "fa_b200_fwd_splitkv_noalloc: oaccum / lseaccum must be NULL (the partials live in the library's static array)");
  FA_CHECK(caller.o.ptr && caller.softmax_lse, "out and softmax_lse must be set");
  // paths that allocate stream-ordered scratch of their own are outside this entry point
  FA_CHECK(caller.cu_seqlens_q == nullptr && caller.cu_seqlens_k == nullptr,
           "fa_b200_fwd_splitkv_noalloc: packed (cu_seqlens) batches are not supported; use seqused_k for ragged caches");
  FA_CHECK(caller.lt_start == nullptr && caller.num_splits_dynamic == nullptr && !caller.is_e4m3,
           "fa_b200_fwd_splitkv_noalloc: FlashMask, per-batch dynamic splits and FP8 are not supported");
  Params p = caller;
  if (p.num_splits <= 1) {      // nothing to combine, no scratch: same as fa_b200_fwd_splitkv
    run_splitkv(p, s);
    return;
  }
  const size_t rows = (size_t)p.num_splits * p.b * p.seqlen_q * p.h;
  const size_t need = rows * ((size_t)p.d + 1) * sizeof(float);
  if (need > (size_t)FA_B200_NOALLOC_SCRATCH_BYTES)
    throw Error("fa_b200_fwd_splitkv_noalloc: the split partials need " + std::to_string(need) +
                " bytes (num_splits * batch * seqlen_q * heads * (head_size + 1) * 4) but the static scratch array holds " +
                std::to_string((size_t)FA_B200_NOALLOC_SCRATCH_BYTES) +
                " (FA_B200_NOALLOC_SCRATCH_BYTES); lower num_splits or use fa_b200_fwd_splitkv with caller-owned scratch");
  float *base = scratch_device_address();
  p.oaccum = base;
  p.lseaccum = base + rows * (size_t)p.d;     // rows * d floats: a multiple of 8 floats since d % 8 == 0
  run_splitkv(p, s);
  combine_fwd(p, s);
}

}  // namespace fa

extern "C" bool fa_b200_fwd_splitkv_noalloc(const fa_b200_params *p, cudaStream_t stream) {
  try {
    FA_CHECK(p != nullptr, "null params");
    fa::run_splitkv_noalloc(*p, stream);
  } catch (const std::exception &e) {
    flash_attn_set_error(e.what());
    return false;
  } catch (...) {
    flash_attn_set_error(nullptr);
    return false;
  }
  return true;
}
