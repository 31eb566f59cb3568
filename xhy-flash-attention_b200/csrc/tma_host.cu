// tma_host.cu -- host-side CUtensorMap encoding.  cuTensorMapEncodeTiled is fetched through the runtime
// (cudaGetDriverEntryPoint) so the library links against libcudart only.
#include <mutex>

#include "common.h"
#include "sm100.cuh"

namespace fa {

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
    if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) fn = (EncodeTiledFn)p;
  });
  if (!fn) throw Error("cuTensorMapEncodeTiled is not available from the CUDA driver");
  return fn;
}

CUtensorMap make_tma_4d(const TmaDesc4 &d) {
  CUtensorMap m;
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {  // autograd worker threads start without a current driver context
    FA_CUDA(cudaFree(nullptr));
    ctx_bound = true;
  }
  cuuint64_t dims[4] = {d.dims[0], d.dims[1], d.dims[2], d.dims[3]};
  cuuint64_t strides[3] = {d.strides[0], d.strides[1], d.strides[2]};
  cuuint32_t box[4] = {d.box[0], d.box[1], d.box[2], d.box[3]};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = encode_fn()(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void *>(d.ptr), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    throw Error("cuTensorMapEncodeTiled failed with code " + std::to_string((int)r) + " (dims " +
                std::to_string(d.dims[0]) + "," + std::to_string(d.dims[1]) + "," + std::to_string(d.dims[2]) + "," +
                std::to_string(d.dims[3]) + "; strides " + std::to_string(d.strides[0]) + "," +
                std::to_string(d.strides[1]) + "," + std::to_string(d.strides[2]) + ")");
  return m;
}

}  // namespace fa
