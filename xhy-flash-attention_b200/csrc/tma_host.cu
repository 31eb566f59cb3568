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
  CUresult r = encode_fn()(&m, d.elem_bytes == 1 ? CU_TENSOR_MAP_DATA_TYPE_UINT8 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void *>(d.ptr), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    throw Error("cuTensorMapEncodeTiled failed with code " + std::to_string((int)r) + " (dims " +
                std::to_string(d.dims[0]) + "," + std::to_string(d.dims[1]) + "," + std::to_string(d.dims[2]) + "," +
                std::to_string(d.dims[3]) + "; strides " + std::to_string(d.strides[0]) + "," +
                std::to_string(d.strides[1]) + "," + std::to_string(d.strides[2]) + ")");
  return m;
}

// ---- tensor maps whose row extent is only known on the device ----------------------------------------------------
// The flat varlen symbols (flash_attn_varlen_fwd / _bwd) carry cu_seqlens but no total row counts, and the row extent
// of a packed tensor's map must be exact (the tail tile of the last sequence relies on the out-of-bounds zero fill).
// Instead of reading cu_seqlens[b] back (a stream synchronisation on every call, impossible under graph capture),
// a one-warp kernel copies the host-encoded maps into global memory and patches their row extent there:
// tensormap.replace in shared memory, then tensormap.cp_fenceproxy to global (the scheme CUTLASS uses for grouped GEMM).
// The attention kernels read such maps through a pointer after a fence.proxy.tensormap acquire.
struct PatchArgs {
  CUtensorMap maps[4];
  int n;
  int k_like[4];     // 0: rows = cu_seqlens_q[b], 1: rows = cu_seqlens_k[b]
};

__global__ void __launch_bounds__(32) patch_maps_kernel(const __grid_constant__ PatchArgs a, CUtensorMap *dst,
                                                        const int32_t *cu_q, const int32_t *cu_k, int b) {
  __shared__ __align__(128) CUtensorMap sm[4];
  const int lane = threadIdx.x;
  for (int i = 0; i < a.n; ++i)
    reinterpret_cast<uint32_t *>(&sm[i])[lane] = reinterpret_cast<const uint32_t *>(&a.maps[i])[lane];
  __syncwarp();
  for (int i = 0; i < a.n; ++i) {
    int rows = a.k_like[i] ? cu_k[b] : cu_q[b];
    rows = rows < 1 ? 1 : rows;
    const uint32_t sa = (uint32_t)__cvta_generic_to_shared(&sm[i]);
    if (lane == 0)
      asm volatile("tensormap.replace.tile.global_dim.shared::cta.b1024.b32 [%0], 2, %1;" ::"r"(sa), "r"(rows) : "memory");
    __syncwarp();
    asm volatile("tensormap.cp_fenceproxy.global.shared::cta.tensormap::generic.release.gpu.sync.aligned [%0], [%1], 128;"
                 ::"l"(dst + i), "r"(sa) : "memory");
  }
}

// Returns a stream-ordered device allocation holding n patched maps (free it with cudaFreeAsync after the consumer).
CUtensorMap *patch_tma_rows(const CUtensorMap *maps, const int *k_like, int n, const int32_t *cu_q, const int32_t *cu_k,
                            int b, cudaStream_t s) {
  FA_CHECK(n >= 1 && n <= 4, "patch_tma_rows: 1..4 maps");
  PatchArgs a;
  a.n = n;
  for (int i = 0; i < n; ++i) { a.maps[i] = maps[i]; a.k_like[i] = k_like[i]; }
  for (int i = n; i < 4; ++i) { a.maps[i] = maps[0]; a.k_like[i] = 0; }
  CUtensorMap *dst = nullptr;
  FA_CUDA(cudaMallocAsync(&dst, 4 * sizeof(CUtensorMap), s));
  patch_maps_kernel<<<1, 32, 0, s>>>(a, dst, cu_q, cu_k, b);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { cudaFreeAsync(dst, s); FA_CUDA(e); }
  note_launch(2);
  return dst;
}

}  // namespace fa
