// handle_abi.cu -- the setter-style params handles of the reference's FlashMask-v2 / FA3 libraries (include/fa_handle.h).
//
// Replaces csrc/flashmask_v2/flash_api.{h,cu} and csrc/flash_attn_v3/flash_api.{h,cu}: the handle is a heap object with
// one slot per field of the reference's Flash_fwd_params / Flash_bwd_params (csrc/flashmask_v2/flash.h:13-231); the run
// functions translate the slots into fa_b200_params and take the same dispatchers as every other entry point.
#include <cstdlib>
#include <cstring>
#include <new>

#include "common.h"
#include "../../include/fa_handle.h"

#define FA_SLOT(type, member) type member;
struct Flash_fwd_params {
  FA_HANDLE_FWD_FIELDS(FA_SLOT)
  FA_HANDLE_MASK_FIELDS(FA_SLOT)
};
struct Flash_bwd_params : public Flash_fwd_params {
  FA_HANDLE_BWD_FIELDS(FA_SLOT)
  FA_HANDLE_BWD_MASK_FIELDS(FA_SLOT)
};
#undef FA_SLOT

namespace {

using fa::Params;

void require(bool ok, const char *what) {
  if (!ok) throw fa::Error(std::string("params handle: ") + what);
}

fa_b200_tensor tens(void *ptr, int64_t bs, int64_t rs, int64_t hs) {
  fa_b200_tensor t;
  t.ptr = ptr; t.batch_stride = bs; t.row_stride = rs; t.head_stride = hs;
  return t;
}

bool decode_shape(const Flash_fwd_params &h) {   // seqlen_q is the maximum over the batch when cu_seqlens_q is given
  return h.lt_start_ptr == nullptr && h.h_k > 0 && h.seqlen_q * (h.h / h.h_k) <= 128 && (h.d == 64 || h.d == 128);
}

// (h, total_q) <-> (b, h, max_seqlen_q): the v2/v3 surface keeps varlen LSE packed (flash_fwd_launch_template.h:161-164)
__global__ void lse_repack_kernel(float *packed, float *padded, const int *cu, int b, int h, int max_sq, int64_t total_q,
                                  int to_packed) {
  const int bi = blockIdx.z, hq = blockIdx.y;
  const int row = blockIdx.x * blockDim.x + threadIdx.x;
  const int r0 = cu[bi], n = cu[bi + 1] - r0;
  if (row >= n) return;
  float *a = packed + (int64_t)hq * total_q + r0 + row;
  float *c = padded + ((int64_t)bi * h + hq) * max_sq + row;
  if (to_packed) *a = *c; else *c = *a;
}

Params common_params(const Flash_fwd_params &h) {
  require(!h.is_fp32, "fp32 inputs are not supported");
  require(h.softcap >= 0.f, "softcap must be >= 0");
  require(h.qv_ptr == nullptr, "qv is not supported");
  require(h.cu_seqlens_knew == nullptr, "cu_seqlens_knew is not supported");
  require((h.knew_ptr == nullptr) == (h.vnew_ptr == nullptr), "knew_ptr and vnew_ptr go together");
  require((h.rotary_cos_ptr == nullptr) == (h.rotary_sin_ptr == nullptr), "rotary_cos_ptr and rotary_sin_ptr go together");
  require(h.rotary_cos_ptr == nullptr || h.knew_ptr != nullptr, "rotary embedding needs knew / vnew");
  require(h.p_dropout == 0.f || h.p_dropout >= 1.f, "dropout is not supported (p_dropout is the keep probability)");
  require(h.dv == 0 || h.dv == h.d, "dv != d is not supported");
  Params p;
  memset(&p, 0, sizeof(p));
  p.abi_version = FA_B200_ABI_VERSION;
  p.is_bf16 = (h.is_bf16 || h.is_e4m3) ? 1 : 0;         // e4m3 inputs: out is bf16 (flash_api.cu: out_type of the FP8 path)
  if (h.is_e4m3) {
    p.is_e4m3 = 1;
    p.q_descale = h.q_descale_ptr; p.k_descale = h.k_descale_ptr; p.v_descale = h.v_descale_ptr;
    p.q_descale_batch_stride = h.q_descale_batch_stride; p.q_descale_head_stride = h.q_descale_head_stride;
    p.k_descale_batch_stride = h.k_descale_batch_stride; p.k_descale_head_stride = h.k_descale_head_stride;
    p.v_descale_batch_stride = h.v_descale_batch_stride; p.v_descale_head_stride = h.v_descale_head_stride;
  }
  p.q = tens(h.q_ptr, h.q_batch_stride, h.q_row_stride, h.q_head_stride);
  p.k = tens(h.k_ptr, h.k_batch_stride, h.k_row_stride, h.k_head_stride);
  p.v = tens(h.v_ptr, h.v_batch_stride, h.v_row_stride, h.v_head_stride);
  p.o = tens(h.o_ptr, h.o_batch_stride, h.o_row_stride, h.o_head_stride);
  p.b = h.b; p.seqlen_q = h.seqlen_q; p.seqlen_k = h.seqlen_k; p.h = h.h; p.h_k = h.h_k; p.d = h.d;
  p.seqlen_q_rounded = fa::round_up(h.seqlen_q, 128); p.seqlen_k_rounded = fa::round_up(h.seqlen_k, 128);
  p.d_rounded = fa::round_up(h.d, 32);
  p.total_q = h.cu_seqlens_q ? h.total_q : -1;
  p.total_k = h.cu_seqlens_k ? h.total_k : -1;
  p.cu_seqlens_q = h.cu_seqlens_q; p.cu_seqlens_k = h.cu_seqlens_k; p.seqused_k = h.seqused_k;
  p.seqused_q = h.seqused_q;
  p.softmax_scale = h.scale_softmax;
  p.lse_empty_is_neg_inf = 1;                          // csrc/flashmask_v2/softmax.h:153
  p.is_causal = h.is_causal ? 1 : 0;
  p.causal_bottom_right = 1;                           // csrc/flashmask_v2/mask.h:85-97
  p.window_left = h.is_local ? h.window_size_left : -1;
  p.window_right = h.is_local ? h.window_size_right : -1;
  p.softcap = h.softcap;
  if (h.page_table) {                                  // csrc/flash_attn_v3/paged_kv.h: k/v are (num_pages, page_size, h_k, d)
    require(h.page_size > 0 && h.num_pages > 0, "page_table needs page_size and num_pages");
    p.page_table = h.page_table; p.page_table_batch_stride = h.page_table_batch_stride;
    p.page_size = h.page_size; p.num_pages = h.num_pages;
  }
  p.kv_batch_idx = h.kv_batch_idx; p.b_k = h.b_k;
  p.leftpad_k = h.leftpad_k;
  if (h.lt_start_ptr) {
    p.lt_start = h.lt_start_ptr; p.lt_end = h.lt_end_ptr; p.ut_start = h.ut_start_ptr; p.ut_end = h.ut_end_ptr;
    p.h_mask = h.h_flashmask > 0 ? h.h_flashmask : 1;
    p.flashmask_maxmin = h.flashmask_maxmin_ptr;
  }
  return p;
}

void set_split(Params &p, const Flash_fwd_params &h) {
  p.num_splits = h.num_splits;
  if (h.num_splits > 1) {
    require(h.oaccum_ptr && h.softmax_lseaccum_ptr, "split-KV needs oaccum_ptr and softmax_lseaccum_ptr");
    const int64_t hd = (int64_t)h.h * h.d, rows = (int64_t)h.seqlen_q;
    require((h.oaccum_head_stride == 0 || h.oaccum_head_stride == h.d) && (h.oaccum_row_stride == 0 || h.oaccum_row_stride == hd) &&
            (h.oaccum_batch_stride == 0 || h.oaccum_batch_stride == rows * hd) &&
            (h.oaccum_split_stride == 0 || h.oaccum_split_stride == (int64_t)h.b * rows * hd),
            "oaccum must be contiguous (splits, b, seqlen_q, h, d)");
    require((h.lseaccum_head_stride == 0 || h.lseaccum_head_stride == rows) &&
            (h.lseaccum_batch_stride == 0 || h.lseaccum_batch_stride == (int64_t)h.h * rows) &&
            (h.lseaccum_split_stride == 0 || h.lseaccum_split_stride == (int64_t)h.b * h.h * rows),
            "softmax_lseaccum must be contiguous (splits, b, h, seqlen_q)");
    p.oaccum = (float *)h.oaccum_ptr; p.lseaccum = (float *)h.softmax_lseaccum_ptr;
  }
}

__global__ void add_lengths_kernel(int *out, const int *in, int add, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (in ? in[i] : 0) + add;
}

// knew / vnew (+ rotary): rotate q and k_new, append k_new / v_new at seqused_k, attend over seqused_k + seqlen_knew
// (csrc/flash_attn_v3/mainloop_fwd_sm90_tma_gmma_ws.hpp:1090,1439,1542,1582).  Returns the stream-ordered scratch that
// the caller frees after the attention launch.
void *append_new_kv(Params &p, const Flash_fwd_params &h, cudaStream_t stream) {
  require(h.seqlen_knew > 0, "seqlen_knew must be set with knew_ptr");
  require(h.knew_batch_stride == h.vnew_batch_stride && h.knew_row_stride == h.vnew_row_stride &&
          h.knew_head_stride == h.vnew_head_stride, "knew and vnew must share their strides");
  const size_t esz = 2;
  const size_t q_bytes = h.rotary_cos_ptr ? (size_t)h.b * h.seqlen_q * h.h * h.d * esz : 0;
  const size_t k_bytes = h.rotary_cos_ptr ? (size_t)h.b * h.seqlen_knew * h.h_k * h.d * esz : 0;
  auto up256 = [](size_t x) { return (x + 255) / 256 * 256; };
  const size_t q_off = up256(sizeof(int) * (size_t)h.b), k_off = q_off + up256(q_bytes);   // [lengths | q_rot | k_rot]
  char *scratch = nullptr;
  FA_CUDA(cudaMallocAsync(&scratch, k_off + up256(k_bytes) + 256, stream));
  int *lens_after = (int *)scratch;
  const void *knew = h.knew_ptr;
  int64_t nbs = h.knew_batch_stride, nrs = h.knew_row_stride, nhs = h.knew_head_stride;
  try {
    if (h.rotary_cos_ptr) {
      require(h.rotary_dim > 0 && h.rotary_dim % 2 == 0 && h.rotary_dim <= h.d, "rotary_dim must be even and <= d");
      void *q_rot = scratch + q_off, *k_rot = scratch + k_off;
      const int64_t qh = h.d, qr = (int64_t)h.h * h.d, qb = qr * h.seqlen_q;
      const int64_t kh = h.d, kr = (int64_t)h.h_k * h.d, kb = kr * h.seqlen_knew;
      const int fixed = !(h.is_causal || h.is_local);
      fa::rotary(h.q_ptr, q_rot, h.rotary_cos_ptr, h.rotary_sin_ptr, h.seqused_k, h.b, h.seqlen_q, h.h, h.d, h.rotary_dim,
                 0x7fffffff, h.is_rotary_interleaved ? 1 : 0, fixed, p.is_bf16, h.q_batch_stride, h.q_row_stride,
                 h.q_head_stride, qb, qr, qh, stream);
      fa::rotary(h.knew_ptr, k_rot, h.rotary_cos_ptr, h.rotary_sin_ptr, h.seqused_k, h.b, h.seqlen_knew, h.h_k, h.d,
                 h.rotary_dim, 0x7fffffff, h.is_rotary_interleaved ? 1 : 0, 0, p.is_bf16, h.knew_batch_stride,
                 h.knew_row_stride, h.knew_head_stride, kb, kr, kh, stream);
      p.q = tens(q_rot, qb, qr, qh);
      knew = k_rot; nbs = kb; nrs = kr; nhs = kh;
    }
    p.seqused_k = h.seqused_k;                        // lengths before the append
    // K and V may be rotated / plain: two single-tensor appends through the same entry
    Params pk = p; pk.v = pk.k;
    fa::append_kv_paged(pk, knew, knew, h.seqlen_knew, nbs, nrs, nhs, stream);
    Params pv = p; pv.k = pv.v;
    fa::append_kv_paged(pv, h.vnew_ptr, h.vnew_ptr, h.seqlen_knew, h.vnew_batch_stride, h.vnew_row_stride,
                        h.vnew_head_stride, stream);
    add_lengths_kernel<<<fa::ceil_div(h.b, 128), 128, 0, stream>>>(lens_after, h.seqused_k, h.seqlen_knew, h.b);
    FA_CUDA(cudaGetLastError());
    p.seqused_k = lens_after;
  } catch (...) {
    cudaFreeAsync(scratch, stream);
    throw;
  }
  return scratch;
}

void run_fwd_handle(Flash_fwd_params *hp, cudaStream_t stream) {
  require(hp != nullptr, "null handle");
  const Flash_fwd_params &h = *hp;
  Params p = common_params(h);
  require(h.softmax_lse_ptr != nullptr, "softmax_lse_ptr must be set");
  set_split(p, h);
  if (h.knew_ptr) {
    require(h.cu_seqlens_q == nullptr && h.cu_seqlens_k == nullptr, "knew / vnew with cu_seqlens is not supported");
    void *scratch = append_new_kv(p, h, stream);
    p.softmax_lse = (float *)h.softmax_lse_ptr;
    try {
      if (h.num_splits > 1 || decode_shape(h)) fa::run_splitkv(p, stream); else fa::run_fwd(p, stream);
    } catch (...) {
      cudaFreeAsync(scratch, stream);
      throw;
    }
    FA_CUDA(cudaFreeAsync(scratch, stream));
    return;
  }
  // varlen: the kernels write LSE as (b, h, max_seqlen_q); this surface wants (h, total_q) (flash_fwd_launch_template.h:161-164)
  auto run_with_lse = [&](auto &&launch) {
    if (!h.cu_seqlens_q) { p.softmax_lse = (float *)h.softmax_lse_ptr; launch(); return; }
    require(h.total_q > 0, "total_q must be set with cu_seqlens_q");
    float *tmp = nullptr;
    FA_CUDA(cudaMallocAsync(&tmp, sizeof(float) * (size_t)h.b * h.h * h.seqlen_q, stream));
    p.softmax_lse = tmp;
    try { launch(); } catch (...) { cudaFreeAsync(tmp, stream); throw; }
    dim3 grid(fa::ceil_div(h.seqlen_q, 128), h.h, h.b);
    lse_repack_kernel<<<grid, 128, 0, stream>>>((float *)h.softmax_lse_ptr, tmp, h.cu_seqlens_q, h.b, h.h, h.seqlen_q,
                                                h.total_q, 1);
    FA_CUDA(cudaGetLastError());
    FA_CUDA(cudaFreeAsync(tmp, stream));
  };
  if (h.num_splits > 1 || decode_shape(h)) {
    if (h.num_splits > 1 && h.num_splits_dynamic_ptr) {
      // per-batch split counts (csrc/flash_attn_v3/flash.h:96-125): computed here unless the caller says the metadata in
      // num_splits_dynamic_ptr is already there (skip_scheduler_metadata_computation), then used by the kernel + combine
      if (!h.skip_scheduler_metadata_computation) fa::prepare_num_splits(p, h.num_splits_dynamic_ptr, h.num_sm, stream);
      p.num_splits_dynamic = h.num_splits_dynamic_ptr;
    }
    if (h.num_splits > 1) { p.softmax_lse = (float *)h.softmax_lse_ptr; fa::run_splitkv(p, stream); }   // partials only
    else run_with_lse([&] { fa::run_splitkv(p, stream); });
    return;
  }
  run_with_lse([&] { fa::run_fwd(p, stream); });
}

void run_combine_handle(Flash_fwd_params *hp, cudaStream_t stream) {
  require(hp != nullptr, "null handle");
  const Flash_fwd_params &h = *hp;
  Params p = common_params(h);
  require(h.num_splits > 1, "run_mha_fwd_combine needs num_splits > 1");
  set_split(p, h);
  require(h.softmax_lse_ptr && p.o.ptr, "o_ptr and softmax_lse_ptr must be set");
  if (!h.cu_seqlens_q) { p.softmax_lse = (float *)h.softmax_lse_ptr; fa::combine_fwd(p, stream); return; }
  require(h.total_q > 0, "total_q must be set with cu_seqlens_q");     // varlen: combine into (b, h, max), repack to (h, total_q)
  float *tmp = nullptr;
  FA_CUDA(cudaMallocAsync(&tmp, sizeof(float) * (size_t)h.b * h.h * h.seqlen_q, stream));
  p.softmax_lse = tmp;
  try { fa::combine_fwd(p, stream); } catch (...) { cudaFreeAsync(tmp, stream); throw; }
  dim3 grid(fa::ceil_div(h.seqlen_q, 128), h.h, h.b);
  lse_repack_kernel<<<grid, 128, 0, stream>>>((float *)h.softmax_lse_ptr, tmp, h.cu_seqlens_q, h.b, h.h, h.seqlen_q, h.total_q, 1);
  FA_CUDA(cudaGetLastError());
  FA_CUDA(cudaFreeAsync(tmp, stream));
}

void run_bwd_handle(Flash_bwd_params *hp, cudaStream_t stream) {
  require(hp != nullptr, "null handle");
  const Flash_bwd_params &h = *hp;
  Params p = common_params(h);
  require(h.do_ptr && h.dq_ptr && h.dk_ptr && h.dv_ptr && h.softmax_lse_ptr, "do / dq / dk / dv / softmax_lse must be set");
  p.dout = tens(h.do_ptr, h.do_batch_stride, h.do_row_stride, h.do_head_stride);
  p.dq = tens(h.dq_ptr, h.dq_batch_stride, h.dq_row_stride, h.dq_head_stride);
  p.dk = tens(h.dk_ptr, h.dk_batch_stride, h.dk_row_stride, h.dk_head_stride);
  p.dv = tens(h.dv_ptr, h.dv_batch_stride, h.dv_row_stride, h.dv_head_stride);
  p.dkv_expanded = 0;                      // dk / dv carry h_k heads; the group is reduced inside the kernel
  p.deterministic = h.deterministic ? 1 : 0;
  // Workspaces: the caller's dq_accum / dsoftmax_sum / softmax_lse_log2 are sized for the reference's sm90 tile shapes
  // (flash_bwd_launch_template.h:42-49), not for the 128-row tiles used here, so this path takes stream-ordered
  // scratch of its own size instead.
  const bool varlen = h.cu_seqlens_q != nullptr;
  if (varlen) require(h.total_q > 0, "total_q must be set with cu_seqlens_q");
  const size_t acc_rows = varlen ? (size_t)h.h * fa::round_up((int)(h.total_q + 128 * (int64_t)h.b), 128)
                                 : (size_t)h.b * h.h * p.seqlen_q_rounded;
  const size_t n_acc = acc_rows * p.d_rounded, n_d = (size_t)h.b * h.h * p.seqlen_q_rounded;
  const size_t n_lse = varlen ? (size_t)h.b * h.h * h.seqlen_q : 0;
  float *scratch = nullptr;
  FA_CUDA(cudaMallocAsync(&scratch, sizeof(float) * (n_acc + n_d + n_lse), stream));
  p.dq_accum = scratch; p.softmax_d = scratch + n_acc;
  if (varlen) {
    float *tmp = scratch + n_acc + n_d;
    dim3 grid(fa::ceil_div(h.seqlen_q, 128), h.h, h.b);
    lse_repack_kernel<<<grid, 128, 0, stream>>>((float *)h.softmax_lse_ptr, tmp, h.cu_seqlens_q, h.b, h.h, h.seqlen_q,
                                                h.total_q, 0);
    FA_CUDA(cudaGetLastError());
    p.softmax_lse = tmp;
  } else {
    p.softmax_lse = (float *)h.softmax_lse_ptr;
  }
  try {
    fa::run_bwd(p, stream);
  } catch (...) {
    cudaFreeAsync(scratch, stream);
    throw;
  }
  FA_CUDA(cudaFreeAsync(scratch, stream));
}

int num_splits_handle(const Flash_fwd_params &h) {
  const int num_sm = h.num_sm > 0 ? h.num_sm : fa::device_sm_count();
  if (h.h_k <= 0) return 1;
  if (decode_shape(h)) return fa_b200_decode_num_splits(h.b * ((h.seqlen_q * (h.h / h.h_k) + 15) / 16), h.h_k, h.seqlen_k, num_sm);
  return fa_b200_num_splits(h.b, h.h_k, h.seqlen_q * (h.h / h.h_k), h.seqlen_k, h.d, num_sm);
}

}  // namespace

extern "C" {

#define FA_HANDLE_LIFECYCLE(P)                                                                                       \
  Flash_fwd_params *P##_create_fwd_params_handle(void) { return new (std::nothrow) Flash_fwd_params(); }             \
  Flash_bwd_params *P##_create_bwd_params_handle(void) { return new (std::nothrow) Flash_bwd_params(); }             \
  void P##_clear_fwd_params_handle(Flash_fwd_params *h) { if (h) *h = Flash_fwd_params(); }                          \
  void P##_clear_bwd_params_handle(Flash_bwd_params *h) { if (h) *h = Flash_bwd_params(); }                          \
  Flash_fwd_params *P##_cast_to_fwd_params_handle(Flash_bwd_params *h) { return static_cast<Flash_fwd_params *>(h); } \
  void P##_destroy_fwd_params_handle(Flash_fwd_params *h) { delete h; }                                              \
  void P##_destroy_bwd_params_handle(Flash_bwd_params *h) { delete h; }                                              \
  void P##_run_mha_fwd_combine(Flash_fwd_params *h, cudaStream_t s, bool) { run_combine_handle(h, s); }              \
  void P##_run_mha_fwd(Flash_fwd_params *h, cudaStream_t s) { run_fwd_handle(h, s); }                                \
  bool P##_get_pagedkv_tma(Flash_fwd_params *) { return false; }                                                     \
  bool P##_get_pack_gqa(Flash_fwd_params *h) { return h && decode_shape(*h); }                                       \
  int P##_get_num_splits(Flash_fwd_params *h) { return h ? num_splits_handle(*h) : 1; }                              \
  void P##_run_mha_bwd(Flash_bwd_params *h, cudaStream_t s) { run_bwd_handle(h, s); }

FA_HANDLE_LIFECYCLE(flashmaskv2)
FA_HANDLE_LIFECYCLE(fa3)

#define FA_DEF_BOTH(P, type, member)                                                                      \
  type P##_fwd_params_get_##member(const Flash_fwd_params *h) { return h->member; }                       \
  void P##_fwd_params_set_##member(Flash_fwd_params *h, type value) { h->member = value; }          \
  type P##_bwd_params_get_##member(const Flash_bwd_params *h) { return h->member; }                       \
  void P##_bwd_params_set_##member(Flash_bwd_params *h, type value) { h->member = value; }
#define FA_DEF_BWD(P, type, member)                                                                       \
  type P##_bwd_params_get_##member(const Flash_bwd_params *h) { return h->member; }                       \
  void P##_bwd_params_set_##member(Flash_bwd_params *h, type value) { h->member = value; }
#define FA_DEF_BOTH_V2(type, member) FA_DEF_BOTH(flashmaskv2, type, member)
#define FA_DEF_BWD_V2(type, member) FA_DEF_BWD(flashmaskv2, type, member)
#define FA_DEF_BOTH_FA3(type, member) FA_DEF_BOTH(fa3, type, member)
#define FA_DEF_BWD_FA3(type, member) FA_DEF_BWD(fa3, type, member)

FA_HANDLE_FWD_FIELDS(FA_DEF_BOTH_V2)
FA_HANDLE_MASK_FIELDS(FA_DEF_BOTH_V2)
FA_HANDLE_BWD_FIELDS(FA_DEF_BWD_V2)
FA_HANDLE_BWD_MASK_FIELDS(FA_DEF_BWD_V2)
FA_HANDLE_FWD_FIELDS(FA_DEF_BOTH_FA3)
FA_HANDLE_BWD_FIELDS(FA_DEF_BWD_FA3)

}  // extern "C"
