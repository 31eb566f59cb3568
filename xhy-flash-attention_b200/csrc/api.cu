// api.cu -- the extern "C" surface of libflashattn_b200.so (include/fa_capi.h, include/fa_b200.h).
//
// Error convention follows the reference (csrc/capi/flash_attn.cu:45-67): every body is wrapped in
// try/catch, the message is kept in thread-local storage and the call returns false.
// There is no CPU path here: every entry point enqueues CUDA kernels on the caller's stream.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <algorithm>
#include <atomic>
#include <mutex>
#include <vector>

#include "common.h"
#include "../../include/fa_capi.h"

namespace fa {

static thread_local std::unique_ptr<char[]> tls_error;
static thread_local int tls_last_kind = 0;
static thread_local int tls_fwd_reserve = 0;   // fa_b200_set_forward_sm_reserve
static std::atomic<int64_t> g_launches{0};  // process-wide: backward runs on autograd worker threads

void note_launch(int kind) {
  tls_last_kind = kind;
  g_launches.fetch_add(1, std::memory_order_relaxed);
}

// Optional per-kernel timing (fa_b200_profile_*).  Process-wide on purpose: a backward launched from autograd's worker
// thread is read back from the thread that enabled profiling.  One mutex serialises the few host-side touches.
static std::atomic<bool> g_profile{false};
static std::mutex g_profile_mu;
static cudaEvent_t g_events[2][2] = {{nullptr, nullptr}, {nullptr, nullptr}};
static bool g_marked[2] = {false, false};

void profile_mark(int slot, bool begin, cudaStream_t s) {
  if (!g_profile.load(std::memory_order_relaxed)) return;
  std::lock_guard<std::mutex> lk(g_profile_mu);
  cudaEvent_t &e = g_events[slot][begin ? 0 : 1];
  if (!e) FA_CUDA(cudaEventCreate(&e));
  FA_CUDA(cudaEventRecord(e, s));
  if (!begin) g_marked[slot] = true;
}

// A call that leaves the tcgen05 envelope runs on the generic SIMT kernels, two to three orders of magnitude slower.
// That must not be silent: the reason is kept per thread (fa_b200_last_fallback_reason), counted process-wide
// (fa_b200_fallback_count) and printed to stderr once per distinct reason unless FA_B200_QUIET is set.
static thread_local std::string tls_fallback_reason;
static std::atomic<int64_t> g_fallbacks{0};
static void note_fallback(const char *pass, const std::string &why) {
  tls_fallback_reason = std::string(pass) + ": " + why;
  g_fallbacks.fetch_add(1, std::memory_order_relaxed);
  static std::mutex mu;
  static std::vector<std::string> seen;
  static const bool quiet = getenv("FA_B200_QUIET") != nullptr;
  if (quiet) return;
  std::lock_guard<std::mutex> lk(mu);
  for (const auto &x : seen) if (x == tls_fallback_reason) return;
  seen.push_back(tls_fallback_reason);
  fprintf(stderr, "[libflashattn_b200] %s -> generic SIMT kernels (slow path; this message is printed once per reason)\n",
          tls_fallback_reason.c_str());
}

struct DevInfo { int sm_count; int cc_major; int cc_minor; };
static DevInfo dev_info() {
  // once-initialised per-device cache (the reference keeps the same kind of cache: csrc/capi/cuda_utils.cu:12-48)
  static std::mutex mu;
  static DevInfo cache[64];
  static bool have[64] = {};
  int dev = 0;
  FA_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> g(mu);
  if (dev < 64 && have[dev]) return cache[dev];
  DevInfo d;
  FA_CUDA(cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, dev));
  FA_CUDA(cudaDeviceGetAttribute(&d.cc_major, cudaDevAttrComputeCapabilityMajor, dev));
  FA_CUDA(cudaDeviceGetAttribute(&d.cc_minor, cudaDevAttrComputeCapabilityMinor, dev));
  if (dev < 64) { cache[dev] = d; have[dev] = true; }
  return d;
}
int device_sm_count() { return dev_info().sm_count; }
int forward_grid_sms() { return std::max(1, dev_info().sm_count - tls_fwd_reserve); }
int device_cc_major() { return dev_info().cc_major; }

static void check_common(const Params &p) {
  FA_CHECK(p.abi_version == FA_B200_ABI_VERSION, "fa_b200_params.abi_version mismatch");
  FA_CHECK(p.b > 0, "batch size must be positive");
  FA_CHECK(p.h > 0 && p.h_k > 0 && p.h % p.h_k == 0, "num_heads must be a multiple of num_heads_k");
  FA_CHECK(p.d > 0 && p.d % 8 == 0 && p.d <= 256, "head_size must be a multiple of 8 and <= 256");
  FA_CHECK(p.q.ptr && p.k.ptr && p.v.ptr, "q, k, v must be set");
  FA_CHECK(device_cc_major() >= 10, "libflashattn_b200 needs a Blackwell (sm_100a) device");
  if (p.attn_mask) FA_CHECK(p.mask_head_mod > 0 && p.mask_seq_mod > 0, "attn_mask needs mask_head_mod and mask_seq_mod");
  FA_CHECK(p.softcap >= 0.f, "softcap must be >= 0");
  if (p.attn_bias || p.ds_out) FA_CHECK(p.bias_batch_mod > 0 && p.bias_rows > 0, "attn_bias / ds_out need bias_batch_mod and bias_rows");
  if (p.page_table) {
    FA_CHECK(p.page_size > 0 && p.num_pages > 0 && p.page_table_batch_stride > 0, "page_table needs page_size, num_pages and its batch stride");
    FA_CHECK(!p.kv_batch_idx, "page_table and kv_batch_idx are exclusive");
    FA_CHECK(!p.cu_seqlens_k, "page_table with cu_seqlens_k is not supported");
  }
  if (p.lt_start) {
    FA_CHECK(p.h_mask > 0 && p.h % p.h_mask == 0, "flashmask heads must divide num_heads");
  } else {
    FA_CHECK(!p.lt_end && !p.ut_start && !p.ut_end, "flashmask: lt_start is mandatory when any vector is given");
  }
}

void run_fwd(const Params &p, cudaStream_t s) {
  check_common(p);
  FA_CHECK(p.o.ptr && p.softmax_lse, "out and softmax_lse must be set");
  std::string why;
  FA_CHECK(p.p_dropout >= 0.f && p.p_dropout < 1.f, "p_dropout must be in [0, 1)");
  if (p.is_e4m3) {
    // FP8 forward (csrc/flashmask_v2/flash_api.cu:107-131): tensor-core kernel or nothing -- there is no byte-operand
    // generic kernel to fall back to
    FA_CHECK(!p.force_generic && p.s_dmask == nullptr && p.alibi_slopes == nullptr && p.page_table == nullptr &&
             p.kv_batch_idx == nullptr && p.leftpad_k == nullptr && p.attn_bias == nullptr,
             "FP8 forward: plain / causal / local / varlen calls only");
    if (!sm100_fwd_supported(p, &why)) throw Error("FP8 forward: " + why);
    tls_fallback_reason.clear();
    sm100_fwd(p, s);
    return;
  }
  if (p.force_generic) { generic_fwd(p, s); return; }      // asked for by the caller (tests, FA1 symbols): not a fallback
  // return_softmax: out and LSE come from the same tcgen05 kernels as without it, then one more tcgen05 pass writes
  // S_dmask from Q, K and the finished LSE (sm100_probs.cu).  Only logits that are more than scale * q.k (dense mask,
  // softcap, ALiBi, FA1 bias) keep the element-wise generic pass for it.
  if (p.s_dmask != nullptr && sm100_probs_eligible(p)) {
    Params main_pass = p;
    main_pass.s_dmask = nullptr;
    if (sm100_fwd_supported(main_pass, &why)) {
      tls_fallback_reason.clear();
      sm100_fwd(main_pass, s);
      sm100_probs(p, s);
      return;
    }
    note_fallback("forward", why);
    generic_fwd(p, s);
    return;
  }
  // ALiBi and the KV-cache addressing modes outside the decode kernel (inference features) run on the generic kernels.
  // Dropout, the dense attn_mask and softcap have tcgen05 instantiations.
  const char *special = p.s_dmask != nullptr ? "return_softmax (S_dmask) with a dense mask / softcap / ALiBi / bias"
                        : p.alibi_slopes != nullptr ? "alibi_slopes outside the decode kernel"
                        : p.page_table != nullptr ? "paged KV outside the decode kernel"
                        : p.kv_batch_idx != nullptr ? "kv_batch_idx outside the decode kernel"
                        : p.leftpad_k != nullptr ? "leftpad_k outside the decode kernel"
                        : p.attn_bias != nullptr ? "attn_bias (FA1 bias path)" : nullptr;
  if (special) why = special;
  if (!special && sm100_fwd_supported(p, &why)) { tls_fallback_reason.clear(); sm100_fwd(p, s); return; }
  note_fallback("forward", why);
  generic_fwd(p, s);
}

void run_bwd(const Params &p, cudaStream_t s) {
  check_common(p);
  FA_CHECK(p.o.ptr && p.softmax_lse && p.dout.ptr && p.dq.ptr && p.dk.ptr && p.dv.ptr, "bwd tensors must be set");
  FA_CHECK(p.softmax_d, "softmax_d workspace must be set");
  FA_CHECK(!p.is_e4m3, "FP8 is forward-only (as in the reference: csrc/flashmask_v2/flash_api.cu:107-131 has no e4m3 backward)");
  std::string why;
  FA_CHECK(p.p_dropout >= 0.f && p.p_dropout < 1.f, "p_dropout must be in [0, 1)");
  FA_CHECK(!p.page_table && !p.kv_batch_idx && !p.leftpad_k, "backward: KV-cache addressing (page_table / kv_batch_idx / leftpad_k) is forward-only");
  const char *special = p.alibi_slopes != nullptr ? "alibi_slopes"
                        : (p.attn_bias != nullptr || p.ds_out != nullptr) ? "attn_bias / dbias (FA1 bias path)" : nullptr;
  if (p.force_generic) { generic_bwd(p, s); return; }
  if (special) why = special;
  if (!special && sm100_bwd_supported(p, &why)) { tls_fallback_reason.clear(); sm100_bwd(p, s); return; }
  note_fallback("backward", why);
  generic_bwd(p, s);
}

void run_splitkv(const Params &p, cudaStream_t s) {
  check_common(p);
  FA_CHECK(!p.is_e4m3, "FP8 split-KV is not supported");
  if (p.num_splits > 1) FA_CHECK(p.oaccum && p.lseaccum, "split-KV needs oaccum and lseaccum");
  else FA_CHECK(p.o.ptr && p.softmax_lse, "out and softmax_lse must be set");
  splitkv_fwd(p, s);
}

}  // namespace fa

using fa::Params;

#define FA_BEGIN try {
#define FA_END                                                \
  }                                                           \
  catch (const std::exception &e) {                           \
    flash_attn_set_error(e.what());                           \
    return false;                                             \
  }                                                           \
  catch (...) {                                               \
    flash_attn_set_error(nullptr);                            \
    return false;                                             \
  }                                                           \
  return true;

extern "C" {

void flash_attn_set_error(const char *msg) {
  if (msg == nullptr || *msg == '\0') msg = "unknown error";
  size_t n = strlen(msg);
  fa::tls_error.reset(new char[n + 1]);
  memcpy(fa::tls_error.get(), msg, n + 1);
}

const char *flash_attn_error(void) { return fa::tls_error ? fa::tls_error.get() : ""; }

void fa_b200_profile_enable(int on) {
  std::lock_guard<std::mutex> lk(fa::g_profile_mu);
  fa::g_profile.store(on != 0);
  fa::g_marked[0] = fa::g_marked[1] = false;
}

// why the last fa_b200_fwd / fa_b200_bwd / flash_attn_* call of this thread left the tcgen05 kernels ("" if it did not)
const char *fa_b200_last_fallback_reason(void) { return fa::tls_fallback_reason.c_str(); }
int64_t fa_b200_fallback_count(int reset) { return reset ? fa::g_fallbacks.exchange(0) : fa::g_fallbacks.load(); }

// elapsed device time (ms) of the last forward kernel (slot 0) / backward main kernel (slot 1) launched on this
// thread while profiling was enabled; -1 if none.  Blocks until that kernel has finished.
float fa_b200_profile_last_ms(int slot) {
  std::lock_guard<std::mutex> lk(fa::g_profile_mu);
  if (slot < 0 || slot > 1 || !fa::g_marked[slot]) return -1.f;
  float ms = -1.f;
  if (cudaEventSynchronize(fa::g_events[slot][1]) != cudaSuccess) return -1.f;
  if (cudaEventElapsedTime(&ms, fa::g_events[slot][0], fa::g_events[slot][1]) != cudaSuccess) return -1.f;
  return ms;
}

int fa_b200_abi_version(void) { return FA_B200_ABI_VERSION; }
int fa_b200_last_kernel_kind(void) { return fa::tls_last_kind; }
void fa_b200_set_forward_sm_reserve(int num_sms) { fa::tls_fwd_reserve = num_sms > 0 ? num_sms : 0; }
int64_t fa_b200_launch_count(int reset) {
  int64_t v = reset ? fa::g_launches.exchange(0) : fa::g_launches.load();
  return v;
}

bool fa_b200_fwd(const fa_b200_params *p, cudaStream_t stream) {
  FA_BEGIN
  FA_CHECK(p != nullptr, "null params");
  fa::run_fwd(*p, stream);
  FA_END
}

bool fa_b200_bwd(const fa_b200_params *p, cudaStream_t stream) {
  FA_BEGIN
  FA_CHECK(p != nullptr, "null params");
  fa::run_bwd(*p, stream);
  FA_END
}

bool fa_b200_fwd_splitkv(const fa_b200_params *p, cudaStream_t stream) {
  FA_BEGIN
  FA_CHECK(p != nullptr, "null params");
  fa::run_splitkv(*p, stream);
  FA_END
}

bool fa_b200_fwd_combine(const fa_b200_params *p, cudaStream_t stream) {
  FA_BEGIN
  FA_CHECK(p != nullptr, "null params");
  FA_CHECK(p->oaccum && p->lseaccum && p->o.ptr && p->softmax_lse && p->num_splits >= 1, "combine needs partials and outputs");
  fa::combine_fwd(*p, stream);
  FA_END
}

// Restates num_splits_heuristic (csrc/flashmask_v2/heuristics.h:22-59): smallest split count reaching
// 85% of the best wave efficiency, no split when the grid already fills 80% of the SMs.
int fa_b200_num_splits(int batch, int num_heads_k, int seqlen_q_packed, int seqlen_k, int head_size, int num_sm) {
  const int block_n = 128, block_m = 64;
  int num_n_blocks = (seqlen_k + block_n - 1) / block_n;
  int num_m_blocks = (seqlen_q_packed + block_m - 1) / block_m;
  int total_mblocks = batch * num_heads_k * num_m_blocks;
  if (num_sm <= 0) num_sm = 148;
  if (total_mblocks >= 0.8f * num_sm) return 1;
  if (num_n_blocks <= 4) return 1;
  int max_splits = 128;
  if (max_splits > num_sm) max_splits = num_sm;
  if (max_splits > num_n_blocks) max_splits = num_n_blocks;
  float best = 0.f;
  float eff[128];
  for (int s = 1; s <= max_splits; ++s) {
    float n_waves = float(total_mblocks * s) / num_sm;
    float e = n_waves / ceilf(n_waves);
    eff[s - 1] = e;
    if (e > best) best = e;
  }
  for (int s = 1; s <= max_splits; ++s)
    if (eff[s - 1] >= 0.85f * best) return s;
  (void)head_size;
  return 1;
}

int fa_b200_decode_num_splits(int batch, int num_heads_k, int seqlen_k, int num_sm) {
  if (num_sm <= 0) num_sm = 148;
  const int slots = 2 * num_sm;               // decode_kernel: 2 resident CTAs per SM
  const int base = batch * num_heads_k;
  const int nblk = (seqlen_k + 127) / 128;
  const int overhead = 2;                     // per-CTA prologue + merge, in units of one 128-key block
  int best_s = 1;
  long best_cost = -1;
  const int max_s = nblk < 64 ? nblk : 64;
  for (int s = 1; s <= max_s; ++s) {
    const int per = (nblk + s - 1) / s;
    if (s > 1 && per < 2) break;
    const long waves = ((long)base * s + slots - 1) / slots;
    const long cost = waves * (per + overhead);
    if (best_cost < 0 || cost < best_cost) { best_cost = cost; best_s = s; }
  }
  return best_s;
}

bool fa_b200_prepare_num_splits(const fa_b200_params *p, int32_t *num_splits_dynamic_out, int num_sm, cudaStream_t stream) {
  FA_BEGIN
  FA_CHECK(p && p->abi_version == FA_B200_ABI_VERSION, "fa_b200_params.abi_version mismatch");
  fa::prepare_num_splits(*p, num_splits_dynamic_out, num_sm, stream);
  FA_END
}

bool fa_b200_append_kv(const void *k_new, const void *v_new, void *k_cache, void *v_cache,
                       const int32_t *cache_seqlens, int batch, int seqlen_new, int seqlen_cache,
                       int num_heads_k, int head_size, int is_bf16,
                       int64_t nbs, int64_t nrs, int64_t nhs, int64_t cbs, int64_t crs, int64_t chs,
                       cudaStream_t stream) {
  FA_BEGIN
  FA_CHECK(k_new && v_new && k_cache && v_cache, "append_kv: null tensor");
  FA_CHECK(head_size % 8 == 0, "append_kv: head_size must be a multiple of 8");
  fa::append_kv(k_new, v_new, k_cache, v_cache, cache_seqlens, batch, seqlen_new, seqlen_cache, num_heads_k,
                head_size, is_bf16, nbs, nrs, nhs, cbs, crs, chs, stream);
  FA_END
}

bool fa_b200_flashmask_prepare(const fa_b200_params *p, int block_n, cudaStream_t stream) {
  FA_BEGIN
  FA_CHECK(p && p->lt_start && p->flashmask_maxmin, "flashmask_prepare needs lt_start and the maxmin workspace");
  fa::flashmask_prepare(*p, block_n, stream);
  FA_END
}

bool fa_b200_append_kv_paged(const fa_b200_params *p, const void *k_new, const void *v_new, int seqlen_new,
                             int64_t new_batch_stride, int64_t new_row_stride, int64_t new_head_stride,
                             cudaStream_t stream) {
  FA_BEGIN
  FA_CHECK(p && p->abi_version == FA_B200_ABI_VERSION, "fa_b200_params.abi_version mismatch");
  FA_CHECK(k_new && v_new && p->k.ptr && p->v.ptr, "append_kv: null tensor");
  fa::append_kv_paged(*p, k_new, v_new, seqlen_new, new_batch_stride, new_row_stride, new_head_stride, stream);
  FA_END
}

bool fa_b200_rotary(const void *x, void *out, const void *cos, const void *sin, const int32_t *seqlen_offsets,
                    int batch, int seqlen, int num_heads, int head_size, int rotary_dim, int seqlen_ro,
                    int interleaved, int fixed_position, int is_bf16,
                    int64_t x_batch_stride, int64_t x_row_stride, int64_t x_head_stride,
                    int64_t out_batch_stride, int64_t out_row_stride, int64_t out_head_stride, cudaStream_t stream) {
  FA_BEGIN
  FA_CHECK(x && out && cos && sin, "rotary: null tensor");
  fa::rotary(x, out, cos, sin, seqlen_offsets, batch, seqlen, num_heads, head_size, rotary_dim, seqlen_ro, interleaved,
             fixed_position, is_bf16, x_batch_stride, x_row_stride, x_head_stride, out_batch_stride, out_row_stride,
             out_head_stride, stream);
  FA_END
}

bool fa_b200_lse_merge(void *o, float *lse, const void *o_step, const float *lse_step, int64_t rows, int heads,
                       int head_size, int is_bf16, int o_is_fp32, int64_t lse_row_stride, int64_t lse_head_stride,
                       int64_t lse_step_row_stride, int64_t lse_step_head_stride, cudaStream_t stream) {
  FA_BEGIN
  FA_CHECK(o && lse && o_step && lse_step, "lse_merge: null tensor");
  fa::lse_merge(o, lse, o_step, lse_step, 1, rows, heads, head_size, is_bf16, o_is_fp32, lse_row_stride,
                lse_head_stride, lse_step_row_stride, lse_step_head_stride, 0, 0, 0, 0, stream);
  FA_END
}

bool fa_b200_lse_merge_batched(void *o, float *lse, const void *o_step, const float *lse_step, int batch, int64_t rows,
                               int heads, int head_size, int is_bf16, int o_is_fp32, int64_t o_batch_stride,
                               int64_t o_step_batch_stride, int64_t lse_batch_stride, int64_t lse_row_stride,
                               int64_t lse_head_stride, int64_t lse_step_batch_stride, int64_t lse_step_row_stride,
                               int64_t lse_step_head_stride, cudaStream_t stream) {
  FA_BEGIN
  FA_CHECK(o && lse && o_step && lse_step, "lse_merge: null tensor");
  fa::lse_merge(o, lse, o_step, lse_step, batch, rows, heads, head_size, is_bf16, o_is_fp32, lse_row_stride,
                lse_head_stride, lse_step_row_stride, lse_step_head_stride, o_batch_stride, o_step_batch_stride,
                lse_batch_stride, lse_step_batch_stride, stream);
  FA_END
}

}  // extern "C"
