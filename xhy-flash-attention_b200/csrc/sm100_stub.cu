// temporary stubs (replaced by sm100_fwd.cu / sm100_bwd.cu / decode.cu)
#include "common.h"
namespace fa {
#if 0
bool sm100_fwd_supported(const Params &, std::string *) { return false; }
void sm100_fwd(const Params &, cudaStream_t) {}
#endif
#if 0
bool sm100_bwd_supported(const Params &, std::string *) { return false; }
void sm100_bwd(const Params &, cudaStream_t) {}
#endif
#if 1
void splitkv_fwd(const Params &p, cudaStream_t s) { generic_splitkv_fwd(p, s); }
#endif
}
