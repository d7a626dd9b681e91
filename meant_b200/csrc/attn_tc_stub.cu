// Placeholder until the tcgen05 kernels land: reports "unsupported" so MEANT_IMPL_AUTO takes
// the exact path.  Replaced by attn_tc.cu.
#include "attn_common.cuh"
namespace meant {
int attn_tc_supported(const AttnArgs&, int, int) { return 0; }
int attn_tc_fwd(const AttnArgs&, cudaStream_t) { return MEANT_EUNSUPPORTED; }
int attn_tc_bwd(const AttnArgs&, cudaStream_t) { return MEANT_EUNSUPPORTED; }
}  // namespace meant
