// C-ABI glue: launch counter, version, attention dispatch (exact CUDA-core path vs tcgen05 path).
#include "attn_common.cuh"
#include <atomic>

namespace meant {
int attn_exact_fwd(const AttnArgs& a, int dtype, cudaStream_t st);
int attn_exact_bwd(const AttnArgs& a, int dtype, cudaStream_t st);
int attn_tc_supported(const AttnArgs& a, int dtype, int backward);
int attn_tc_fwd(const AttnArgs& a, cudaStream_t st);
int attn_tc_bwd(const AttnArgs& a, cudaStream_t st);
int attn_prerotate(const AttnArgs& a, int dtype, cudaStream_t st);
}  // namespace meant

static std::atomic<int64_t> g_launches{0};

extern "C" void meant_count_launch_(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
extern "C" int64_t meant_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
extern "C" void meant_reset_launch_count(void) { g_launches.store(0, std::memory_order_relaxed); }
namespace meant { long long attn_tma_fallbacks(); }
extern "C" int64_t meant_attn_tma_fallbacks(void) { return meant::attn_tma_fallbacks(); }

extern "C" int meant_version(void) { return 5; }

extern "C" int meant_attn_tcgen05_supported(const meant_attn_args* args, int dtype, int backward) {
  return args ? meant::attn_tc_supported(*args, dtype, backward) : 0;
}

extern "C" int meant_attn_fwd(const meant_attn_args* args, int dtype, int impl, meant_stream_t stream) {
  if (!args) return MEANT_EARG;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const bool tc_ok = meant::attn_tc_supported(*args, dtype, 0) != 0;
  if (impl == MEANT_IMPL_TCGEN05 && !tc_ok) return MEANT_EUNSUPPORTED;
  if (impl == MEANT_IMPL_TCGEN05 || (impl == MEANT_IMPL_AUTO && tc_ok)) return meant::attn_tc_fwd(*args, st);
  return meant::attn_exact_fwd(*args, dtype, st);
}

extern "C" int meant_attn_bwd(const meant_attn_args* args, int dtype, int impl, meant_stream_t stream) {
  if (!args) return MEANT_EARG;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const bool tc_ok = meant::attn_tc_supported(*args, dtype, 1) != 0;
  if (impl == MEANT_IMPL_TCGEN05 && !tc_ok) return MEANT_EUNSUPPORTED;
  if (impl == MEANT_IMPL_TCGEN05 || (impl == MEANT_IMPL_AUTO && tc_ok)) return meant::attn_tc_bwd(*args, st);
  return meant::attn_exact_bwd(*args, dtype, st);
}

extern "C" int meant_attn_prerotate(const meant_attn_args* args, int dtype, meant_stream_t stream) {
  if (!args) return MEANT_EARG;
  return meant::attn_prerotate(*args, dtype, reinterpret_cast<cudaStream_t>(stream));
}
