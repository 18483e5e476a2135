// placeholder until the tcgen05 contraction lands (next commit): reports "unsupported" so AUTO stays on the FFMA path
#include "search.cuh"
namespace lipvq {
bool search_tc_supported(int64_t, int, int) { return false; }
float search_tc_gamma(int D) { return (float)(D + 4) * 1.2e-5f; }
int search_tc_launch(const float*, int64_t, int, const float*, int, const SearchPlan&, char*, cudaStream_t) {
  set_error("tcgen05 search path not built");
  return LIPVQ_ERR_UNSUPPORTED;
}
}  // namespace lipvq
