// tcgen05 interaction kernels (placeholder until the tensor-core path lands; the dispatcher
// in api.cu falls through to interact_simt.cu while this reports "unsupported").
#include "common.cuh"

namespace b200 {
bool interact_tc_supported(int64_t, int64_t, int64_t, int) { return false; }
int interact_fwd_tc(const void*, void*, int64_t, int64_t, int64_t, int, cudaStream_t) {
  set_error("interact_fwd_tc: not built");
  return B200_EUNSUPPORTED;
}
int interact_bwd_tc(const void*, const void*, void*, int64_t, int64_t, int64_t, int, cudaStream_t) {
  set_error("interact_bwd_tc: not built");
  return B200_EUNSUPPORTED;
}
}  // namespace b200
