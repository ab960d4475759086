// tcgen05 / TMA path for the dense half-step (placeholder until the kernel lands).
#include "common.cuh"
namespace cnmf {
bool tc_supported(int64_t, int64_t, int, int) { return false; }
int64_t tc_workspace_bytes(int64_t, int64_t, int, int) { return 0; }
int launch_tc_half_step(const float*, int64_t, int64_t, int64_t, const float*, int64_t, float*, int64_t,
                        const float*, const uint8_t*, int, int, int, float, float, float, double*, void*,
                        int64_t, cudaStream_t) {
  set_error("tcgen05 path not built");
  return CNMF_B200_ERR_UNSUPPORTED;
}
}  // namespace cnmf
