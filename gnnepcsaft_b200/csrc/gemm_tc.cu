// tcgen05 + TMEM GEMM for the dense transforms (placeholder until the kernel lands: reports
// "not supported" so every shape runs on the fp32 kernel).
#include "common.cuh"

namespace gnnx {

bool tc_supported(int64_t, int64_t, int64_t, int, int, int64_t, int64_t, int64_t, const void*, const void*, const void*, int) {
    return false;
}
size_t tc_workspace_bytes(int64_t, int64_t, int64_t, int, int, int) { return 0; }
int gemm_tc(const float*, int64_t, int, const float*, int64_t, int, float*, int64_t, int64_t, int64_t, int64_t, const float*,
            int, int, void*, size_t, cudaStream_t) {
    return set_error(GNNX_ERR_UNSUPPORTED, "gemm_tc: tensor-core path not built");
}

}  // namespace gnnx
