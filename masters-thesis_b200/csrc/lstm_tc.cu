// tcgen05 LSTM engines (B200_MODE_TC_PRECISE / B200_MODE_TC_FAST) -- placeholder until the kernels land.
#include "common.cuh"

namespace b200 {

bool tc_available() { return false; }

size_t tc_workspace_bytes(int, int64_t, int, int, int, int) { return 0; }

int tc_fwd(int, const float*, int, int, int, const float*, int64_t, int, float*, float*, float, uint64_t, int, void*, size_t,
           cudaStream_t) {
    set_error("tcgen05 LSTM engine not built into this library");
    return B200_ERR_UNSUPPORTED;
}

int tc_bwd_layer(int, int, const float*, int, int, int, const float*, int64_t, int, const float*, const float*, const float*,
                 float*, float, uint64_t, int, void*, size_t, cudaStream_t) {
    set_error("tcgen05 LSTM engine not built into this library");
    return B200_ERR_UNSUPPORTED;
}

}  // namespace b200
