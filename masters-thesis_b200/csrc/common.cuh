// Shared helpers for libb200lstm.so: error plumbing, Philox4x32-10, warp reductions.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/b200_lstm.h"

namespace b200 {

// ---- error plumbing ------------------------------------------------------------------------
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define B200_CHECK_ARG(cond, ...)                 \
    do {                                          \
        if (!(cond)) {                            \
            b200::set_error(__VA_ARGS__);         \
            return B200_ERR_INVALID;              \
        }                                         \
    } while (0)

#define B200_CUDA(call)                                            \
    do {                                                           \
        cudaError_t _e = (call);                                   \
        if (_e != cudaSuccess) return b200::cuda_fail(_e, #call);  \
    } while (0)

#define B200_LAUNCH_CHECK(name)                                           \
    do {                                                                  \
        cudaError_t _e = cudaGetLastError();                              \
        if (_e != cudaSuccess) return b200::cuda_fail(_e, "launch " name); \
    } while (0)

// ---- optional per-kernel timing (bench.py's roofline leg): CUDA events recorded on the launching
// stream around each named launch while enabled; zero cost when disabled.
void timing_begin(const char* name, cudaStream_t st);
void timing_end(cudaStream_t st);
struct TimedLaunch {
    cudaStream_t st;
    TimedLaunch(const char* name, cudaStream_t s) : st(s) { timing_begin(name, s); }
    ~TimedLaunch() { timing_end(st); }
};

inline int num_sms() {
    static int cached[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (cached[dev] == 0) {
        int n = 0;
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        cached[dev] = n > 0 ? n : 148;
    }
    return cached[dev];
}

// ---- Philox4x32-10 (Salmon et al., SC'11); spec mirrored in oracle/lstm_oracle.py -----------
struct U4 {
    uint32_t x, y, z, w;
};

__host__ __device__ __forceinline__ U4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                      uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0;
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
        uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
        uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0;
        c1 = lo1;
        c2 = n2;
        c3 = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return U4{c0, c1, c2, c3};
}

// Inter-layer dropout: one Philox block covers 8 consecutive hidden units of one (row, t, layer);
// unit j of the block keeps iff its 16-bit lane >= round(p * 65536).  Identical in fwd, bwd and
// b200_dropout_mask.  counter = (row_lo, row_hi, t | layer << 16, unit / 8).
__host__ __device__ __forceinline__ uint32_t dropout_threshold(float p) {
    float t = p * 65536.0f + 0.5f;
    return t <= 0.f ? 0u : (t >= 65535.f ? 65535u : (uint32_t)t);
}
__device__ __forceinline__ U4 dropout_block(uint64_t seed, int layer, int64_t row, int t, int unit_block) {
    return philox4x32_10((uint32_t)row, (uint32_t)((uint64_t)row >> 32), (uint32_t)t | ((uint32_t)layer << 16),
                         (uint32_t)unit_block, (uint32_t)seed, (uint32_t)(seed >> 32));
}
__device__ __forceinline__ uint32_t dropout_lane16(const U4& b, int j) {  // j in 0..7
    uint32_t w = (j < 2) ? b.x : (j < 4) ? b.y : (j < 6) ? b.z : b.w;
    return (w >> (16 * (j & 1))) & 0xFFFFu;
}

// ---- reductions -------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ float sigmoid_precise(float x) { return 1.0f / (1.0f + __expf(-x)); }
__device__ __forceinline__ float tanh_precise(float x) {
    // 1 - 2/(exp(2x)+1): exp overflow -> inf -> 1, underflow -> -1; ~2 ulp of __expf + division
    float e = __expf(2.0f * x);
    return 1.0f - 2.0f / (e + 1.0f);
}

}  // namespace b200
