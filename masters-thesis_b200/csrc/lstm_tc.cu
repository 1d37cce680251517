// tcgen05 LSTM engine (B200_MODE_TC_PRECISE, B200_MODE_TC_FAST): the recurrent gate GEMM on the
// 5th-generation tensor cores with the accumulator in TMEM and the LSTM cell fused into the epilogue.
//
// Replaces self.lstm(lstm_input) -- nn.LSTM forward (src/model.py:104, module 88-94; oneDNN
// mkldnn_rnn_layer on the host, cuDNN _cudnn_rnn on a GPU) one layer per launch.
//
// Structure of `lstm_tc_fwd_kernel<NSPLIT, NTILE, LAYER0>` (persistent, one CTA per SM):
//   * a tile is 128 consecutive sequences; the CTA keeps NTILE tiles in flight ("slots"), each with
//     its own 128-lane x 256-column fp32 accumulator in TMEM (gates i,f,g,o of 64 hidden units, in
//     the natural torch row order of W), its own h operand buffer and its own x operand buffer
//   * the layer's weights stay resident in shared memory for the whole launch as bf16 hi (+ lo) images
//     in the UMMA interleaved K-major layout (tc_common.cuh); `tc_pack_weights_kernel` builds the images
//   * warp roles: 4 epilogue warps per slot (thread = one sequence: TMEM -> registers -> sigma/tanh/cell
//     -> h back to shared memory as the next step's A operand, h and c to the stash), 2 loader warps
//     (x_t operand of the next step, coalesced, dropout applied), 1 MMA warp (one thread issues)
//   * three mbarriers per slot: x_full (loader -> MMA), h_ready (epilogue -> MMA: h_t written, TMEM
//     drained), acc_full (tcgen05.commit -> epilogue and loader: gates ready, operands consumed)
//   * NSPLIT = 2 (precise): every product is hi*hi + lo*hi + hi*lo on bf16 halves of the fp32 operands,
//     fp32 accumulate -- three tcgen05.mma per K-step, error ~2^-16 per product; NSPLIT = 1 (fast): bf16.
// Per step and tile the tensor pipe does (KX + 64)/16 K-steps x (3 | 1) MMAs of 128x256x16.
//
// The stash format is the same transparent fp32 [L][T][N][H] as the CUDA-core engine, so backward can
// run on either engine.
#include "common.cuh"
#include "tc_common.cuh"

namespace b200 {

using namespace tc;

constexpr int TH = 64;              // hidden size of this engine
constexpr int TG = 4 * TH;          // gate columns = UMMA N
constexpr int TM = 128;             // sequences per tile = UMMA M
constexpr int LOADER_WARPS = 2;

__host__ __device__ constexpr int kx_of(bool layer0) { return layer0 ? 16 : TH; }
// bytes of one operand image with `rows` rows and K columns of bf16
__host__ __device__ constexpr uint32_t op_bytes(int rows, int K) { return (uint32_t)rows * K * 2; }
__host__ __device__ constexpr uint32_t sbo_of(int K) { return (uint32_t)(K / 8) * LBO_BYTES; }
__device__ __forceinline__ uint32_t op_offset(int row, int k, int K) {
    return (uint32_t)(row >> 3) * sbo_of(K) + (uint32_t)(k >> 3) * LBO_BYTES + (uint32_t)(row & 7) * 16 + (uint32_t)(k & 7) * 2;
}

// Packed weights of one layer in global memory (built every step by tc_pack_weights_kernel):
//   [W_in hi][W_in lo][W_hh hi][W_hh lo] operand images, then bias' (256 floats, see below)
struct PackGeom {
    int kin, kx;                    // true input width, padded K of the x operand (16 or 64)
    uint32_t win_bytes, whh_bytes;  // bytes of ONE image
};
__host__ __device__ inline PackGeom pack_geom(bool layer0, int F) {
    PackGeom g;
    g.kin = layer0 ? F : TH;
    g.kx = kx_of(layer0);
    g.win_bytes = op_bytes(TG, g.kx);
    g.whh_bytes = op_bytes(TG, TH);
    return g;
}
__host__ __device__ inline size_t pack_bytes(bool layer0) {
    return 2 * (size_t)op_bytes(TG, kx_of(layer0)) + 2 * (size_t)op_bytes(TG, TH) + TG * sizeof(float);
}

// bias' folds the bias into the activation argument: sigma(a+b) = 1/(1+2^(-log2e a - log2e b)),
// tanh(a+b) = 1 - 2/(1+2^(2 log2e a + 2 log2e b)); the epilogue computes fma(a, scale, bias').
constexpr float LOG2E = 1.4426950408889634f;

__global__ void __launch_bounds__(256)
tc_pack_weights_kernel(const float* __restrict__ w_ih, const float* __restrict__ w_hh, const float* __restrict__ b_ih,
                       const float* __restrict__ b_hh, int kin, int kx, uint8_t* __restrict__ out) {
    const uint32_t win_bytes = op_bytes(TG, kx), whh_bytes = op_bytes(TG, TH);
    uint8_t* win_hi = out;
    uint8_t* win_lo = out + win_bytes;
    uint8_t* whh_hi = out + 2 * win_bytes;
    uint8_t* whh_lo = whh_hi + whh_bytes;
    float* biasp = reinterpret_cast<float*>(whh_lo + whh_bytes);
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, nthr = gridDim.x * blockDim.x;
    for (int i = tid; i < TG * kx; i += nthr) {
        const int j = i / kx, k = i - j * kx;
        const float v = k < kin ? w_ih[j * kin + k] : 0.f;
        const __nv_bfloat16 h = __float2bfloat16_rn(v);
        const __nv_bfloat16 l = __float2bfloat16_rn(v - __bfloat162float(h));
        const uint32_t o = op_offset(j, k, kx);
        *reinterpret_cast<__nv_bfloat16*>(win_hi + o) = h;
        *reinterpret_cast<__nv_bfloat16*>(win_lo + o) = l;
    }
    for (int i = tid; i < TG * TH; i += nthr) {
        const int j = i / TH, k = i - j * TH;
        const float v = w_hh[i];
        const __nv_bfloat16 h = __float2bfloat16_rn(v);
        const __nv_bfloat16 l = __float2bfloat16_rn(v - __bfloat162float(h));
        const uint32_t o = op_offset(j, k, TH);
        *reinterpret_cast<__nv_bfloat16*>(whh_hi + o) = h;
        *reinterpret_cast<__nv_bfloat16*>(whh_lo + o) = l;
    }
    for (int j = tid; j < TG; j += nthr) {
        const float b = b_ih[j] + b_hh[j];
        const int gate = j / TH;
        biasp[j] = gate == 2 ? 2.0f * LOG2E * b : -LOG2E * b;
    }
}

// --------------------------------------------------------------------------------------------- forward
template <int NSPLIT, int NTILE, bool LAYER0>
struct FwdSmem {
    static constexpr int KX = kx_of(LAYER0);
    static constexpr uint32_t WIN = op_bytes(TG, KX), WHH = op_bytes(TG, TH);
    static constexpr uint32_t XB = op_bytes(TM, KX), HB = op_bytes(TM, TH);
    // offsets (bytes) from the 128-byte aligned base
    static constexpr uint32_t off_win = 0;                                  // [NSPLIT] images
    static constexpr uint32_t off_whh = off_win + NSPLIT * WIN;             // [NSPLIT]
    static constexpr uint32_t off_x = off_whh + NSPLIT * WHH;               // [NTILE][NSPLIT]
    static constexpr uint32_t off_h = off_x + NTILE * NSPLIT * XB;          // [NTILE][NSPLIT]
    static constexpr uint32_t off_bias = off_h + NTILE * NSPLIT * HB;       // 256 floats
    static constexpr uint32_t off_bar = off_bias + TG * 4;                  // 3 * NTILE mbarriers
    static constexpr uint32_t off_tmem = off_bar + 3 * NTILE * 8;
    static constexpr uint32_t total = off_tmem + 16 + 128;                  // + alignment slack
};

template <int NSPLIT, int NTILE, bool LAYER0>
__global__ void __launch_bounds__(32 * (4 * NTILE + LOADER_WARPS + 1), 1)
lstm_tc_fwd_kernel(const uint8_t* __restrict__ packed, const float* __restrict__ xin, int kin, int64_t N, int T,
                   float* __restrict__ h_out, float* __restrict__ c_out, int layer, bool drop, uint64_t seed, float p) {
    using S = FwdSmem<NSPLIT, NTILE, LAYER0>;
    constexpr int KX = S::KX;
    constexpr int EPI_WARPS = 4 * NTILE;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
    float* bias_s = reinterpret_cast<float*>(smem + S::off_bias);
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S::off_bar);      // [slot][0 x_full, 1 h_ready, 2 acc_full]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S::off_tmem);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // ---- one-time set-up: weights -> smem, barriers, TMEM --------------------------------------------
    {
        // the packed image is [win hi][win lo][whh hi][whh lo][bias']; for NSPLIT == 1 only the hi images are staged
        const uint32_t win = S::WIN, whh = S::WHH;
        const uint4* src = reinterpret_cast<const uint4*>(packed);
        uint4* dst = reinterpret_cast<uint4*>(smem);
        for (uint32_t i = threadIdx.x; i < (NSPLIT * win) / 16; i += blockDim.x) dst[S::off_win / 16 + i] = src[i];
        const uint4* src2 = reinterpret_cast<const uint4*>(packed + 2 * win);
        for (uint32_t i = threadIdx.x; i < (NSPLIT * whh) / 16; i += blockDim.x) dst[S::off_whh / 16 + i] = src2[i];
        const float* bsrc = reinterpret_cast<const float*>(packed + 2 * win + 2 * whh);
        for (int i = threadIdx.x; i < TG; i += blockDim.x) bias_s[i] = bsrc[i];
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < NTILE; ++s) {
            mbar_init(&bars[s * 3 + 0], LOADER_WARPS);
            mbar_init(&bars[s * 3 + 1], 4);
            mbar_init(&bars[s * 3 + 2], 1);
        }
        fence_barrier_init();
    }
    if (warp == EPI_WARPS + LOADER_WARPS) tmem_alloc(tmem_slot, NTILE * TG);
    fence_proxy_async();          // weights written with generic stores, read by the tensor core (async proxy)
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const int64_t ntiles = (N + TM - 1) / TM;
    const int64_t ngroups = (ntiles + NTILE - 1) / NTILE;
    const int64_t my_groups = blockIdx.x < ngroups ? (ngroups - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int64_t total_steps = my_groups * T;             // per slot

    if (warp < EPI_WARPS) {
        // ================================ epilogue: thread = one sequence =================================
        const int slot = warp >> 2, wq = warp & 3;
        const int r = wq * 32 + lane;                                        // row in tile = TMEM lane
        const uint32_t taddr = tmem_base + ((uint32_t)(wq * 32) << 16) + (uint32_t)(slot * TG);
        uint8_t* h_hi = smem + S::off_h + (slot * NSPLIT) * S::HB;
        uint8_t* h_lo = h_hi + S::HB;
        uint64_t* h_ready = &bars[slot * 3 + 1];
        uint64_t* acc_full = &bars[slot * 3 + 2];
        const uint32_t row_off = (uint32_t)(r >> 3) * sbo_of(TH) + (uint32_t)(r & 7) * 16;
        float cst[TH];
        int64_t g = 0;
        for (int64_t grp = blockIdx.x; grp < ngroups; grp += gridDim.x) {
            const int64_t tile = grp * NTILE + slot;
            const int64_t row = tile * TM + r;
            const bool live = tile < ntiles && row < N;
#pragma unroll
            for (int u = 0; u < TH; ++u) cst[u] = 0.f;
            for (int t = 0; t < T; ++t, ++g) {
                mbar_wait(acc_full, (uint32_t)(g & 1));
                tc_fence_after();
                float* hdst = h_out + ((int64_t)t * N + row) * TH;
                float* cdst = c_out + ((int64_t)t * N + row) * TH;
#pragma unroll
                for (int uc = 0; uc < TH / 8; ++uc) {
                    float ai[8], af[8], ag[8], ao[8];
                    tmem_ld8(taddr + 0 * TH + uc * 8, ai);
                    tmem_ld8(taddr + 1 * TH + uc * 8, af);
                    tmem_ld8(taddr + 2 * TH + uc * 8, ag);
                    tmem_ld8(taddr + 3 * TH + uc * 8, ao);
                    tmem_ld_wait();
                    float hv[8], cv[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int u = uc * 8 + j;
                        // sigma(x) = 1/(1 + 2^(-log2e x)),  tanh(x) = 1 - 2/(1 + 2^(2 log2e x)); bias folded into bias_s
                        const float ei = exp2f(fmaf(ai[j], -LOG2E, bias_s[0 * TH + u]));
                        const float ef = exp2f(fmaf(af[j], -LOG2E, bias_s[1 * TH + u]));
                        const float eg = exp2f(fmaf(ag[j], 2.f * LOG2E, bias_s[2 * TH + u]));
                        const float eo = exp2f(fmaf(ao[j], -LOG2E, bias_s[3 * TH + u]));
                        const float ig = __fdividef(1.f, 1.f + ei), fg = __fdividef(1.f, 1.f + ef);
                        const float gg = 1.f - __fdividef(2.f, 1.f + eg), og = __fdividef(1.f, 1.f + eo);
                        const float c = fmaf(fg, cst[u], ig * gg);
                        const float ec = exp2f(2.f * LOG2E * c);
                        const float h = og * (1.f - __fdividef(2.f, 1.f + ec));
                        cst[u] = c;
                        hv[j] = h;
                        cv[j] = c;
                    }
                    uint4 hi4, lo4;
                    split2(hv[0], hv[1], hi4.x, lo4.x);
                    split2(hv[2], hv[3], hi4.y, lo4.y);
                    split2(hv[4], hv[5], hi4.z, lo4.z);
                    split2(hv[6], hv[7], hi4.w, lo4.w);
                    *reinterpret_cast<uint4*>(h_hi + row_off + uc * LBO_BYTES) = hi4;
                    if (NSPLIT == 2) *reinterpret_cast<uint4*>(h_lo + row_off + uc * LBO_BYTES) = lo4;
                    if (live) {
                        st_global_v8(hdst + uc * 8, hv);
                        st_global_v8(cdst + uc * 8, cv);
                    }
                }
                tc_fence_before();            // TMEM reads of this step are complete (wait::ld above)
                fence_proxy_async();          // h_t visible to the tensor core
                __syncwarp();
                if (lane == 0) mbar_arrive(h_ready);
            }
        }
    } else if (warp < EPI_WARPS + LOADER_WARPS) {
        // ================================ loader: x_t operand of every slot ==================================
        const int lw = warp - EPI_WARPS;
        const uint32_t thr = dropout_threshold(p);
        const float inv_keep = 1.0f / (1.0f - p);
        int64_t g = 0;
        for (int64_t grp = blockIdx.x; grp < ngroups; grp += gridDim.x) {
            for (int t = 0; t < T; ++t, ++g) {
#pragma unroll
                for (int slot = 0; slot < NTILE; ++slot) {
                    const int64_t tile = grp * NTILE + slot;
                    uint8_t* x_hi = smem + S::off_x + (slot * NSPLIT) * S::XB;
                    uint8_t* x_lo = x_hi + S::XB;
                    if (g > 0) mbar_wait(&bars[slot * 3 + 2], (uint32_t)((g - 1) & 1));   // MMA of the previous step consumed x
                    if (LAYER0) {
                        // 128 rows x [x0..x_{kin-1}, 0...]: lane handles rows lw*64 + lane and +32
#pragma unroll
                        for (int rr = 0; rr < 2; ++rr) {
                            const int r = lw * 64 + rr * 32 + lane;
                            const int64_t row = tile * TM + r;
                            float v[8];
#pragma unroll
                            for (int k = 0; k < 8; ++k) v[k] = 0.f;
                            if (tile < ntiles && row < N) {
                                const float* src = xin + ((int64_t)t * N + row) * kin;
                                for (int k = 0; k < kin; ++k) v[k] = src[k];
                            }
                            uint4 hi4, lo4;
                            split2(v[0], v[1], hi4.x, lo4.x);
                            split2(v[2], v[3], hi4.y, lo4.y);
                            split2(v[4], v[5], hi4.z, lo4.z);
                            split2(v[6], v[7], hi4.w, lo4.w);
                            const uint32_t ro = (uint32_t)(r >> 3) * sbo_of(KX) + (uint32_t)(r & 7) * 16;
                            const uint4 z = make_uint4(0, 0, 0, 0);
                            *reinterpret_cast<uint4*>(x_hi + ro) = hi4;
                            *reinterpret_cast<uint4*>(x_hi + ro + LBO_BYTES) = z;
                            if (NSPLIT == 2) {
                                *reinterpret_cast<uint4*>(x_lo + ro) = lo4;
                                *reinterpret_cast<uint4*>(x_lo + ro + LBO_BYTES) = z;
                            }
                        }
                    } else {
                        // 16 row groups of 8 rows; this warp takes 8 of them.  lane = (row in group, quarter)
                        const int rg_r = lane >> 2, q = lane & 3;
#pragma unroll 2
                        for (int rgi = 0; rgi < 8; ++rgi) {
                            const int rg = lw * 8 + rgi;
                            const int r = rg * 8 + rg_r;
                            const int64_t row = tile * TM + r;
                            const bool live = tile < ntiles && row < N;
                            const float* src = xin + ((int64_t)t * N + row) * TH;
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const int k = 4 * q + 16 * i;
                                float4 v = live ? *reinterpret_cast<const float4*>(src + k) : make_float4(0.f, 0.f, 0.f, 0.f);
                                if (drop) {
                                    U4 b = dropout_block(seed, layer, row, t, k >> 3);
                                    const int j0 = k & 7;
                                    v.x = dropout_lane16(b, j0 + 0) >= thr ? v.x * inv_keep : 0.f;
                                    v.y = dropout_lane16(b, j0 + 1) >= thr ? v.y * inv_keep : 0.f;
                                    v.z = dropout_lane16(b, j0 + 2) >= thr ? v.z * inv_keep : 0.f;
                                    v.w = dropout_lane16(b, j0 + 3) >= thr ? v.w * inv_keep : 0.f;
                                }
                                uint2 hi2, lo2;
                                split2(v.x, v.y, hi2.x, lo2.x);
                                split2(v.z, v.w, hi2.y, lo2.y);
                                const uint32_t o = (uint32_t)rg * sbo_of(KX) + (uint32_t)(k >> 3) * LBO_BYTES + (uint32_t)rg_r * 16 +
                                                   (uint32_t)(k & 7) * 2;
                                *reinterpret_cast<uint2*>(x_hi + o) = hi2;
                                if (NSPLIT == 2) *reinterpret_cast<uint2*>(x_lo + o) = lo2;
                            }
                        }
                    }
                    fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&bars[slot * 3 + 0]);
                }
            }
        }
    } else {
        // ================================ MMA issuer (one thread) =============================================
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_bf16(TM, TG);
            const uint32_t s_base = smem_u32(smem);
            for (int64_t g = 0; g < total_steps; ++g) {
                const bool first = (g % T) == 0;              // t == 0: h_{-1} = 0, skip the recurrent K-steps
#pragma unroll
                for (int slot = 0; slot < NTILE; ++slot) {
                    mbar_wait(&bars[slot * 3 + 0], (uint32_t)(g & 1));                    // x_t staged
                    if (g > 0) mbar_wait(&bars[slot * 3 + 1], (uint32_t)((g - 1) & 1));   // h_{t-1} staged, TMEM drained
                    tc_fence_after();
                    const uint32_t d = tmem_base + (uint32_t)(slot * TG);
                    const uint32_t x_hi = s_base + S::off_x + (slot * NSPLIT) * S::XB, x_lo = x_hi + S::XB;
                    const uint32_t h_hi = s_base + S::off_h + (slot * NSPLIT) * S::HB, h_lo = h_hi + S::HB;
                    const uint32_t win_hi = s_base + S::off_win, win_lo = win_hi + S::WIN;
                    const uint32_t whh_hi = s_base + S::off_whh, whh_lo = whh_hi + S::WHH;
                    uint32_t acc = 0;
#pragma unroll
                    for (int ks = 0; ks < KX / 16; ++ks) {
                        const uint32_t ko = ks * 2 * LBO_BYTES;
                        umma_bf16(d, make_desc(x_hi + ko, LBO_BYTES, sbo_of(KX)), make_desc(win_hi + ko, LBO_BYTES, sbo_of(KX)), idesc, acc);
                        acc = 1;
                        if (NSPLIT == 2) {
                            umma_bf16(d, make_desc(x_lo + ko, LBO_BYTES, sbo_of(KX)), make_desc(win_hi + ko, LBO_BYTES, sbo_of(KX)), idesc, 1);
                            umma_bf16(d, make_desc(x_hi + ko, LBO_BYTES, sbo_of(KX)), make_desc(win_lo + ko, LBO_BYTES, sbo_of(KX)), idesc, 1);
                        }
                    }
                    if (!first) {
#pragma unroll
                        for (int ks = 0; ks < TH / 16; ++ks) {
                            const uint32_t ko = ks * 2 * LBO_BYTES;
                            umma_bf16(d, make_desc(h_hi + ko, LBO_BYTES, sbo_of(TH)), make_desc(whh_hi + ko, LBO_BYTES, sbo_of(TH)), idesc, 1);
                            if (NSPLIT == 2) {
                                umma_bf16(d, make_desc(h_lo + ko, LBO_BYTES, sbo_of(TH)), make_desc(whh_hi + ko, LBO_BYTES, sbo_of(TH)), idesc, 1);
                                umma_bf16(d, make_desc(h_hi + ko, LBO_BYTES, sbo_of(TH)), make_desc(whh_lo + ko, LBO_BYTES, sbo_of(TH)), idesc, 1);
                            }
                        }
                    }
                    umma_commit(&bars[slot * 3 + 2]);
                }
            }
        }
        __syncwarp();
    }
    // ---- teardown ----------------------------------------------------------------------------------------
    tc_fence_before();
    __syncthreads();
    if (warp == EPI_WARPS + LOADER_WARPS) tmem_dealloc(tmem_base, NTILE * TG);
}

// --------------------------------------------------------------------------------------------- self test
// One 128 x 256 x K GEMM through exactly the descriptors, instruction and TMEM load path of the kernel above.
__global__ void __launch_bounds__(128, 1)
tc_selftest_kernel(const float* __restrict__ A, const float* __restrict__ B, int K, float* __restrict__ D) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~(uintptr_t)127);
    uint8_t* a_hi = smem;
    uint8_t* a_lo = a_hi + op_bytes(TM, K);
    uint8_t* b_hi = a_lo + op_bytes(TM, K);
    uint8_t* b_lo = b_hi + op_bytes(TG, K);
    uint64_t* bar = reinterpret_cast<uint64_t*>(b_lo + op_bytes(TG, K));
    uint32_t* tslot = reinterpret_cast<uint32_t*>(bar + 1);
    for (int i = threadIdx.x; i < TM * K; i += blockDim.x) {
        const int r = i / K, k = i - r * K;
        const float v = A[i];
        const __nv_bfloat16 h = __float2bfloat16_rn(v);
        *reinterpret_cast<__nv_bfloat16*>(a_hi + op_offset(r, k, K)) = h;
        *reinterpret_cast<__nv_bfloat16*>(a_lo + op_offset(r, k, K)) = __float2bfloat16_rn(v - __bfloat162float(h));
    }
    for (int i = threadIdx.x; i < TG * K; i += blockDim.x) {
        const int r = i / K, k = i - r * K;
        const float v = B[i];
        const __nv_bfloat16 h = __float2bfloat16_rn(v);
        *reinterpret_cast<__nv_bfloat16*>(b_hi + op_offset(r, k, K)) = h;
        *reinterpret_cast<__nv_bfloat16*>(b_lo + op_offset(r, k, K)) = __float2bfloat16_rn(v - __bfloat162float(h));
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
    if (warp == 0) tmem_alloc(tslot, TG);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tslot;
    if (threadIdx.x == 0) {
        constexpr uint32_t idesc = make_idesc_bf16(TM, TG);
        uint32_t acc = 0;
        for (int ks = 0; ks < K / 16; ++ks) {
            const uint32_t ko = ks * 2 * LBO_BYTES;
            umma_bf16(tmem_base, make_desc(smem_u32(a_hi) + ko, LBO_BYTES, sbo_of(K)), make_desc(smem_u32(b_hi) + ko, LBO_BYTES, sbo_of(K)), idesc, acc);
            acc = 1;
            umma_bf16(tmem_base, make_desc(smem_u32(a_lo) + ko, LBO_BYTES, sbo_of(K)), make_desc(smem_u32(b_hi) + ko, LBO_BYTES, sbo_of(K)), idesc, 1);
            umma_bf16(tmem_base, make_desc(smem_u32(a_hi) + ko, LBO_BYTES, sbo_of(K)), make_desc(smem_u32(b_lo) + ko, LBO_BYTES, sbo_of(K)), idesc, 1);
        }
        umma_commit(bar);
    }
    mbar_wait(bar, 0);
    tc_fence_after();
    const int r = warp * 32 + lane;
    for (int c0 = 0; c0 < TG; c0 += 8) {
        float v[8];
        tmem_ld8(tmem_base + ((uint32_t)(warp * 32) << 16) + c0, v);
        tmem_ld_wait();
        for (int j = 0; j < 8; ++j) D[r * TG + c0 + j] = v[j];
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, TG);
}

// --------------------------------------------------------------------------------------------- host side
size_t ffma_workspace_bytes(int64_t N, int T, int L);
int ffma_bwd_layer(int layer, const float* params, int F, int H, int L, const float* x, int64_t N, int T,
                   const float* stash_h, const float* stash_c, const float* dhT, float* grads, float p, uint64_t seed,
                   int training, void* workspace, size_t ws_bytes, cudaStream_t st);

bool tc_available() { return true; }

static size_t tc_pack_total(int L) {
    size_t tot = 0;
    for (int l = 0; l < L; ++l) tot += (pack_bytes(l == 0) + 255) & ~(size_t)255;
    return tot;
}

// workspace = [FFMA backward workspace][packed weights of every layer]
size_t tc_workspace_bytes(int mode, int64_t N, int T, int F, int H, int L) {
    (void)mode; (void)F; (void)H;
    return ((ffma_workspace_bytes(N, T, L) + 255) & ~(size_t)255) + tc_pack_total(L);
}

template <int NSPLIT, int NTILE, bool LAYER0>
static int launch_fwd(const uint8_t* packed, const float* xin, int kin, int64_t N, int T, float* h_out, float* c_out,
                      int layer, bool drop, uint64_t seed, float p, cudaStream_t st) {
    using S = FwdSmem<NSPLIT, NTILE, LAYER0>;
    auto kern = lstm_tc_fwd_kernel<NSPLIT, NTILE, LAYER0>;
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::total));
    const int64_t ntiles = (N + TM - 1) / TM, ngroups = (ntiles + NTILE - 1) / NTILE;
    const int grid = (int)(ngroups < num_sms() ? ngroups : num_sms());
    const int threads = 32 * (4 * NTILE + LOADER_WARPS + 1);
    {
        TimedLaunch _t("lstm_tc_fwd_kernel", st);
        kern<<<grid, threads, S::total, st>>>(packed, xin, kin, N, T, h_out, c_out, layer, drop, seed, p);
    }
    B200_LAUNCH_CHECK("lstm_tc_fwd_kernel");
    return B200_OK;
}

int tc_fwd(int mode, const float* params, int F, int H, int L, const float* x, int64_t N, int T, float* stash_h,
           float* stash_c, float p, uint64_t seed, int training, void* workspace, size_t ws_bytes, cudaStream_t st) {
    if (H != TH) { set_error("tcgen05 engine supports hidden=64 only, got %d", H); return B200_ERR_UNSUPPORTED; }
    B200_CHECK_ARG(workspace != nullptr, "lstm_fwd (tcgen05): workspace required for the packed weights");
    const size_t need = tc_pack_total(L);
    B200_CHECK_ARG(ws_bytes >= need, "lstm_fwd (tcgen05): workspace too small (%zu < %zu)", ws_bytes, need);
    // packed weights live at the END of the workspace so that the backward workspace layout is untouched
    uint8_t* pk = (uint8_t*)workspace + ((ws_bytes - need) & ~(size_t)255);
    const bool drop = training && p > 0.f;
    const int64_t lsz = (int64_t)T * N * TH;
    const bool precise = mode == B200_MODE_TC_PRECISE;
    for (int l = 0; l < L; ++l) {
        const int kin = l == 0 ? F : TH;
        const float* base = params + b200_param_offset(F, TH, L, l);
        const float* w_ih = base;
        const float* w_hh = w_ih + (int64_t)TG * kin;
        const float* b_ih = w_hh + (int64_t)TG * TH;
        const float* b_hh = b_ih + TG;
        { TimedLaunch _t("tc_pack_weights_kernel", st); tc_pack_weights_kernel<<<32, 256, 0, st>>>(w_ih, w_hh, b_ih, b_hh, kin, kx_of(l == 0), pk); }
        B200_LAUNCH_CHECK("tc_pack_weights_kernel");
        const float* xin = l == 0 ? x : stash_h + (l - 1) * lsz;
        float* ho = stash_h + l * lsz;
        float* co = stash_c + l * lsz;
        int rc;
        if (l == 0) {
            rc = precise ? launch_fwd<2, 2, true>(pk, xin, kin, N, T, ho, co, l, false, seed, p, st)
                         : launch_fwd<1, 2, true>(pk, xin, kin, N, T, ho, co, l, false, seed, p, st);
        } else {
            rc = precise ? launch_fwd<2, 1, false>(pk, xin, kin, N, T, ho, co, l, drop, seed, p, st)
                         : launch_fwd<1, 2, false>(pk, xin, kin, N, T, ho, co, l, drop, seed, p, st);
        }
        if (rc) return rc;
        pk += (pack_bytes(l == 0) + 255) & ~(size_t)255;
    }
    return B200_OK;
}

int tc_bwd_layer(int mode, int layer, const float* params, int F, int H, int L, const float* x, int64_t N, int T,
                 const float* stash_h, const float* stash_c, const float* dhT, float* grads, float p, uint64_t seed,
                 int training, void* workspace, size_t ws_bytes, cudaStream_t st) {
    (void)mode;
    // backward currently runs on the fp32 CUDA-core engine (same stash format); reserve the packed-weight tail
    const size_t tail = tc_pack_total(L);
    const size_t usable = ws_bytes > tail ? (ws_bytes - tail) & ~(size_t)255 : 0;
    return ffma_bwd_layer(layer, params, F, H, L, x, N, T, stash_h, stash_c, dhT, grads, p, seed, training, workspace, usable, st);
}

}  // namespace b200

using namespace b200;

// D[128][256] = A[128][K] * B[256][K]^T through the tcgen05 path (bf16x3 split); K multiple of 16, K <= 64.
extern "C" int b200_tc_selftest(const float* A, const float* B, int K, float* D, void* stream) {
    B200_CHECK_ARG(A && B && D && K >= 16 && K <= 64 && K % 16 == 0, "tc_selftest: bad arguments");
    const size_t smem = 2 * (size_t)op_bytes(TM, K) + 2 * (size_t)op_bytes(TG, K) + 64 + 128;
    B200_CUDA(cudaFuncSetAttribute(tc_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    tc_selftest_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(A, B, K, D);
    B200_LAUNCH_CHECK("tc_selftest_kernel");
    return B200_OK;
}
