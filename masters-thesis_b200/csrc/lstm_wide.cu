// Wide-model LSTM engine: hidden sizes other than 64 (BASELINE.json configs[2]: hidden 256, 8 layers).
//
// At H = 256 one layer's weights (W_ih, W_hh: 2 x 1024 x 256) no longer fit one SM's shared memory next to the
// operands, and a 128-row tile's gates (128 x 1024 fp32) no longer fit TMEM, so the persistent fused kernels of
// lstm_tc.cu do not apply.  This engine is the cuDNN-style decomposition on hand-written tcgen05 GEMMs:
//
//   forward  layer l:  G[T*N x 4H]  = dropout(h^{l-1})[T*N x H] . W_ih^T            one GEMM over all steps (l > 0)
//            step t:   G_t[N x 4H] += h_{t-1}[N x H] . W_hh^T                       GEMM (accumulate), t > 0
//                      cell kernel: + bias (+ x_t . W_ih^T for layer 0, K = F <= 8 on CUDA cores), activations,
//                                   c_t, h_t; G_t is overwritten with the ACTIVATED gates (the backward stash)
//   backward step t:   cell kernel: (dh_t, dc_t, gates, c) -> dG_t (overwrites G_t), dc_{t-1}
//                      dh_{t-1}[N x H] = dG_t[N x 4H] . W_hh                        GEMM, B operand read MN-major
//            layer l:  dW_hh = dG[1..T-1]^T . h[0..T-2]     dW_ih = dG^T . dropout(h^{l-1})    split-K GEMMs, both MN-major
//                      dX = dG . W_ih (then the dropout mask) = d_out of layer l-1;   db = column sums of dG
//
// One GEMM kernel serves all of it: C[M x N] (+)= op(A) . op(B) with fp32 operands in global memory that are split
// into bf16 hi (+ lo) images while being staged (the "interleaved" UMMA layout of tc_common.cuh), fp32 accumulation
// in TMEM, 128 x 256 output tiles, K chunks of 64 double-buffered in shared memory: all threads stage chunk k+1
// while the tensor core works on chunk k.  NSPLIT = 2 (B200_MODE_TC_PRECISE) issues hi*hi + lo*hi + hi*lo, NSPLIT = 1
// (B200_MODE_TC_FAST) is the plain bf16 GEMM with fp32 accumulation that configs[2] names.
//
// Replaces, for hidden != 64: nn.LSTM forward/backward (src/model.py:88-109, autograd), as lstm_tc.cu does for 64.
#include "common.cuh"
#include "tc_common.cuh"

namespace b200 {
using namespace tc;

namespace {

constexpr int GM = 128;            // output tile rows = UMMA M
constexpr int GN = 256;            // output tile columns (max UMMA N)
constexpr int GK = 64;             // K chunk per pipeline stage
constexpr int G_THREADS = 256;
constexpr int MAX_SPLITS = 32;

// operand variants (how A and B are stored in global memory, all row-major fp32)
//   0: A [M x K], B [N x K]      C = A . B^T          both images K-major
//   1: A [M x K], B [K x N]      C = A . B            B image read MN-major
//   2: A [K x M], B [K x N]      C = A^T . B          both images read MN-major (K = rows: split-K over CTAs)
struct GemmArgs {
    const float* A; const float* B; float* C;
    int64_t M; int N; int64_t K;
    int64_t lda, ldb, ldc;
    int accumulate;                // C += (ignored when part != nullptr)
    int64_t k_per_split;           // multiple of GK
    float* part;                   // split-K partials [splits][M x N] (variant 2), or nullptr
    // pre-packed B (variants 0, 1): operand images of every (N tile, K chunk) as wide_pack_b_kernel lays them out; then B is
    // staged by the TMA engine (one bulk copy per image) instead of by the threads
    const uint8_t* bp_hi; const uint8_t* bp_lo;
    // fused LSTM cell (forward recurrent step, FUSE kernels only): C is the gate buffer G_t [M][4H] in torch gate order holding
    // the input projection (has_x) and receiving the ACTIVATED gates; one N tile = the four gates of 64 hidden units
    int H; int has_x; int F;
    const float* b_ih; const float* b_hh; const float* c_prev; float* c_out; float* h_out;
    const float* x_t; const float* w_ih0;
};

__device__ __forceinline__ int blk_rr(int lane) { return (lane >> 1) & 7; }
__device__ __forceinline__ int blk_q(int lane) { return (lane & 1) | ((lane >> 3) & 2); }

// A warp moves NRG row groups (8 rows each) x 64 columns of a row-major fp32 matrix (leading dimension ld) into an
// operand image: lane = (row in group, column quarter); rows >= rows_live read as zero.
template <int NRG>
__device__ __forceinline__ void blk_load(float4 (&v)[NRG][4], const float* base, int64_t ld, int64_t rows_live, int rg0, int lane) {
    const int rr = blk_rr(lane), q = blk_q(lane);
#pragma unroll
    for (int rgi = 0; rgi < NRG; ++rgi) {
        const int r = (rg0 + rgi) * 8 + rr;
        const float* src = base + (int64_t)r * ld + 4 * q;
#pragma unroll
        for (int i = 0; i < 4; ++i)
            v[rgi][i] = r < rows_live ? *reinterpret_cast<const float4*>(src + 16 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
}
// image: rows x cols bf16, 8 x 8 core matrices of 128 contiguous bytes, sbo = bytes between 8-row groups
template <int NSPLIT, int NRG>
__device__ __forceinline__ void blk_store(const float4 (&v)[NRG][4], uint8_t* img_hi, uint8_t* img_lo, uint32_t sbo, int col0,
                                          int rg0, int lane) {
    const int rr = blk_rr(lane), q = blk_q(lane);
#pragma unroll
    for (int rgi = 0; rgi < NRG; ++rgi) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int c = col0 + 4 * q + 16 * i;
            const float4 x = v[rgi][i];
            uint2 hi2, lo2;
            split2(x.x, x.y, hi2.x, lo2.x);
            split2(x.z, x.w, hi2.y, lo2.y);
            const uint32_t o = (uint32_t)(rg0 + rgi) * sbo + (uint32_t)(c >> 3) * LBO_BYTES + (uint32_t)rr * 16 + (uint32_t)(c & 7) * 2;
            *reinterpret_cast<uint2*>(img_hi + o) = hi2;
            if (NSPLIT == 2) *reinterpret_cast<uint2*>(img_lo + o) = lo2;
        }
    }
}

template <int NSPLIT>
struct GemmSmem {
    static constexpr uint32_t A_IMG = GM * GK * 2, B_IMG = GN * GK * 2;           // one bf16 image
    static constexpr uint32_t STAGE = NSPLIT * (A_IMG + B_IMG);
    static constexpr uint32_t off_bar = 2 * STAGE;                                // 2 stage barriers + done + 2 B-full
    static constexpr uint32_t off_tmem = off_bar + 5 * 8;
    static constexpr uint32_t total = off_tmem + 16 + 128;
};

__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

template <int NSPLIT, int VAR, bool FUSE = false>
__global__ void __launch_bounds__(G_THREADS, 1) wide_gemm_kernel(GemmArgs a) {
    using S = GemmSmem<NSPLIT>;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + S::off_bar);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S::off_tmem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // the N tiles of one M tile are neighbouring CTAs (they run at the same time and share the A rows through L2: with the M
    // tile as the fast index the input-projection GEMM read A four times from DRAM, ncu 6.3 GB against 1.6 GB)
    const int ntiles = (a.N + GN - 1) / GN;
    const int nt_idx = (int)(blockIdx.x % ntiles);
    const int64_t m0 = (int64_t)(blockIdx.x / ntiles) * GM;
    const int n0 = nt_idx * GN;
    const int NT = a.N - n0 < GN ? a.N - n0 : GN;          // multiple of 64
    const int64_t k_beg = (int64_t)blockIdx.z * a.k_per_split;
    const int64_t k_end = k_beg + a.k_per_split < a.K ? k_beg + a.k_per_split : a.K;
    const int nchunks = (int)((k_end - k_beg + GK - 1) / GK);

    if (threadIdx.x == 0) {
        mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); mbar_init(&bars[2], 1); mbar_init(&bars[3], 1); mbar_init(&bars[4], 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(tmem_slot, GN);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const uint32_t sbo_k = (uint32_t)(GK / 8) * LBO_BYTES;           // K-major image of [rows x 64]
    const uint32_t sbo_am = (uint32_t)(GM / 8) * LBO_BYTES;          // MN-major A image [64 x 128]
    const uint32_t sbo_bm = (uint32_t)(NT / 8) * LBO_BYTES;          // MN-major B image [64 x NT]
    const uint32_t idesc = make_idesc_bf16(GM, NT, VAR == 2 ? 1 : 0, VAR == 0 ? 0 : 1);

    for (int kc = 0; kc < nchunks; ++kc) {
        const int s = kc & 1;
        uint8_t* a_hi = smem + s * S::STAGE;
        uint8_t* a_lo = a_hi + S::A_IMG;
        uint8_t* b_hi = a_hi + NSPLIT * S::A_IMG;
        uint8_t* b_lo = b_hi + S::B_IMG;
        const int64_t k0 = k_beg + (int64_t)kc * GK;
        if (kc >= 2) mbar_wait(&bars[s], (uint32_t)(((kc >> 1) - 1) & 1));      // MMAs of chunk kc - 2 have read this stage
        // ---- stage B ------------------------------------------------------------------------------------------
        if (VAR != 2 && a.bp_hi) {            // pre-packed images: one bulk copy each, issued before the threads stage A
            if (threadIdx.x == 0) {
                const uint32_t img = (uint32_t)NT * GK * 2;
                const int64_t chunk = k0 / GK, nchunk_all = a.K / GK;
                // image index: variant 0 [N tile][K chunk], variant 1 [K chunk][N tile]; every image padded to GN x GK
                const int64_t idx = VAR == 0 ? (int64_t)nt_idx * nchunk_all + chunk : chunk * ntiles + nt_idx;
                mbar_arrive_expect_tx(&bars[3 + s], NSPLIT * img);
                bulk_copy_g2s(b_hi, a.bp_hi + idx * S::B_IMG, img, &bars[3 + s]);
                if (NSPLIT == 2) bulk_copy_g2s(b_lo, a.bp_lo + idx * S::B_IMG, img, &bars[3 + s]);
            }
        } else if (VAR == 0) {                       // [NT rows x 64 k]: NT / 8 row groups, pairs round-robin over the warps
            for (int pr = warp; pr < NT / 16; pr += G_THREADS / 32) {
                float4 v[2][4];
                blk_load<2>(v, a.B + (int64_t)n0 * a.ldb + k0, a.ldb, NT, 2 * pr, lane);
                blk_store<NSPLIT, 2>(v, b_hi, b_lo, sbo_k, 0, 2 * pr, lane);
            }
        } else {                              // [64 k-rows x NT cols]: (column block, row-group pair) units
            for (int u = warp; u < (NT / 64) * 4; u += G_THREADS / 32) {
                const int cb = u >> 2, pr = u & 3;
                float4 v[2][4];
                blk_load<2>(v, a.B + k0 * a.ldb + n0 + cb * 64, a.ldb, k_end - k0, 2 * pr, lane);
                blk_store<NSPLIT, 2>(v, b_hi, b_lo, sbo_bm, cb * 64, 2 * pr, lane);
            }
        }
        // ---- stage A ------------------------------------------------------------------------------------------
        if (VAR != 2) {                       // [128 rows x 64 k]: 16 row groups, 2 per warp
            float4 v[2][4];
            blk_load<2>(v, a.A + m0 * a.lda + k0, a.lda, a.M - m0, 2 * warp, lane);
            blk_store<NSPLIT, 2>(v, a_hi, a_lo, sbo_k, 0, 2 * warp, lane);
        } else {                              // [64 k-rows x 128 m-cols]: 2 column blocks x 4 row-group pairs
            const int cb = warp >> 2, pr = warp & 3;
            float4 v[2][4];
            blk_load<2>(v, a.A + k0 * a.lda + m0 + cb * 64, a.lda, k_end - k0, 2 * pr, lane);
            blk_store<NSPLIT, 2>(v, a_hi, a_lo, sbo_am, cb * 64, 2 * pr, lane);
        }
        fence_proxy_async();
        __syncthreads();
        if (threadIdx.x == 0) {
            if (VAR != 2 && a.bp_hi) mbar_wait(&bars[3 + s], (uint32_t)((kc >> 1) & 1));      // B images have landed
            tc_fence_after();
            const uint32_t ah = smem_u32(a_hi), al = smem_u32(a_lo), bh = smem_u32(b_hi), bl = smem_u32(b_lo);
#pragma unroll
            for (int ks = 0; ks < GK / 16; ++ks) {
                uint64_t dah, dal, dbh, dbl;
                if (VAR != 2) { dah = make_desc(ah + ks * 2 * LBO_BYTES, LBO_BYTES, sbo_k); dal = make_desc(al + ks * 2 * LBO_BYTES, LBO_BYTES, sbo_k); }
                else { dah = make_desc(ah + ks * 2 * sbo_am, sbo_am, LBO_BYTES); dal = make_desc(al + ks * 2 * sbo_am, sbo_am, LBO_BYTES); }
                if (VAR == 0) { dbh = make_desc(bh + ks * 2 * LBO_BYTES, LBO_BYTES, sbo_k); dbl = make_desc(bl + ks * 2 * LBO_BYTES, LBO_BYTES, sbo_k); }
                else { dbh = make_desc(bh + ks * 2 * sbo_bm, sbo_bm, LBO_BYTES); dbl = make_desc(bl + ks * 2 * sbo_bm, sbo_bm, LBO_BYTES); }
                umma_bf16(tmem_base, dah, dbh, idesc, (kc > 0 || ks > 0) ? 1u : 0u);
                if (NSPLIT == 2) {
                    umma_bf16(tmem_base, dal, dbh, idesc, 1);
                    umma_bf16(tmem_base, dah, dbl, idesc, 1);
                }
            }
            umma_commit(&bars[s]);
            if (kc == nchunks - 1) umma_commit(&bars[2]);
        }
    }
    // ---- epilogue: TMEM -> C (warp = TMEM lane quarter x column half) ------------------------------------------
    if (nchunks > 0) {
        mbar_wait(&bars[2], 0u);
        tc_fence_after();
    }
    if (FUSE) {
        // accumulator columns of this tile: [i | f | g | o] x 64 units (gate-interleaved weight images, pack variant 3);
        // warp = TMEM lane quarter x unit half; the cell update of wide_cell_fwd_kernel on the accumulators
        const int lq = warp & 3, uh = warp >> 2, H = a.H;
        const int64_t row = m0 + lq * 32 + lane;
        const uint32_t tl = tmem_base + ((uint32_t)(lq * 32) << 16);
#pragma unroll 1
        for (int uc = 0; uc < 4; ++uc) {
            const int ul = uh * 32 + uc * 8;                      // first unit of the chunk inside the tile
            float acc[4][8];
#pragma unroll
            for (int k = 0; k < 4; ++k) tmem_ld8(tl + (uint32_t)(k * 64 + ul), acc[k]);
            tmem_ld_wait();
            if (row < a.M) {
                const int u0 = nt_idx * 64 + ul;                  // global hidden unit
                float* g = a.C + row * a.ldc;
                float cp[8], cv[8], hv[8];
                if (a.c_prev) {
                    const float4* c4 = reinterpret_cast<const float4*>(a.c_prev + row * H + u0);
                    const float4 c0 = c4[0], c1 = c4[1];
                    cp[0] = c0.x; cp[1] = c0.y; cp[2] = c0.z; cp[3] = c0.w; cp[4] = c1.x; cp[5] = c1.y; cp[6] = c1.z; cp[7] = c1.w;
                } else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) cp[j] = 0.f;
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int col = k * H + u0;
                    if (a.has_x) {
                        const float4* x4 = reinterpret_cast<const float4*>(g + col);
                        const float4 x0 = x4[0], x1 = x4[1];
                        acc[k][0] += x0.x; acc[k][1] += x0.y; acc[k][2] += x0.z; acc[k][3] += x0.w;
                        acc[k][4] += x1.x; acc[k][5] += x1.y; acc[k][6] += x1.z; acc[k][7] += x1.w;
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        float v = acc[k][j] + (a.b_ih[col + j] + a.b_hh[col + j]);
                        if (a.w_ih0) for (int f = 0; f < a.F; ++f) v = fmaf(a.x_t[row * a.F + f], a.w_ih0[(int64_t)(col + j) * a.F + f], v);
                        acc[k][j] = k == 2 ? tanhf(v) : sigmoidf_(v);
                    }
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    cv[j] = acc[1][j] * cp[j] + acc[0][j] * acc[2][j];
                    hv[j] = acc[3][j] * tanhf(cv[j]);
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    float4* o4 = reinterpret_cast<float4*>(g + k * H + u0);
                    o4[0] = make_float4(acc[k][0], acc[k][1], acc[k][2], acc[k][3]);
                    o4[1] = make_float4(acc[k][4], acc[k][5], acc[k][6], acc[k][7]);
                }
                float4* co = reinterpret_cast<float4*>(a.c_out + row * H + u0);
                co[0] = make_float4(cv[0], cv[1], cv[2], cv[3]); co[1] = make_float4(cv[4], cv[5], cv[6], cv[7]);
                float4* ho = reinterpret_cast<float4*>(a.h_out + row * H + u0);
                ho[0] = make_float4(hv[0], hv[1], hv[2], hv[3]); ho[1] = make_float4(hv[4], hv[5], hv[6], hv[7]);
            }
        }
    } else {
        const int lq = warp & 3, ch = warp >> 2;
        const int64_t row = m0 + lq * 32 + lane;
        float* crow = a.part ? a.part + ((int64_t)blockIdx.z * a.M + row) * a.N + n0 : a.C + row * a.ldc + n0;
        for (int c0 = ch * (NT / 2); c0 < (ch + 1) * (NT / 2); c0 += 8) {
            float v[8];
            if (nchunks > 0) {
                tmem_ld8(tmem_base + ((uint32_t)(lq * 32) << 16) + (uint32_t)c0, v);
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = 0.f;
            }
            if (row < a.M) {
                float4* dst = reinterpret_cast<float4*>(crow + c0);
                float4 lo = make_float4(v[0], v[1], v[2], v[3]), hi = make_float4(v[4], v[5], v[6], v[7]);
                if (!a.part && a.accumulate) {
                    const float4 o0 = dst[0], o1 = dst[1];
                    lo.x += o0.x; lo.y += o0.y; lo.z += o0.z; lo.w += o0.w;
                    hi.x += o1.x; hi.y += o1.y; hi.z += o1.z; hi.w += o1.w;
                }
                dst[0] = lo;
                dst[1] = hi;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, GN);
}

// Operand images of a weight matrix W for use as the B operand, one GN x GK image per (N tile, K chunk), hi and lo:
//   variant 0 (W is [N x K], C = A . W^T): image rows = n within the tile, cols = k within the chunk (K-major)
//   variant 1 (W is [K x N], C = A . W):   image rows = k within the chunk, cols = n within the tile (read MN-major)
__global__ void __launch_bounds__(256)
wide_pack_b_kernel(const float* __restrict__ W, int variant, int N, int64_t K, uint8_t* __restrict__ out_hi, uint8_t* __restrict__ out_lo) {
    const int ntile = (N + GN - 1) / GN;
    const int64_t nchunk = K / GK;
    const int64_t total = (int64_t)N * K;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t n, k;
        if (variant != 1) { n = i / K; k = i - n * K; } else { k = i / N; n = i - k * N; }
        int nt = (int)(n / GN), nl = (int)(n - (int64_t)nt * GN);
        if (variant == 3) {                   // W row n = gate * H + u  ->  tile u / 64, row gate * 64 + u % 64   (N = 4H)
            const int Hh = N / 4, gate = (int)(n / Hh), u = (int)(n - (int64_t)gate * Hh);
            nt = u / 64;
            nl = gate * 64 + (u & 63);
        }
        const int64_t kc = k / GK;
        const int kl = (int)(k - kc * GK);
        const int NT = N - nt * GN < GN ? N - nt * GN : GN;
        const int64_t img = variant != 1 ? (int64_t)nt * nchunk + kc : kc * ntile + nt;
        const int r = variant != 1 ? nl : kl, c = variant != 1 ? kl : nl, cols = variant != 1 ? GK : NT;
        const uint32_t o = (uint32_t)(r >> 3) * (uint32_t)(cols / 8) * LBO_BYTES + (uint32_t)(c >> 3) * LBO_BYTES + (uint32_t)(r & 7) * 16 + (uint32_t)(c & 7) * 2;
        const float v = W[i];
        const __nv_bfloat16 h = __float2bfloat16_rn(v);
        *reinterpret_cast<__nv_bfloat16*>(out_hi + img * (GN * GK * 2) + o) = h;
        *reinterpret_cast<__nv_bfloat16*>(out_lo + img * (GN * GK * 2) + o) = __float2bfloat16_rn(v - __bfloat162float(h));
    }
}

// C[i] = sum over splits of part[z][i] in fixed order (bit-identical run to run)
__global__ void __launch_bounds__(256) wide_splitk_reduce_kernel(const float* __restrict__ part, int splits, int64_t n, float* __restrict__ C) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float s = 0.f;
        for (int z = 0; z < splits; ++z) s += part[(int64_t)z * n + i];
        C[i] = s;
    }
}

template <int NSPLIT, int VAR, bool FUSE = false>
int launch_gemm_t(GemmArgs a, int splits, cudaStream_t st, const char* name) {
    using S = GemmSmem<NSPLIT>;
    auto kern = wide_gemm_kernel<NSPLIT, VAR, FUSE>;
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::total));
    dim3 grid((unsigned)(((a.M + GM - 1) / GM) * ((a.N + GN - 1) / GN)), 1, (unsigned)splits);
    {
        TimedLaunch _t(name, st);
        kern<<<grid, G_THREADS, S::total, st>>>(a);
    }
    B200_LAUNCH_CHECK("wide_gemm_kernel");
    return B200_OK;
}

}  // namespace

// C (+)= op(A) . op(B); see the variant table above.  Split-K (variant 2 only) needs `part`.
size_t wide_pack_bytes(int N, int64_t K) { return (size_t)((N + GN - 1) / GN) * (size_t)(K / GK) * GN * GK * 2; }   // one of hi / lo

int wide_pack_b(const float* W, int variant, int N, int64_t K, uint8_t* out_hi, uint8_t* out_lo, cudaStream_t st) {
    B200_CHECK_ARG(variant == 0 || variant == 1 || variant == 3, "wide_pack_b: variant must be 0, 1 or 3 (gate-interleaved 0)");
    B200_CHECK_ARG(variant != 3 || N % 256 == 0, "wide_pack_b: gate-interleaved packing needs N = 4H with H a multiple of 64");
    B200_CHECK_ARG(W && out_hi && out_lo && N % 64 == 0 && K % GK == 0, "wide_pack_b: bad arguments");
    int64_t grid = ((int64_t)N * K + 255) / 256;
    if (grid > (int64_t)num_sms() * 8) grid = (int64_t)num_sms() * 8;
    { TimedLaunch _t("wide_pack_b_kernel", st); wide_pack_b_kernel<<<(unsigned)grid, 256, 0, st>>>(W, variant, N, K, out_hi, out_lo); }
    B200_LAUNCH_CHECK("wide_pack_b_kernel");
    return B200_OK;
}

int wide_gemm(int nsplit, int variant, const float* A, const float* B, float* C, int64_t M, int N, int64_t K, int64_t lda,
              int64_t ldb, int64_t ldc, int accumulate, float* part, size_t part_floats, cudaStream_t st, const char* name,
              const uint8_t* bp_hi = nullptr, const uint8_t* bp_lo = nullptr) {
    B200_CHECK_ARG(nsplit == 1 || nsplit == 2, "wide_gemm: nsplit must be 1 or 2");
    B200_CHECK_ARG(variant >= 0 && variant <= 2, "wide_gemm: variant must be 0, 1 or 2");
    B200_CHECK_ARG(A && B && C && M > 0 && N > 0 && K > 0, "wide_gemm: bad arguments (M=%lld N=%d K=%lld)", (long long)M, N, (long long)K);
    B200_CHECK_ARG(N % 64 == 0, "wide_gemm: N must be a multiple of 64, got %d", N);
    B200_CHECK_ARG(variant == 2 || K % GK == 0, "wide_gemm: K must be a multiple of 64 for variants 0 and 1, got %lld", (long long)K);
    B200_CHECK_ARG(variant != 2 || M % GM == 0, "wide_gemm: M must be a multiple of 128 for variant 2, got %lld", (long long)M);
    B200_CHECK_ARG(lda % 4 == 0 && ldb % 4 == 0 && ldc % 4 == 0, "wide_gemm: leading dimensions must be multiples of 4 floats");
    B200_CHECK_ARG((((uintptr_t)A | (uintptr_t)B | (uintptr_t)C) & 15) == 0, "wide_gemm: pointers must be 16-byte aligned");
    GemmArgs a{A, B, C, M, N, K, lda, ldb, ldc, accumulate, 0, nullptr, variant == 2 ? nullptr : bp_hi, variant == 2 ? nullptr : bp_lo,
               0, 0, 0, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    int splits = 1;
    const int64_t chunks = (K + GK - 1) / GK;
    if (variant == 2 && part) {
        const int64_t tiles = ((M + GM - 1) / GM) * ((N + GN - 1) / GN);
        int64_t want = (2 * (int64_t)num_sms() + tiles - 1) / tiles;          // about two waves of CTAs
        if (want > MAX_SPLITS) want = MAX_SPLITS;
        if (want > chunks) want = chunks;
        const int64_t fit = (int64_t)(part_floats / ((size_t)M * N));
        if (want > fit) want = fit;
        if (want > 1) splits = (int)want;
    }
    a.k_per_split = ((chunks + splits - 1) / splits) * GK;
    splits = (int)((K + a.k_per_split - 1) / a.k_per_split);
    if (splits > 1) a.part = part;
    B200_CHECK_ARG(splits == 1 || ldc == N, "wide_gemm: split-K needs a dense C");
    int rc;
    if (nsplit == 2) rc = variant == 0 ? launch_gemm_t<2, 0>(a, splits, st, name) : variant == 1 ? launch_gemm_t<2, 1>(a, splits, st, name) : launch_gemm_t<2, 2>(a, splits, st, name);
    else rc = variant == 0 ? launch_gemm_t<1, 0>(a, splits, st, name) : variant == 1 ? launch_gemm_t<1, 1>(a, splits, st, name) : launch_gemm_t<1, 2>(a, splits, st, name);
    if (rc) return rc;
    if (splits > 1) {
        const int64_t n = M * N;
        int64_t grid = (n + 255) / 256;
        if (grid > (int64_t)num_sms() * 8) grid = (int64_t)num_sms() * 8;
        { TimedLaunch _t("wide_splitk_reduce_kernel", st); wide_splitk_reduce_kernel<<<(unsigned)grid, 256, 0, st>>>(part, splits, n, C); }
        B200_LAUNCH_CHECK("wide_splitk_reduce_kernel");
    }
    return B200_OK;
}

// Forward recurrent step with the LSTM cell fused into the GEMM epilogue: G_t (input projection, torch gate order) += h_{t-1} . W_hh^T,
// bias, activations, c_t, h_t; G_t receives the activated gates.  W_hh comes as gate-interleaved operand images (pack variant 3).
int wide_rec_fwd_fused(int nsplit, const float* h_prev, float* Gt, int64_t N, int H, int has_x, const float* b_ih, const float* b_hh,
                       const float* c_prev, float* c_out, float* h_out, const float* x_t, const float* w_ih0, int F,
                       const uint8_t* bp_hi, const uint8_t* bp_lo, cudaStream_t st) {
    B200_CHECK_ARG(H % 64 == 0 && bp_hi && bp_lo, "wide_rec_fwd_fused: needs H % 64 == 0 and packed weights");
    GemmArgs a{h_prev, nullptr, Gt, N, 4 * H, H, H, H, 4LL * H, 0, (int64_t)H, nullptr, bp_hi, bp_lo,
               H, has_x, F, b_ih, b_hh, c_prev, c_out, h_out, x_t, w_ih0};
    return nsplit == 2 ? launch_gemm_t<2, 0, true>(a, 1, st, "wide_gemm_rec_fwd_cell") : launch_gemm_t<1, 0, true>(a, 1, st, "wide_gemm_rec_fwd_cell");
}

namespace {

// ------------------------------------------------------------------------------------------- pointwise kernels
// forward cell of step t of one layer: G_t [N][4H] holds the GEMM part of the gate pre-activations (nothing for
// layer 0 at t = 0); adds bias (and x_t . W_ih^T for layer 0), writes activated gates, c_t, h_t
__global__ void __launch_bounds__(256)
wide_cell_fwd_kernel(float* __restrict__ G, const float* __restrict__ b_ih, const float* __restrict__ b_hh,
                     const float* __restrict__ x_t, const float* __restrict__ w_ih0, int F, int has_gemm,
                     const float* __restrict__ c_prev, float* __restrict__ c_out, float* __restrict__ h_out, int64_t N, int H) {
    const int64_t total = N * H;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t n = i / H;
        const int u = (int)(i - n * H);
        float* g = G + n * 4 * H;
        float pre[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int j = k * H + u;
            float v = (has_gemm ? g[j] : 0.f) + (b_ih[j] + b_hh[j]);
            if (w_ih0) for (int f = 0; f < F; ++f) v = fmaf(x_t[n * F + f], w_ih0[j * F + f], v);
            pre[k] = v;
        }
        const float ig = sigmoidf_(pre[0]), fg = sigmoidf_(pre[1]), gg = tanhf(pre[2]), og = sigmoidf_(pre[3]);
        const float c = fg * (c_prev ? c_prev[i] : 0.f) + ig * gg;
        g[u] = ig; g[H + u] = fg; g[2 * H + u] = gg; g[3 * H + u] = og;
        c_out[i] = c;
        h_out[i] = og * tanhf(c);
    }
}

// backward cell of step t: activated gates in G_t are overwritten with the pre-activation gradients
// bsum (nullable): per-thread-group partial column sums of dG = the bias gradient.  The grid stride is a multiple of H (256 threads
// per block, H in {128, 256}), so a thread sees ONE hidden unit u for all its rows: it adds its four gate-gradient sums of this
// launch to its own slot bsum[group][gate * H + u], group = global thread index / H -- read-modify-write of a slot no other thread
// touches, the same grid for every step of a layer: deterministic, and the 100 GB-per-layer column-sum pass over dG disappears.
__global__ void __launch_bounds__(256)
wide_cell_bwd_kernel(float* __restrict__ G, const float* __restrict__ c_t, const float* __restrict__ c_prev,
                     const float* __restrict__ d_out_t, const float* __restrict__ dhT, const float* __restrict__ dh_rec,
                     float* __restrict__ dc, int first, int64_t N, int H, float* __restrict__ bsum) {
    const int64_t total = N * H;
    float bs0 = 0.f, bs1 = 0.f, bs2 = 0.f, bs3 = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t n = i / H;
        const int u = (int)(i - n * H);
        float* g = G + n * 4 * H;
        const float ig = g[u], fg = g[H + u], gg = g[2 * H + u], og = g[3 * H + u];
        float dh = 0.f;
        if (d_out_t) dh += d_out_t[i];
        if (dhT) dh += dhT[i];
        if (dh_rec) dh += dh_rec[i];
        const float tc = tanhf(c_t[i]);
        const float dcc = (first ? 0.f : dc[i]) + dh * og * (1.f - tc * tc);
        const float cp = c_prev ? c_prev[i] : 0.f;
        dc[i] = dcc * fg;
        const float d0 = dcc * gg * ig * (1.f - ig), d1 = dcc * cp * fg * (1.f - fg), d2 = dcc * ig * (1.f - gg * gg), d3 = dh * tc * og * (1.f - og);
        g[u] = d0;
        g[H + u] = d1;
        g[2 * H + u] = d2;
        g[3 * H + u] = d3;
        bs0 += d0; bs1 += d1; bs2 += d2; bs3 += d3;
    }
    if (bsum != nullptr) {
        const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
        if (tid < total) {
            float* o = bsum + (tid / H) * 4 * H + (tid % H);
            o[0] += bs0; o[H] += bs1; o[2 * H] += bs2; o[3 * H] += bs3;
        }
    }
}

// dst = src * keep / (1 - p) with the keep mask of the LSTM kernels (common.cuh: one Philox block per 8 units)
__global__ void __launch_bounds__(256)
wide_dropout_kernel(uint64_t seed, int layer, int64_t N, int T, int H, float p, const float* src, float* dst) {   // src may alias dst
    const uint32_t thr = dropout_threshold(p);
    const float scale = 1.0f / (1.0f - p);
    const int HB = H / 8;
    const int64_t total = (int64_t)T * N * HB;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int ub = (int)(i % HB);
        const int64_t rt = i / HB;
        const int64_t row = rt % N;
        const int t = (int)(rt / N);
        const U4 b = dropout_block(seed, layer, row, t, ub);
        const float4* s4 = reinterpret_cast<const float4*>(src + rt * H + ub * 8);
        float4 a0 = s4[0], a1 = s4[1];
        a0.x = dropout_lane16(b, 0) >= thr ? a0.x * scale : 0.f; a0.y = dropout_lane16(b, 1) >= thr ? a0.y * scale : 0.f;
        a0.z = dropout_lane16(b, 2) >= thr ? a0.z * scale : 0.f; a0.w = dropout_lane16(b, 3) >= thr ? a0.w * scale : 0.f;
        a1.x = dropout_lane16(b, 4) >= thr ? a1.x * scale : 0.f; a1.y = dropout_lane16(b, 5) >= thr ? a1.y * scale : 0.f;
        a1.z = dropout_lane16(b, 6) >= thr ? a1.z * scale : 0.f; a1.w = dropout_lane16(b, 7) >= thr ? a1.w * scale : 0.f;
        float4* d4 = reinterpret_cast<float4*>(dst + rt * H + ub * 8);
        d4[0] = a0;
        d4[1] = a1;
    }
}

// column sums of dG [R][C] (bias gradient) and, for layer 0, dG^T . x (x [R][F], F <= 8): per-CTA partials over a slab
// of rows, then a fixed-order reduce.  part [nslab][C][F + 1]: column F = the plain sum
constexpr int CS_MAXF = 8;
__global__ void __launch_bounds__(256)
wide_colsum_kernel(const float* __restrict__ dG, const float* __restrict__ x, int F, int64_t R, int C, int64_t rows_per_slab,
                   float* __restrict__ part) {
    const int64_t r0 = (int64_t)blockIdx.y * rows_per_slab;
    const int64_t r1 = r0 + rows_per_slab < R ? r0 + rows_per_slab : R;
    for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < C; j += gridDim.x * blockDim.x) {
        float acc[CS_MAXF + 1];
#pragma unroll
        for (int f = 0; f <= CS_MAXF; ++f) acc[f] = 0.f;
        for (int64_t r = r0; r < r1; ++r) {
            const float g = dG[r * C + j];
            acc[CS_MAXF] += g;
#pragma unroll
            for (int f = 0; f < CS_MAXF; ++f) if (f < F) acc[f] = fmaf(g, x[r * F + f], acc[f]);
        }
        float* o = part + ((int64_t)blockIdx.y * C + j) * (F + 1);
        for (int f = 0; f < F; ++f) o[f] = acc[f];
        o[F] = acc[CS_MAXF];
    }
}
__global__ void __launch_bounds__(256)
wide_colsum_reduce_kernel(const float* __restrict__ part, int nslab, int C, int F, float* __restrict__ gw_ih0,
                          float* __restrict__ gb_ih, float* __restrict__ gb_hh) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < C * (F + 1); i += gridDim.x * blockDim.x) {
        const int j = i / (F + 1), f = i - j * (F + 1);
        float s = 0.f;
        for (int z = 0; z < nslab; ++z) s += part[((int64_t)z * C + j) * (F + 1) + f];
        if (f < F) { if (gw_ih0) gw_ih0[j * F + f] = s; }
        else { gb_ih[j] = s; gb_hh[j] = s; }
    }
}

inline unsigned ew_grid(int64_t n) {
    int64_t g = (n + 255) / 256;
    const int64_t cap = (int64_t)num_sms() * 16;
    return (unsigned)(g < cap ? (g > 0 ? g : 1) : cap);
}

struct WideWs {
    float* G;        // [L][T][N][4H] gates (forward: activated; backward: overwritten with dG layer by layer)
    float* xd;       // [T][N][H]     dropped input of the layer being processed
    float* dx[2];    // [T][N][H]     d_out ping-pong between layers
    float* dh_rec;   // [N][H]
    float* dc;       // [N][H]
    float* part;     // split-K partials / column-sum partials
    size_t part_floats;
    float* bsum;     // [groups][4H] bias-gradient partials accumulated by the backward cell kernel (layers > 0)
    uint8_t* pack[4];   // B-operand images of the current layer's weights: [W_ih, W_hh] x [hi, lo] (K-major in forward, MN-major in backward)
};
constexpr int COLSUM_SLABS = 64;

size_t wide_part_floats(int H, int F) {
    const size_t gemm = (size_t)MAX_SPLITS * 4 * H * H;
    const size_t cs = (size_t)COLSUM_SLABS * 4 * H * (CS_MAXF + 1);
    (void)F;
    return gemm > cs ? gemm : cs;
}
size_t al256(size_t b) { return (b + 255) & ~(size_t)255; }
// the backward cell kernel runs on at most num_sms * 16 blocks of 256 threads = groups of H threads with 4H floats each
size_t wide_bsum_floats() { return (size_t)num_sms() * 16 * 256 * 4; }
// one of hi / lo of a [4H x H] weight as B operand: [N = 4H][K = H] (variant 0) or [K = 4H][N = H] (variant 1); images are padded
// to GN columns, so the two counts differ when H < GN
size_t wide_pack_region(int H) {
    const size_t a = wide_pack_bytes(4 * H, H), b = wide_pack_bytes(H, 4 * H);
    return al256(a > b ? a : b);
}

bool wide_ws(void* workspace, size_t ws_bytes, int64_t N, int T, int F, int H, int L, WideWs* w) {
    uint8_t* p = (uint8_t*)workspace;
    const size_t lsz = (size_t)T * N * H * 4;
    size_t off = 0;
    w->G = (float*)(p + off); off += al256((size_t)L * lsz * 4);
    w->xd = (float*)(p + off); off += al256(lsz);
    w->dx[0] = (float*)(p + off); off += al256(lsz);
    w->dx[1] = (float*)(p + off); off += al256(lsz);
    w->dh_rec = (float*)(p + off); off += al256((size_t)N * H * 4);
    w->dc = (float*)(p + off); off += al256((size_t)N * H * 4);
    w->part = (float*)(p + off);
    w->part_floats = wide_part_floats(H, F);
    off += al256(w->part_floats * 4);
    w->bsum = (float*)(p + off);
    off += al256(wide_bsum_floats() * 4);
    const size_t pk = wide_pack_region(H);
    for (int i = 0; i < 4; ++i) { w->pack[i] = p + off; off += pk; }
    return workspace != nullptr && off <= ws_bytes;
}

struct LayerParams { const float *w_ih, *w_hh, *b_ih, *b_hh; int in; };
LayerParams layer_params(const float* params, int F, int H, int l) {
    int64_t off = 0;
    for (int i = 0; i < l; ++i) off += 4LL * H * (i == 0 ? F : H) + 4LL * H * H + 8LL * H;
    const int in = l == 0 ? F : H;
    LayerParams lp;
    lp.in = in;
    lp.w_ih = params + off;
    lp.w_hh = lp.w_ih + 4LL * H * in;
    lp.b_ih = lp.w_hh + 4LL * H * H;
    lp.b_hh = lp.b_ih + 4LL * H;
    return lp;
}

}  // namespace

// bf16 mode (B200_MODE_TC_FAST) at hidden 128 / 256: the operand-image engine of lstm_wide2.cu
bool wide2_supported(int H, int F);
size_t wide2_workspace_bytes(int64_t N, int T, int F, int H, int L);
int wide2_fwd(const float* params, int F, int H, int L, const float* x, int64_t N, int T, float* stash_h, float* stash_c, float p,
              uint64_t seed, int training, void* workspace, size_t ws_bytes, cudaStream_t st);
int wide2_bwd_layer(int layer, const float* params, int F, int H, int L, const float* x, int64_t N, int T, const float* stash_h,
                    const float* stash_c, const float* dhT, float* grads, float p, uint64_t seed, int training, void* workspace,
                    size_t ws_bytes, cudaStream_t st);
static int g_wide_images = 1;         // option "wide_images": 0 = bf16 mode on the fp32-storage path below (comparison runs)
void wide_set_images(int on) { g_wide_images = on; }
bool wide_uses_images(int mode, int H, int F) { return mode == B200_MODE_TC_FAST && g_wide_images != 0 && wide2_supported(H, F); }

size_t wide_workspace_bytes(int mode, int64_t N, int T, int F, int H, int L) {
    if (wide_uses_images(mode, H, F)) return wide2_workspace_bytes(N, T, F, H, L);
    const size_t lsz = (size_t)T * N * H * 4;
    return al256((size_t)L * lsz * 4) + 3 * al256(lsz) + 2 * al256((size_t)N * H * 4) + al256(wide_part_floats(H, F) * 4) +
           al256(wide_bsum_floats() * 4) + 4 * wide_pack_region(H) + 256;
}

static int g_wide_fuse = 0;
void wide_set_fuse(int on) { g_wide_fuse = on; }

static int wide_check(const char* who, int H, int F) {
    B200_CHECK_ARG(H % 64 == 0 && H >= 128, "%s: the wide engine needs a hidden size that is a multiple of 64 and >= 128, got %d", who, H);
    B200_CHECK_ARG(F <= CS_MAXF, "%s: at most %d input features", who, CS_MAXF);
    return B200_OK;
}

int wide_fwd(int mode, const float* params, int F, int H, int L, const float* x, int64_t N, int T, float* stash_h,
             float* stash_c, float p, uint64_t seed, int training, void* workspace, size_t ws_bytes, cudaStream_t st) {
    int rc = wide_check("lstm_fwd", H, F);
    if (rc) return rc;
    if (wide_uses_images(mode, H, F)) return wide2_fwd(params, F, H, L, x, N, T, stash_h, stash_c, p, seed, training, workspace, ws_bytes, st);
    WideWs w;
    B200_CHECK_ARG(wide_ws(workspace, ws_bytes, N, T, F, H, L, &w), "lstm_fwd (wide): workspace too small (%zu bytes, need %zu)", ws_bytes,
                   wide_workspace_bytes(mode, N, T, F, H, L));
    const int nsplit = mode == B200_MODE_TC_PRECISE ? 2 : 1;
    const bool drop = training && p > 0.f;
    const int64_t lsz = (int64_t)T * N * H, NH = N * H, G4 = 4LL * H;
    // option wide_fuse = 1: LSTM cell in the epilogue of the recurrent GEMM.  Measured (hidden 256 x 8 layers, 256 windows): no gain in
    // bf16 mode (85 ms against 47 + 38 ms for GEMM + cell kernel) and slower with three products (125 against 102 ms): the
    // thread-per-row epilogue runs after the MMAs of its CTA instead of beside another kernel's, so it stays off by default.
    const bool fuse = g_wide_fuse != 0;      // b200_set_option("wide_fuse", 1)
    for (int l = 0; l < L; ++l) {
        const LayerParams lp = layer_params(params, F, H, l);
        float* G = w.G + (int64_t)l * lsz * 4;
        float* hs = stash_h + (int64_t)l * lsz;
        float* cs = stash_c + (int64_t)l * lsz;
        // this layer's weights as K-major operand images (once per step; every GEMM below then stages B with bulk copies)
        if (l > 0) { rc = wide_pack_b(lp.w_ih, 0, (int)G4, H, w.pack[0], w.pack[1], st); if (rc) return rc; }
        rc = wide_pack_b(lp.w_hh, fuse ? 3 : 0, (int)G4, H, w.pack[2], w.pack[3], st);
        if (rc) return rc;
        if (l > 0) {
            const float* xin = stash_h + (int64_t)(l - 1) * lsz;
            if (drop) {
                { TimedLaunch _t("wide_dropout_kernel", st); wide_dropout_kernel<<<ew_grid((int64_t)T * N * (H / 8)), 256, 0, st>>>(seed, l, N, T, H, p, xin, w.xd); }
                B200_LAUNCH_CHECK("wide_dropout_kernel");
                xin = w.xd;
            }
            rc = wide_gemm(nsplit, 0, xin, lp.w_ih, G, (int64_t)T * N, (int)G4, H, H, H, G4, 0, nullptr, 0, st, "wide_gemm_xproj", w.pack[0], w.pack[1]);
            if (rc) return rc;
        }
        for (int t = 0; t < T; ++t) {
            float* Gt = G + (int64_t)t * N * G4;
            if (t > 0 && fuse) {
                rc = wide_rec_fwd_fused(nsplit, hs + (int64_t)(t - 1) * NH, Gt, N, H, l > 0 ? 1 : 0, lp.b_ih, lp.b_hh, cs + (int64_t)(t - 1) * NH,
                                        cs + (int64_t)t * NH, hs + (int64_t)t * NH, l == 0 ? x + (int64_t)t * N * F : nullptr,
                                        l == 0 ? lp.w_ih : nullptr, F, w.pack[2], w.pack[3], st);
                if (rc) return rc;
                continue;
            }
            if (t > 0) {
                rc = wide_gemm(nsplit, 0, hs + (int64_t)(t - 1) * NH, lp.w_hh, Gt, N, (int)G4, H, H, H, G4, l > 0 ? 1 : 0, nullptr, 0, st,
                               "wide_gemm_rec_fwd", w.pack[2], w.pack[3]);
                if (rc) return rc;
            }
            {
                TimedLaunch _t("wide_cell_fwd_kernel", st);
                wide_cell_fwd_kernel<<<ew_grid(NH), 256, 0, st>>>(Gt, lp.b_ih, lp.b_hh, l == 0 ? x + (int64_t)t * N * F : nullptr,
                                                                  l == 0 ? lp.w_ih : nullptr, F, (l > 0 || t > 0) ? 1 : 0,
                                                                  t > 0 ? cs + (int64_t)(t - 1) * NH : nullptr, cs + (int64_t)t * NH,
                                                                  hs + (int64_t)t * NH, N, H);
            }
            B200_LAUNCH_CHECK("wide_cell_fwd_kernel");
        }
    }
    return B200_OK;
}

// Backward of one layer (layers must be called top-down; the forward's gate stash in the workspace is consumed).
int wide_bwd_layer(int mode, int layer, const float* params, int F, int H, int L, const float* x, int64_t N, int T,
                   const float* stash_h, const float* stash_c, const float* dhT, float* grads, float p, uint64_t seed,
                   int training, void* workspace, size_t ws_bytes, cudaStream_t st) {
    int rc = wide_check("lstm_bwd", H, F);
    if (rc) return rc;
    if (wide_uses_images(mode, H, F))
        return wide2_bwd_layer(layer, params, F, H, L, x, N, T, stash_h, stash_c, dhT, grads, p, seed, training, workspace, ws_bytes, st);
    WideWs w;
    B200_CHECK_ARG(wide_ws(workspace, ws_bytes, N, T, F, H, L, &w), "lstm_bwd (wide): workspace too small (%zu bytes)", ws_bytes);
    const int nsplit = mode == B200_MODE_TC_PRECISE ? 2 : 1;
    const bool drop = training && p > 0.f;
    const int64_t lsz = (int64_t)T * N * H, NH = N * H, G4 = 4LL * H;
    const int l = layer;
    const LayerParams lp = layer_params(params, F, H, l);
    const int64_t poff = lp.w_ih - params;
    float* gw_ih = grads + poff;
    float* gw_hh = gw_ih + 4LL * H * lp.in;
    float* gb_ih = gw_hh + 4LL * H * H;
    float* gb_hh = gb_ih + 4LL * H;
    float* G = w.G + (int64_t)l * lsz * 4;
    const float* hs = stash_h + (int64_t)l * lsz;
    const float* cs = stash_c + (int64_t)l * lsz;
    const bool top = l == L - 1;
    const float* d_out = top ? nullptr : w.dx[((L - 1 - l) & 1) ^ 1];        // written by layer l + 1
    float* d_in = w.dx[(L - 1 - l) & 1];

    // W_hh [4H x H] and W_ih [4H x H] (layers > 0) as MN-major operand images for dG . W
    rc = wide_pack_b(lp.w_hh, 1, H, G4, w.pack[2], w.pack[3], st);
    if (rc) return rc;
    if (l > 0) { rc = wide_pack_b(lp.w_ih, 1, H, G4, w.pack[0], w.pack[1], st); if (rc) return rc; }
    // layers > 0: the bias gradient is accumulated by the cell kernel itself (layer 0 keeps the column-sum pass, which also yields dW_ih)
    const unsigned cell_grid = ew_grid(NH);
    const int64_t bgroups = ((int64_t)cell_grid * 256 < NH ? (int64_t)cell_grid * 256 : NH + H - 1) / H;
    float* bsum = l > 0 ? w.bsum : nullptr;
    if (bsum) B200_CUDA(cudaMemsetAsync(bsum, 0, sizeof(float) * (size_t)bgroups * G4, st));
    for (int t = T - 1; t >= 0; --t) {
        float* Gt = G + (int64_t)t * N * G4;
        {
            TimedLaunch _t("wide_cell_bwd_kernel", st);
            wide_cell_bwd_kernel<<<cell_grid, 256, 0, st>>>(Gt, cs + (int64_t)t * NH, t > 0 ? cs + (int64_t)(t - 1) * NH : nullptr,
                                                              d_out ? d_out + (int64_t)t * NH : nullptr, (top && t == T - 1) ? dhT : nullptr,
                                                              t < T - 1 ? w.dh_rec : nullptr, w.dc, t == T - 1 ? 1 : 0, N, H, bsum);
        }
        B200_LAUNCH_CHECK("wide_cell_bwd_kernel");
        if (t > 0) {
            rc = wide_gemm(nsplit, 1, Gt, lp.w_hh, w.dh_rec, N, H, G4, G4, H, H, 0, nullptr, 0, st, "wide_gemm_rec_bwd", w.pack[2], w.pack[3]);
            if (rc) return rc;
        }
    }
    // weight gradients: dW_hh = dG[1..T-1]^T . h[0..T-2]
    if (T > 1) {
        rc = wide_gemm(nsplit, 2, G + N * G4, hs, gw_hh, G4, H, (int64_t)(T - 1) * N, G4, H, H, 0, w.part, w.part_floats, st, "wide_gemm_wgrad_hh");
        if (rc) return rc;
    } else {
        B200_CUDA(cudaMemsetAsync(gw_hh, 0, sizeof(float) * 4 * H * H, st));
    }
    const int64_t R = (int64_t)T * N;
    if (l > 0) {
        const float* xin = stash_h + (int64_t)(l - 1) * lsz;
        if (drop) {
            { TimedLaunch _t("wide_dropout_kernel", st); wide_dropout_kernel<<<ew_grid(R * (H / 8)), 256, 0, st>>>(seed, l, N, T, H, p, xin, w.xd); }
            B200_LAUNCH_CHECK("wide_dropout_kernel");
            xin = w.xd;
        }
        rc = wide_gemm(nsplit, 2, G, xin, gw_ih, G4, H, R, G4, H, H, 0, w.part, w.part_floats, st, "wide_gemm_wgrad_ih");
        if (rc) return rc;
        // d_out of the layer below: dX = dG . W_ih, then the mask that was applied to this layer's input
        rc = wide_gemm(nsplit, 1, G, lp.w_ih, d_in, R, H, G4, G4, H, H, 0, nullptr, 0, st, "wide_gemm_dgrad_x", w.pack[0], w.pack[1]);
        if (rc) return rc;
        if (drop) {
            { TimedLaunch _t("wide_dropout_kernel", st); wide_dropout_kernel<<<ew_grid(R * (H / 8)), 256, 0, st>>>(seed, l, N, T, H, p, d_in, d_in); }
            B200_LAUNCH_CHECK("wide_dropout_kernel");
        }
    }
    // bias gradients (and layer 0's dW_ih, K = F on CUDA cores)
    if (bsum) {
        { TimedLaunch _t("wide_colsum_reduce_kernel", st); wide_colsum_reduce_kernel<<<(unsigned)((G4 + 255) / 256), 256, 0, st>>>(bsum, (int)bgroups, (int)G4, 0, nullptr, gb_ih, gb_hh); }
        B200_LAUNCH_CHECK("wide_colsum_reduce_kernel");
    } else {
        const int Fx = l == 0 ? F : 0;
        int64_t rows_per = (R + COLSUM_SLABS - 1) / COLSUM_SLABS;
        const int nslab = (int)((R + rows_per - 1) / rows_per);
        dim3 grid((unsigned)((G4 + 255) / 256), (unsigned)nslab);
        { TimedLaunch _t("wide_colsum_kernel", st); wide_colsum_kernel<<<grid, 256, 0, st>>>(G, l == 0 ? x : nullptr, Fx, R, (int)G4, rows_per, w.part); }
        B200_LAUNCH_CHECK("wide_colsum_kernel");
        { TimedLaunch _t("wide_colsum_reduce_kernel", st); wide_colsum_reduce_kernel<<<(unsigned)((G4 * (Fx + 1) + 255) / 256), 256, 0, st>>>(w.part, nslab, (int)G4, Fx, l == 0 ? gw_ih : nullptr, gb_ih, gb_hh); }
        B200_LAUNCH_CHECK("wide_colsum_reduce_kernel");
    }
    return B200_OK;
}

}  // namespace b200

using namespace b200;

// Diagnostics: the engine's GEMM on caller-supplied matrices (fp32, row-major, device).  variant 0: C = A[MxK] . B[NxK]^T;
// 1: C = A[MxK] . B[KxN]; 2: C = A[KxM]^T . B[KxN] (split-K when workspace is given).  nsplit 2 = bf16 hi/lo, 1 = plain bf16.
extern "C" int b200_wide_gemm(int nsplit, int variant, const float* A, const float* B, float* C, int64_t M, int N, int64_t K,
                              int accumulate, void* workspace, size_t workspace_bytes, void* stream) {
    const int64_t lda = variant == 2 ? M : K, ldb = variant == 0 ? K : N;
    if (variant != 2 && workspace && N > 0 && K > 0 && N % 64 == 0 && K % 64 == 0) {      // B through pre-packed images + TMA bulk copies
        const size_t pk = (wide_pack_bytes(N, K) + 255) & ~(size_t)255;
        if (workspace_bytes >= 2 * pk) {
            uint8_t* hi = (uint8_t*)workspace;
            uint8_t* lo = hi + pk;
            int rc = wide_pack_b(B, variant, N, K, hi, lo, (cudaStream_t)stream);
            if (rc) return rc;
            return wide_gemm(nsplit, variant, A, B, C, M, N, K, lda, ldb, N, accumulate, nullptr, 0, (cudaStream_t)stream, "wide_gemm_kernel", hi, lo);
        }
    }
    return wide_gemm(nsplit, variant, A, B, C, M, N, K, lda, ldb, N, accumulate, (float*)workspace, workspace_bytes / sizeof(float),
                     (cudaStream_t)stream, "wide_gemm_kernel");
}
