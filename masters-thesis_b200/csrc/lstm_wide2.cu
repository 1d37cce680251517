// Wide-model LSTM engine, bf16 mode (B200_MODE_TC_FAST, BASELINE.json configs[2]: hidden 256, 8 layers, bf16 GEMMs with fp32
// accumulation).  Every operand of every GEMM lives in global memory as ready-made UMMA operand images, so the kernels move
// them with bulk copies only (no thread touches an operand on its way to the tensor core), and the LSTM cell runs in the
// epilogue of the forward GEMM.  lstm_wide.cu keeps the fp32-storage path of B200_MODE_TC_PRECISE.
//
// Activation chunk = a [128 rows x 128 columns] bf16 operand image in the interleaved (SWIZZLE_NONE) layout of tc_common.cuh, 32 KB:
//     byte(r, c) = (r / 8) * 2048 + (c / 8) * 128 + (r % 8) * 16 + (c % 8) * 2
// Read K-major it is a [128 x K = 128] operand (two K = 64 halves at +0 / +1024, SBO 2048); read MN-major the same bytes are a
// [K = 128 rows x 128] operand whose 64-row halves are 16 KB of contiguous memory -- so every kernel stages its operands with a
// few large bulk copies.  (A first version with [128 x 64] chunks had to assemble the MN-major stages of the weight gradient from
// 64 copies of 1 KB; ncu showed the loader warp serialising them, 4.0 k clk per stage against 1.0 k clk of MMAs.)
// With NT = ceil(N / 128) row tiles, HC = H / 128, UG = H / 32:
//   Himg[l]  [T][NT][HC] chunks   h_t of layer l (written by the forward epilogue); under dropout a second, masked set is the
//                                 input of layer l + 1 (the 1/(1-p) lives in the packed W_ih, as in lstm_tc.cu)
//   X0img    [T][NT][1]  chunk    the F input features of layer 0, columns >= F zero
//   Gimg[l]  [T][NT][UG] chunks   activated gates (forward) -> pre-activation gradients dG (backward, in place); gate columns
//                                 are permuted: g' = (u / 32) * 128 + gate * 32 + u % 32, so that the four gates of 32 hidden
//                                 units are one chunk (one N tile of the forward GEMM, one M tile of the weight gradient)
//   c, dc (fp32), dh, dx (bf16)    in the "row-slot" layout [tile][u / 8][128 rows][8]: the thread that owns a row (TMEM lane) moves
//                                 32 or 16 bytes and a warp 1 KB / 512 B of contiguous memory (row-major [N][H] costs 32 cache lines per
//                                 warp access)
//   Wf[l]    [UG][KC] images      [128 rows g' x 64] K-major (16 KB each) of [W_ih | W_hh], KC = 2 * ((1 or HC) + HC) K-chunks of 64
//   Wd[l]    per part (h, x): [4H / 64][64 g' rows x H columns] MN-major images for dG . W
//
//   forward  step t: w2_fwd_kernel    gates = [x_t | h_{t-1}] . Wf^T, cell in the epilogue; persistent CTAs, each keeps the weights
//                                     of one unit group (128 KB) in shared memory and streams the A chunks of its row tiles
//   backward step t: w2_cell_bwd_kernel  (dh, dc, gates, c_{t-1}) -> dG, dc_{t-1}, bias-gradient partials; operands staged through shared
//                                          memory with bulk copies (every block's operands are contiguous in the layouts above)
//                    w2_dgrad_kernel      [dh_{t-1} | dx_t] = dG_t . Wd          two 128 x 256 accumulators per CTA
//            layer:  w2_wgrad_kernel      d[W_ih | W_hh] = dG^T . [x | h_{t-1}]  over all steps, split-K, both operands MN-major
//
// Replaces, for hidden != 64 in bf16 mode: nn.LSTM forward / backward under autocast (src/model.py:88-109, train.py:172
// precision), as lstm_tc.cu does for hidden 64.
#include <cuda_bf16.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace b200 {
using namespace tc;

namespace {

constexpr int TM = 128;
constexpr uint32_t CHUNK = 16384, RG_BYTES = 1024;          // weight image [128 x 64]
constexpr uint32_t WCHUNK = 32768, WRG_BYTES = 2048;         // activation chunk [128 x 128]
constexpr int W2_THREADS = 320;                 // 8 epilogue warps, loader warp, MMA warp

__host__ __device__ inline size_t al256(size_t b) { return (b + 255) & ~(size_t)255; }
__host__ __device__ inline uint32_t chunk_off(int r, int c) {
    return (uint32_t)(r >> 3) * 1024u + (uint32_t)(c >> 3) * 128u + (uint32_t)(r & 7) * 16u + (uint32_t)(c & 7) * 2u;
}
__host__ __device__ inline uint32_t wchunk_off(int r, int c) {
    return (uint32_t)(r >> 3) * 2048u + (uint32_t)(c >> 3) * 128u + (uint32_t)(r & 7) * 16u + (uint32_t)(c & 7) * 2u;
}

// fp32 row-slot layout: float index of (row, unit) and of the 8-float slot (tile, octet, r)
__host__ __device__ inline size_t slot_off(int64_t tile, int octet, int r, int H) { return (((size_t)tile * (H >> 3) + octet) * 128 + r) * 8; }

struct Geom {
    int F, H, L, T;
    int64_t N;
    int NT, HC, UG, GC;                                                       // GC = 4H / 64 K-chunks of the data gradient
    __host__ __device__ int xc(int l) const { return l == 0 ? 1 : HC; }       // activation chunks of the layer input
    __host__ __device__ int kc(int l) const { return 2 * (xc(l) + HC); }      // K = 64 weight images per unit group
};
inline Geom make_geom(int64_t N, int T, int F, int H, int L) {
    Geom g;
    g.F = F; g.H = H; g.L = L; g.T = T; g.N = N;
    g.NT = (int)((N + TM - 1) / TM);
    g.HC = H / 128; g.UG = H / 32; g.GC = H / 16;
    return g;
}

// ---- two-CTA cluster helpers (forward kernel, MC variant).  The multicast copies read the activation images of THIS GPU's
// workspace only: TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's
// memory hung the GPU.
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void bulk_copy_g2s_mc(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar, uint16_t mask) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;" ::"r"(smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "h"(mask)
                 : "memory");
}
__device__ __forceinline__ void umma_commit_mc(uint64_t* bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}

// tanh on the XU pipe, one MUFU (relative error 2^-11: an eighth of the bf16 rounding the activations get when they are stored).
// The bf16 mode computes sigma(x) = 0.5 + 0.5 tanh(x / 2) with it: 5 MUFU and ~11 FP32 operations per cell instead of the 7 and ~30 of
// the exp2 / reciprocal form the fp32-accurate engines use.  (Measured: same kernel time and same accuracy against the oracle as the
// exp2 form -- the forward kernel is not bound by its epilogue arithmetic -- kept for the shorter code.)
__device__ __forceinline__ float tanha(float x) { float y; asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

__device__ __forceinline__ uint32_t pack_bf2(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint4 pack_bf8(const float (&v)[8]) {
    return make_uint4(pack_bf2(v[0], v[1]), pack_bf2(v[2], v[3]), pack_bf2(v[4], v[5]), pack_bf2(v[6], v[7]));
}
__device__ __forceinline__ void unpack_bf8(const uint4 q, float (&v)[8]) {
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        v[2 * i] = __uint_as_float(w[i] << 16);
        v[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
    }
}
__device__ __forceinline__ void ld8(const float* p, float (&v)[8]) {
    const float4 a = reinterpret_cast<const float4*>(p)[0], b = reinterpret_cast<const float4*>(p)[1];
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void st8(float* p, const float (&v)[8]) {
    reinterpret_cast<float4*>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
    reinterpret_cast<float4*>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
}
// keep mask of 8 consecutive units (bit j = keep unit j) -> AND masks for the four bf16 pairs
__device__ __forceinline__ uint4 and_mask8(uint4 q, uint32_t kb) {
    uint32_t km[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) km[j] = (((kb >> (2 * j)) & 1u) ? 0x0000FFFFu : 0u) | (((kb >> (2 * j + 1)) & 1u) ? 0xFFFF0000u : 0u);
    return make_uint4(q.x & km[0], q.y & km[1], q.z & km[2], q.w & km[3]);
}

// ------------------------------------------------------------------------------------------------ small kernels
// Packed weights of one layer: Wf chunks, Wd images (h part, then x part), bias' [UG][4][32] (see w2_fwd_kernel).
struct PackLayout {
    size_t wf, wd_h, wd_x, bias, total;
};
inline PackLayout pack_layout(const Geom& g, int l) {
    PackLayout p;
    p.wf = 0;
    p.wd_h = al256((size_t)g.UG * g.kc(l) * CHUNK);
    p.wd_x = p.wd_h + al256((size_t)g.GC * 64 * g.H * 2);
    p.bias = p.wd_x + (l > 0 ? al256((size_t)g.GC * 64 * g.H * 2) : 0);
    p.total = p.bias + al256((size_t)4 * g.H * 4);
    return p;
}

// in_scale: 1/(1-p) of the inter-layer dropout folded into W_ih (layers > 0)
__global__ void __launch_bounds__(256)
w2_pack_kernel(const float* __restrict__ w_ih, const float* __restrict__ w_hh, const float* __restrict__ b_ih, const float* __restrict__ b_hh,
               int in, int H, int XC /* K = 64 images of the x part */, float in_scale, uint8_t* __restrict__ wf, uint8_t* __restrict__ wd_h, uint8_t* __restrict__ wd_x,
               float* __restrict__ biasp) {
    const int KC = XC + H / 64, G4 = 4 * H;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nthr = (int64_t)gridDim.x * blockDim.x;
    const int KW = KC * 64;
    for (int64_t i = tid; i < (int64_t)G4 * KW; i += nthr) {
        const int gp = (int)(i / KW), k = (int)(i - (int64_t)gp * KW);            // gp = permuted gate row
        const int ug = gp >> 7, gate = (gp >> 5) & 3, u = ug * 32 + (gp & 31);
        const int g = gate * H + u;
        float v;
        const bool xpart = k < XC * 64;
        if (xpart) v = k < in ? w_ih[(int64_t)g * in + k] * in_scale : 0.f;
        else v = w_hh[(int64_t)g * H + (k - XC * 64)];
        const __nv_bfloat16 b = __float2bfloat16_rn(v);
        // forward image: chunk (ug, k / 64), row gp % 128, column k % 64
        *reinterpret_cast<__nv_bfloat16*>(wf + ((size_t)ug * KC + (k >> 6)) * CHUNK + chunk_off(gp & 127, k & 63)) = b;
        // dgrad image: K-chunk gp / 64, k-row gp % 64, column n of the part: [8 row groups][ncols / 8 cores][128 B]
        const int kcx = gp >> 6, kr = gp & 63;
        if (xpart) {
            if (wd_x != nullptr && k < H) {
                const int n = k;
                *reinterpret_cast<__nv_bfloat16*>(wd_x + (size_t)kcx * 64 * H * 2 + (size_t)(kr >> 3) * (H / 8) * 128 + (size_t)(n >> 3) * 128 + (kr & 7) * 16 + (n & 7) * 2) = b;
            }
        } else {
            const int n = k - XC * 64;
            *reinterpret_cast<__nv_bfloat16*>(wd_h + (size_t)kcx * 64 * H * 2 + (size_t)(kr >> 3) * (H / 8) * 128 + (size_t)(n >> 3) * 128 + (kr & 7) * 16 + (n & 7) * 2) = b;
        }
    }
    for (int64_t gp = tid; gp < G4; gp += nthr) {
        const int ug = (int)gp >> 7, gate = ((int)gp >> 5) & 3, u = ug * 32 + ((int)gp & 31);
        const float b = b_ih[gate * H + u] + b_hh[gate * H + u];
        biasp[gp] = gate == 2 ? b : 0.5f * b;                           // sigma(a + b) = 0.5 + 0.5 tanh(0.5 a + 0.5 b)
    }
}

// layer-0 input as activation chunks: X0img[t][tile] (columns >= F and rows >= N zero); x is [T][N][F] fp32
__global__ void __launch_bounds__(256)
w2_x0_kernel(const float* __restrict__ x, int64_t N, int T, int F, int NT, uint8_t* __restrict__ out) {
    // one thread per (t, tile, row, 8-column core)
    const int64_t total = (int64_t)T * NT * TM * 16;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int core = (int)(i & 15), r = (int)((i >> 4) & 127);
        const int64_t tt = i >> 11;                                   // t * NT + tile
        const int64_t tile = tt % NT, t = tt / NT, row = tile * TM + r;
        float v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int c = core * 8 + j;
            v[j] = (row < N && c < F) ? x[((int64_t)t * N + row) * F + c] : 0.f;
        }
        *reinterpret_cast<uint4*>(out + (size_t)tt * WCHUNK + wchunk_off(r, core * 8)) = pack_bf8(v);
    }
}

// keep bits of the input of `layer`: bits[t][row][ug] (uint32: bit j = keep unit ug * 32 + j), Philox blocks of common.cuh
__global__ void __launch_bounds__(256)
w2_bits_kernel(uint64_t seed, int layer, int64_t N, int T, int UG, float p, uint32_t* __restrict__ bits) {
    const uint32_t thr = dropout_threshold(p);
    const int64_t total = (int64_t)T * N * UG;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int ug = (int)(i % UG);
        const int64_t tr = i / UG, row = tr % N;
        const int t = (int)(tr / N);
        uint32_t w = 0;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const U4 blk = dropout_block(seed, layer, row, t, ug * 4 + b);
#pragma unroll
            for (int j = 0; j < 8; ++j) w |= (dropout_lane16(blk, j) >= thr ? 1u : 0u) << (b * 8 + j);
        }
        bits[i] = w;
    }
}

// The backward pass reads what the forward pass left in the workspace (gates, images, cell state, keep bits, packed weights).  The
// forward writes a 32-byte tag {magic, seed if dropout is on, N, T | L | H | dropout threshold}; every backward launch checks it against
// its own arguments and aborts (trap) on a foreign, reused or differently sized workspace instead of producing gradients from stale data.
struct W2Tag { uint64_t v[4]; };
inline W2Tag make_tag(uint64_t seed, bool drop, float p, int64_t N, int T, int L, int H) {
    W2Tag t;
    t.v[0] = 0x77326c73746d3032ull;                       // "w2lstm02"
    t.v[1] = drop ? seed : 0ull;
    t.v[2] = (uint64_t)N;
    t.v[3] = ((uint64_t)(uint32_t)T << 48) | ((uint64_t)(uint32_t)L << 40) | ((uint64_t)(uint32_t)H << 24) | (drop ? dropout_threshold(p) + 1u : 0u);
    return t;
}
__global__ void w2_tag_kernel(uint64_t* tag, W2Tag t) {
    for (int i = 0; i < 4; ++i) tag[i] = t.v[i];
}

// hidden states of one layer as fp32 [T][N][H] (b200_lstm_hidden_states)
__global__ void __launch_bounds__(256)
w2_unpack_kernel(const uint8_t* __restrict__ himg, int64_t N, int T, int H, int NT, float* __restrict__ out) {
    const int HC = H / 128;
    const int64_t total = (int64_t)T * N * (H / 8);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int c8 = (int)(i % (H / 8));
        const int64_t tr = i / (H / 8), row = tr % N, t = tr / N;
        const int64_t tile = row / TM;
        const int r = (int)(row % TM), col = c8 * 8;
        const uint4 q = *reinterpret_cast<const uint4*>(himg + (((size_t)t * NT + tile) * HC + (col >> 7)) * WCHUNK + wchunk_off(r, col & 127));
        float v[8];
        unpack_bf8(q, v);
        st8(out + (t * N + row) * H + col, v);
    }
}


// ------------------------------------------------------------------------------------------------ forward step
// One LSTM step of one layer for all rows.  CTA c owns unit group ug = c % UG (32 hidden units = 128 permuted gate columns, its
// KC weight chunks stay in shared memory for the whole launch) and the row tiles m = c / UG, c / UG + gridDim.x / UG, ...
// Per tile: XC + HC bulk copies of 32 KB bring [x_t | h_{t-1}], 8 MMAs each (M 128, N 128, K 16) accumulate the gates into one of four
// TMEM slots, and one of three epilogue warp groups (taking tiles in turn) applies the cell and writes c_t, the activated gates and its
// 32 columns of the h_t image.  Activations through tanh.approx (see tanha).
struct FwdArgs {
    const uint8_t* wf; const float* bias;
    const uint8_t* xin; const uint8_t* hprev; uint8_t* hout; uint8_t* hmask; const uint32_t* keep;
    const float* c_prev; float* c_out; uint8_t* gates; float* h_last;
    int64_t N; int NT, XC, HC, UG, GC, H;
    int dbg;       // timing experiments (option "wide_dbg", results are wrong when set): 1 = no cell math / stores, 2 = no operand loads,
                   // 8 = cell math but no stores, 16 = no c_{t-1} loads
    unsigned long long* trace;   // option "wide_dbg" bit 32: clock64 stamps of CTA 0, trace[(role * 256 + tile) * 4 + event] (tools/wide2_trace.py)
};
__device__ __forceinline__ void w2_stamp(unsigned long long* tr, int role, int tile, int ev) {
    if (tr != nullptr && tile < 256) tr[(role * 256 + tile) * 4 + ev] = (unsigned long long)clock64();
}
constexpr int FWD_NST = 3;
constexpr int FWD_GROUPS = 3;                            // epilogue warp groups (4 warps each) taking the tiles of a CTA in turn
constexpr int FWD_THREADS = (FWD_GROUPS * 4 + 2) * 32;
#ifndef W2_PF_DIST
#define W2_PF_DIST 2
#endif
constexpr uint32_t FWD_W_MAX = 8 * CHUNK;
struct FwdSmem {
    static constexpr uint32_t off_w = 0, off_ring = FWD_W_MAX, off_bias = off_ring + FWD_NST * WCHUNK, off_bar = off_bias + 128 * 4;
    // full[NST], empty[NST], acc_full[4], acc_free[4], w_ready
    static constexpr uint32_t nbar = 2 * FWD_NST + 9, off_tmem = off_bar + nbar * 8, total = off_tmem + 16;
};

// MC: launched as clusters of two CTAs = two unit groups that walk the same row tiles.  Each CTA fetches one 64-row half (16 KB) of
// every activation chunk and multicasts it into both CTAs' rings, so that a chunk crosses the L2 -> SM path once per CTA pair; a
// stage is released when the MMAs of BOTH CTAs have read it (multicast tcgen05.commit on the `empty` barriers, count 2).
template <bool MC>
__global__ void __launch_bounds__(FWD_THREADS, 1) w2_fwd_kernel(FwdArgs a) {
    using S = FwdSmem;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + S::off_bar);
    uint64_t* empty = full + FWD_NST;
    uint64_t* acc_full = empty + FWD_NST;
    uint64_t* acc_free = acc_full + 4;
    uint64_t* w_ready = acc_free + 4;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S::off_tmem);
    float* bias_s = reinterpret_cast<float*>(smem + S::off_bias);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ug = (int)(blockIdx.x % a.UG), first = (int)(blockIdx.x / a.UG), stride = (int)(gridDim.x / a.UG);
    const int KC = 2 * (a.XC + a.HC);                   // weight images of this unit group
    const int nk = a.hprev ? a.XC + a.HC : a.XC;        // activation chunks per tile; t = 0: h_{-1} = 0, the recurrent ones are skipped

    if (threadIdx.x == 0) {
        for (int i = 0; i < FWD_NST; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], MC ? 2 : 1); }
        for (int i = 0; i < 4; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_free[i], 4); }
        mbar_init(w_ready, 1);
        fence_barrier_init();
    }
    constexpr int LD_WARP = FWD_GROUPS * 4, MMA_WARP = LD_WARP + 1;
    const uint32_t crank = MC ? (blockIdx.x & 1u) : 0u;         // rank in the CTA pair (cluster dims 2 x 1 x 1)
    if (threadIdx.x < 128) bias_s[threadIdx.x] = a.bias[ug * 128 + threadIdx.x];
    if (warp == MMA_WARP) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    if (MC) cluster_sync_all();                                 // the peer's barriers exist before anything is sent to them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < LD_WARP) {
        // ================================ epilogue: LSTM cell of 32 units of one row ==========================
        const int grp = warp >> 2, lq = warp & 3, r = lq * 32 + lane;
        const uint32_t row_off = (uint32_t)(r >> 3) * WRG_BYTES + (uint32_t)(r & 7) * 16u;
        const uint32_t tl = tmem_base + ((uint32_t)(lq * 32) << 16);
        int i = 0;
        for (int m = first; m < a.NT; m += stride, ++i) {
            if (i % FWD_GROUPS != grp) continue;
            const int slot = i & 3;
            const int64_t row = (int64_t)m * TM + r;
            const bool live = row < a.N;
            float cpv[4][8];                      // c_{t-1} of the row's 32 units, fetched while the MMAs of the tile run
#pragma unroll
            for (int uc = 0; uc < 4; ++uc) {
                if (a.c_prev != nullptr && !(a.dbg & 17)) ld8(a.c_prev + slot_off(m, ug * 4 + uc, r, a.H), cpv[uc]);
                else {
#pragma unroll
                    for (int j = 0; j < 8; ++j) cpv[uc][j] = 0.f;
                }
            }
            uint32_t keep = 0u;                   // keep bits of the row's 32 units (layer above), fetched before the wait as well
            if (a.hmask != nullptr && live) keep = a.keep[row * a.UG + ug];
            unsigned long long* const tr = (blockIdx.x == 0 && lq == 0 && lane == 0) ? a.trace : nullptr;
            w2_stamp(tr, 0, i, 0);
            mbar_wait(&acc_full[slot], (uint32_t)((i >> 2) & 1));
            tc_fence_after();
            w2_stamp(tr, 0, i, 1);
            uint8_t* gbase = a.gates + ((size_t)m * a.UG + ug) * WCHUNK + row_off;
            const size_t hoff = ((size_t)m * a.HC + (ug >> 2)) * WCHUNK + row_off + (uint32_t)((ug & 3) * 4) * 128u;
            uint8_t* hdst = a.hout + hoff;
            uint8_t* mdst = a.hmask ? a.hmask + hoff : nullptr;
#pragma unroll
            for (int uc = 0; uc < 4; ++uc) {
                if (a.dbg & 1) break;
                float ai[8], af[8], ag[8], ao[8];
                const float (&cp)[8] = cpv[uc];
                const uint32_t col = tl + (uint32_t)(slot * 128 + uc * 8);
                tmem_ld8(col + 0 * 32, ai);
                tmem_ld8(col + 1 * 32, af);
                tmem_ld8(col + 2 * 32, ag);
                tmem_ld8(col + 3 * 32, ao);
                tmem_ld_wait();
                float cv[8], hv[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int u = uc * 8 + j;
                    const float gi = fmaf(tanha(fmaf(ai[j], 0.5f, bias_s[0 * 32 + u])), 0.5f, 0.5f);
                    const float gf = fmaf(tanha(fmaf(af[j], 0.5f, bias_s[1 * 32 + u])), 0.5f, 0.5f);
                    const float gg = tanha(ag[j] + bias_s[2 * 32 + u]);
                    const float go = fmaf(tanha(fmaf(ao[j], 0.5f, bias_s[3 * 32 + u])), 0.5f, 0.5f);
                    const float c = fmaf(gf, cp[j], gi * gg);
                    ai[j] = gi; af[j] = gf; ag[j] = gg; ao[j] = go;
                    cv[j] = c;
                    hv[j] = go * tanha(c);
                }
                if (a.dbg & 8) {        // keep the math alive without storing anything (the condition never holds)
                    if (cv[0] == 123.456f && hv[1] == ai[2] + af[3] + ag[4] + ao[5]) a.c_out[0] = 1.f;
                    continue;
                }
                // activated gates: chunk ug, 16-byte slot gate * 4 + uc of this row
                *reinterpret_cast<uint4*>(gbase + (uint32_t)(0 * 4 + uc) * 128u) = pack_bf8(ai);
                *reinterpret_cast<uint4*>(gbase + (uint32_t)(1 * 4 + uc) * 128u) = pack_bf8(af);
                *reinterpret_cast<uint4*>(gbase + (uint32_t)(2 * 4 + uc) * 128u) = pack_bf8(ag);
                *reinterpret_cast<uint4*>(gbase + (uint32_t)(3 * 4 + uc) * 128u) = pack_bf8(ao);
                const uint4 h8 = pack_bf8(hv);
                *reinterpret_cast<uint4*>(hdst + (uint32_t)uc * 128u) = h8;
                if (mdst != nullptr) *reinterpret_cast<uint4*>(mdst + (uint32_t)uc * 128u) = and_mask8(h8, (keep >> (uc * 8)) & 0xFFu);
                st8(a.c_out + slot_off(m, ug * 4 + uc, r, a.H), cv);
                if (live && a.h_last != nullptr) st8(a.h_last + row * a.H + ug * 32 + uc * 8, hv);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_free[slot]);
            w2_stamp(tr, 0, i, 2);
        }
    } else if (warp == LD_WARP) {
        // ================================ loader ================================================================
        if (lane == 0) {
            mbar_arrive_expect_tx(w_ready, (uint32_t)KC * CHUNK);
            for (int kc = 0; kc < KC; ++kc) bulk_copy_g2s(smem + S::off_w + kc * CHUNK, a.wf + ((size_t)ug * KC + kc) * CHUNK, CHUNK, w_ready);
            uint32_t q = 0;
            unsigned long long* const tr = blockIdx.x == 0 ? a.trace : nullptr;
            int li = 0;
            for (int m = first; m < a.NT; m += stride, ++li) {
                w2_stamp(tr, 2, li, 0);
                const int mp = m + W2_PF_DIST * stride;          // L2 prefetch: the layer input of a later tile comes from HBM
                if (W2_PF_DIST > 0 && mp < a.NT) {
                    if (ug < 2 * a.XC) prefetch_l2(a.xin + ((size_t)mp * a.XC + (ug >> 1)) * WCHUNK + (size_t)(ug & 1) * (WCHUNK / 2), WCHUNK / 2);
                    if (a.c_prev != nullptr) prefetch_l2(a.c_prev + slot_off(mp, ug * 4, 0, a.H), 16384);
                }
                for (int kc = 0; kc < nk; ++kc, ++q) {
                    const uint32_t st = q % FWD_NST;
                    if (q >= FWD_NST) mbar_wait(&empty[st], ((q / FWD_NST) - 1) & 1);
                    const uint8_t* src = kc < a.XC ? a.xin + ((size_t)m * a.XC + kc) * WCHUNK : a.hprev + ((size_t)m * a.HC + (kc - a.XC)) * WCHUNK;
                    if (a.dbg & 2) { mbar_arrive(&full[st]); continue; }
                    mbar_arrive_expect_tx(&full[st], WCHUNK);
                    if (MC) bulk_copy_g2s_mc(smem + S::off_ring + st * WCHUNK + crank * (WCHUNK / 2), src + crank * (WCHUNK / 2), WCHUNK / 2, &full[st], (uint16_t)3);
                    else bulk_copy_g2s(smem + S::off_ring + st * WCHUNK, src, WCHUNK, &full[st]);
                    if (kc == 0) w2_stamp(tr, 2, li, 1);
                }
                w2_stamp(tr, 2, li, 2);
            }
        }
        __syncwarp();
    } else {
        // ================================ MMA issuer ============================================================
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_bf16(TM, 128);
            const uint32_t s_base = smem_u32(smem);
            mbar_wait(w_ready, 0u);
            uint32_t q = 0;
            int i = 0;
            for (int m = first; m < a.NT; m += stride, ++i) {
                const int slot = i & 3;
                unsigned long long* const tr = blockIdx.x == 0 ? a.trace : nullptr;
                w2_stamp(tr, 1, i, 0);
                if (i >= 4) mbar_wait(&acc_free[slot], (uint32_t)(((i >> 2) - 1) & 1));
                tc_fence_after();
                w2_stamp(tr, 1, i, 1);
                const uint32_t d = tmem_base + (uint32_t)(slot * 128);
                for (int kc = 0; kc < nk; ++kc, ++q) {
                    const uint32_t st = q % FWD_NST;
                    mbar_wait(&full[st], (q / FWD_NST) & 1);
                    tc_fence_after();
                    if (kc == 0) w2_stamp(tr, 1, i, 2);
                    const uint32_t aa = s_base + S::off_ring + st * WCHUNK, bb = s_base + S::off_w + (uint32_t)(2 * kc) * CHUNK;
#pragma unroll
                    for (int ks = 0; ks < 8; ++ks)          // two K = 64 halves of the activation chunk, one weight image each
                        umma_bf16(d, make_desc(aa + (ks >> 2) * 1024 + (ks & 3) * 256, LBO_BYTES, WRG_BYTES),
                                  make_desc(bb + (ks >> 2) * CHUNK + (ks & 3) * 256, LBO_BYTES, RG_BYTES), idesc, (kc > 0 || ks > 0) ? 1u : 0u);
                    if (MC) umma_commit_mc(&empty[st], (uint16_t)3); else umma_commit(&empty[st]);
                }
                umma_commit(&acc_full[slot]);
                w2_stamp(tr, 1, i, 3);
            }
        }
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    if (MC) cluster_sync_all();                                 // no CTA leaves while its peer may still signal its barriers
    if (warp == MMA_WARP) tmem_dealloc(tmem_base, 512);
}


// ------------------------------------------------------------------------------------------------ backward cell
// One block = the 128 rows x 32 hidden units of (row tile m, unit group ug).  Everything the block touches is contiguous in global
// memory (the gate chunk, 32 KB; the row-slot slabs of c_{t-1} and dc, 16 KB each; of dh and d_out, 8 KB each), so the block
// brings it into shared memory with five bulk copies, computes from there, and sends dG (in place) and dc_{t-1} back with two bulk
// stores: no per-thread global access is left (ncu of the thread-load version: L1 71 % busy at 60 % of DRAM throughput).
// Two blocks per SM overlap each other's copy and compute phases.  Output: dG (pre-activation gradients, bf16, rows >= N zero),
// dc_{t-1}, and the tile's column sums of dG (the bias gradient) added to the block's own slot of bsum.
struct CellArgs {
    uint8_t* G; const float* c_prev; const uint4* dh_rec; const uint4* d_out; const float* dhT; float* dc; float* bsum;
    const uint64_t* tag; W2Tag want;
    int first; int64_t N; int UG, GC, H;
};
constexpr int CELL_THREADS = 256;
struct CellSmem {
    static constexpr uint32_t off_g = 0, off_cp = WCHUNK, off_dc = off_cp + 16384, off_dh = off_dc + 16384, off_do = off_dh + 8192;
    static constexpr uint32_t off_red = off_do + 8192, off_bar = off_red + 8 * 4 * 32 * 4, total = off_bar + 16;
};

__global__ void __launch_bounds__(CELL_THREADS, 2) w2_cell_bwd_kernel(CellArgs a) {
    using S = CellSmem;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + S::off_bar);
    float (*red)[4][32] = reinterpret_cast<float (*)[4][32]>(smem + S::off_red);
    if (blockIdx.x == 0 && threadIdx.x == 0 &&
        (a.tag[0] != a.want.v[0] || a.tag[1] != a.want.v[1] || a.tag[2] != a.want.v[2] || a.tag[3] != a.want.v[3]))
        __trap();                                             // not the workspace this step's forward pass filled
    const int m = (int)(blockIdx.x / a.UG), ug = (int)(blockIdx.x % a.UG);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t* gsrc = a.G + ((size_t)m * a.UG + ug) * WCHUNK;
    const size_t e0 = slot_off(m, ug * 4, 0, a.H);             // first float of the block's four octet slabs (4 x 128 rows x 8)
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        fence_barrier_init();
        const uint32_t bytes = WCHUNK + (a.c_prev ? 16384u : 0u) + (a.first ? 0u : 16384u) + (a.dh_rec ? 8192u : 0u) + (a.d_out ? 8192u : 0u);
        mbar_arrive_expect_tx(bar, bytes);
        bulk_copy_g2s(smem + S::off_g, gsrc, WCHUNK, bar);
        if (a.c_prev) bulk_copy_g2s(smem + S::off_cp, a.c_prev + e0, 16384u, bar);
        if (!a.first) bulk_copy_g2s(smem + S::off_dc, a.dc + e0, 16384u, bar);
        if (a.dh_rec) bulk_copy_g2s(smem + S::off_dh, a.dh_rec + (e0 >> 3), 8192u, bar);
        if (a.d_out) bulk_copy_g2s(smem + S::off_do, a.d_out + (e0 >> 3), 8192u, bar);
    }
    __syncthreads();
    mbar_wait(bar, 0u);
    // a warp step = one 8-row group x the four unit octets of the group: lane = (octet, row in group); the 8 lanes of a quarter warp
    // read 128 contiguous bytes of an image slot
    const int rr = lane & 7, uc = lane >> 3;
    float acc[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) acc[k] = 0.f;
#pragma unroll 1
    for (int it = 0; it < 2; ++it) {
        const int rg = warp * 2 + it, r = rg * 8 + rr;
        const int64_t row = (int64_t)m * TM + r;
        const bool live = row < a.N;
        uint8_t* gbase = smem + S::off_g + (uint32_t)rg * WRG_BYTES + (uint32_t)rr * 16u;
        const uint32_t so = (uint32_t)(uc * 128 + r);          // slot of (octet, row) inside the block's slabs
        float gi[8], gf[8], gg[8], go[8], cp[8], dh[8], dc[8], v[32];
        unpack_bf8(*reinterpret_cast<const uint4*>(gbase + (uint32_t)(0 * 4 + uc) * 128u), gi);
        unpack_bf8(*reinterpret_cast<const uint4*>(gbase + (uint32_t)(1 * 4 + uc) * 128u), gf);
        unpack_bf8(*reinterpret_cast<const uint4*>(gbase + (uint32_t)(2 * 4 + uc) * 128u), gg);
        unpack_bf8(*reinterpret_cast<const uint4*>(gbase + (uint32_t)(3 * 4 + uc) * 128u), go);
#pragma unroll
        for (int j = 0; j < 8; ++j) { cp[j] = 0.f; dh[j] = 0.f; dc[j] = 0.f; }
        {
            float t8[8];
            if (a.c_prev) ld8(reinterpret_cast<const float*>(smem + S::off_cp) + so * 8, cp);
            if (a.dh_rec) unpack_bf8(reinterpret_cast<const uint4*>(smem + S::off_dh)[so], dh);
            if (a.d_out) {
                unpack_bf8(reinterpret_cast<const uint4*>(smem + S::off_do)[so], t8);
#pragma unroll
                for (int j = 0; j < 8; ++j) dh[j] += t8[j];
            }
            if (a.dhT && live) {
                ld8(a.dhT + row * a.H + ug * 32 + uc * 8, t8);
#pragma unroll
                for (int j = 0; j < 8; ++j) dh[j] += t8[j];
            }
            if (!a.first) ld8(reinterpret_cast<const float*>(smem + S::off_dc) + so * 8, dc);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float c = fmaf(gf[j], cp[j], gi[j] * gg[j]);
            const float tc = tanha(c);
            const float dct = live ? fmaf(dh[j] * go[j], 1.f - tc * tc, dc[j]) : 0.f;
            const float d_o = live ? dh[j] * tc : 0.f;
            v[0 * 8 + j] = dct * gg[j] * gi[j] * (1.f - gi[j]);
            v[1 * 8 + j] = dct * cp[j] * gf[j] * (1.f - gf[j]);
            v[2 * 8 + j] = dct * gi[j] * (1.f - gg[j] * gg[j]);
            v[3 * 8 + j] = d_o * go[j] * (1.f - go[j]);
            dc[j] = dct * gf[j];
        }
        st8(reinterpret_cast<float*>(smem + S::off_dc) + so * 8, dc);
        {
            float t8[8];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
#pragma unroll
                for (int j = 0; j < 8; ++j) t8[j] = v[k * 8 + j];
                *reinterpret_cast<uint4*>(gbase + (uint32_t)(k * 4 + uc) * 128u) = pack_bf8(t8);
            }
        }
#pragma unroll
        for (int k = 0; k < 32; ++k) acc[k] += v[k];
    }
    // results back to global memory through the async proxy
    fence_proxy_async();
    // bias gradient: sum the 32 values (4 gates x 8 units of the lane's octet) over the 8 rows of the warp step; lane (uc, rr) ends up
    // with values rr * 4 .. rr * 4 + 3
#pragma unroll
    for (int S2 = 16, LB = 4; S2 >= 4; S2 >>= 1, LB >>= 1) {
        const bool up = (lane & LB) != 0;
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            if (k < S2) {
                const float lo = acc[k], hi = acc[k + S2];
                acc[k] = (up ? hi : lo) + __shfl_xor_sync(0xffffffffu, up ? lo : hi, LB);
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) red[warp][uc][rr * 4 + k] = acc[k];
    __syncthreads();
    if (threadIdx.x == 0) {
        bulk_copy_s2g(gsrc, smem + S::off_g, WCHUNK);
        bulk_copy_s2g(a.dc + e0, smem + S::off_dc, 16384u);
        bulk_commit_wait_read();
    }
    if (threadIdx.x < 128) {
        const int o = threadIdx.x >> 5, idx = threadIdx.x & 31, gate = idx >> 3, j = idx & 7;
        float sum = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) sum += red[w][o][idx];
        a.bsum[(size_t)blockIdx.x * 128 + gate * 32 + o * 8 + j] += sum;
    }
}

// gb_ih = gb_hh = column sums over the row tiles (fixed order)
__global__ void __launch_bounds__(256)
w2_bsum_reduce_kernel(const float* __restrict__ bsum, int NT, int UG, int H, float* __restrict__ gb_ih, float* __restrict__ gb_hh) {
    const int gp = blockIdx.x * blockDim.x + threadIdx.x;
    if (gp >= 4 * H) return;
    const int ug = gp >> 7, gate = (gp >> 5) & 3, u = ug * 32 + (gp & 31);
    float s = 0.f;
    for (int m = 0; m < NT; ++m) s += bsum[((size_t)m * UG + ug) * 128 + (gp & 127)];
    gb_ih[gate * H + u] = s;
    gb_hh[gate * H + u] = s;
}

// ------------------------------------------------------------------------------------------------ data gradient
// out[p] = dG_t . Wd[p]  (p = 0: dh_{t-1} = dG . W_hh, p = 1: dx_t = dG . W_ih, masked with the keep bits of this layer's input;
// the 1/(1-p) is in the packed W_ih).  A task = two row tiles x one part.  Per unit group (K = 128 gate columns) the two dG chunks
// of the tiles arrive with one 32 KB bulk copy each (A ring, 2 slots of 64 KB); the matching [64 x H] halves of the weight image
// arrive separately (B ring, 3 slots of 32 KB); 8 MMAs per half (M 128, N = H, B read MN-major) feed two TMEM accumulators.
struct DgradArgs {
    const uint8_t* dG; const uint8_t* wd[2]; uint4* out[2]; const uint32_t* keep;
    int np; int part[2];
    int64_t N; int NT, H, UG;
    int dbg;       // 1 = no epilogue stores, 2 = no operand loads, 4 = no weight loads only
};
constexpr int DG_NSA = 2, DG_NSB = 3;
struct DgSmem {
    static constexpr uint32_t off_b = DG_NSA * 2 * WCHUNK, off_bar = off_b + DG_NSB * WCHUNK;
    // fullA[2], emptyA[2], fullB[3], emptyB[3], acc_full, acc_free
    static constexpr uint32_t off_tmem = off_bar + (2 * DG_NSA + 2 * DG_NSB + 2) * 8, total = off_tmem + 16;
};

__global__ void __launch_bounds__(W2_THREADS, 1) w2_dgrad_kernel(DgradArgs a) {
    using S = DgSmem;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* fullA = reinterpret_cast<uint64_t*>(smem + S::off_bar);
    uint64_t* emptyA = fullA + DG_NSA;
    uint64_t* fullB = emptyA + DG_NSA;
    uint64_t* emptyB = fullB + DG_NSB;
    uint64_t* acc_full = emptyB + DG_NSB;
    uint64_t* acc_free = acc_full + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S::off_tmem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ntask = ((a.NT + 1) / 2) * a.np;
    const uint32_t b_bytes = (uint32_t)(64 * a.H * 2), rgs = (uint32_t)(a.H / 8) * 128u;

    if (threadIdx.x == 0) {
        for (int i = 0; i < DG_NSA; ++i) { mbar_init(&fullA[i], 1); mbar_init(&emptyA[i], 1); }
        for (int i = 0; i < DG_NSB; ++i) { mbar_init(&fullB[i], 1); mbar_init(&emptyB[i], 1); }
        mbar_init(acc_full, 1);
        mbar_init(acc_free, 8);
        fence_barrier_init();
    }
    if (warp == 9) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < 8) {
        const int j = warp >> 2, lq = warp & 3, r = lq * 32 + lane;
        const uint32_t tl = tmem_base + ((uint32_t)(lq * 32) << 16) + (uint32_t)(j * 256);
        int i = 0;
        for (int task = blockIdx.x; task < ntask; task += gridDim.x, ++i) {
            const int p = (task % a.np) == 0 ? a.part[0] : a.part[1], m = 2 * (task / a.np) + j;
            uint4* const outp = p == 0 ? a.out[0] : a.out[1];       // (selects, not indexed kernel parameters: no local-memory copy)
            const int64_t row = (int64_t)m * TM + r;
            // the row's keep words are fetched while the MMAs of the task run (one global-load latency per 32 columns inside the
            // loop below made the epilogue of the dx tasks ~11 k clk long)
            uint32_t kws[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) kws[k] = 0xFFFFFFFFu;
            if (p == 1 && a.keep != nullptr && m < a.NT && row < a.N) {
#pragma unroll
                for (int k = 0; k < 8; ++k) if (k < a.UG) kws[k] = a.keep[row * a.UG + k];
            }
            mbar_wait(acc_full, (uint32_t)(i & 1));
            tc_fence_after();
            if (m < a.NT && !(a.dbg & 1)) {
#pragma unroll
                for (int cb = 0; cb < 8; ++cb) {
                    const int c0 = cb * 32;
                    if (c0 >= a.H) break;
                    float v[4][8];
#pragma unroll
                    for (int k = 0; k < 4; ++k) tmem_ld8(tl + (uint32_t)(c0 + k * 8), v[k]);
                    const uint32_t kw = kws[cb];
                    tmem_ld_wait();
#pragma unroll
                    for (int k = 0; k < 4; ++k)          // bf16, one 16-byte slot per (row, octet): a warp store is 512 contiguous bytes
                        outp[slot_off(m, (c0 >> 3) + k, r, a.H) >> 3] = and_mask8(pack_bf8(v[k]), (kw >> (k * 8)) & 0xFFu);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(acc_free);
        }
    } else if (warp == 8) {
        if (lane == 0) {
            uint32_t qa = 0, qb = 0;
            for (int task = blockIdx.x; task < ntask; task += gridDim.x) {
                const int p = (task % a.np) == 0 ? a.part[0] : a.part[1], m0 = 2 * (task / a.np);
                const uint8_t* const wdp = p == 0 ? a.wd[0] : a.wd[1];
                const bool two = m0 + 1 < a.NT;
                for (int ug = 0; ug < a.UG; ++ug, ++qa) {
                    const uint32_t sa = qa % DG_NSA;
                    if (qa >= DG_NSA) mbar_wait(&emptyA[sa], ((qa / DG_NSA) - 1) & 1);
                    uint8_t* da = smem + sa * 2 * WCHUNK;
                    if (a.dbg & 2) mbar_arrive(&fullA[sa]);
                    else {
                        mbar_arrive_expect_tx(&fullA[sa], (two ? 2u : 1u) * WCHUNK);
                        bulk_copy_g2s(da, a.dG + ((size_t)m0 * a.UG + ug) * WCHUNK, WCHUNK, &fullA[sa]);
                        if (two) bulk_copy_g2s(da + WCHUNK, a.dG + ((size_t)(m0 + 1) * a.UG + ug) * WCHUNK, WCHUNK, &fullA[sa]);
                    }
                    for (int half = 0; half < 2; ++half, ++qb) {
                        const uint32_t sb = qb % DG_NSB;
                        if (qb >= DG_NSB) mbar_wait(&emptyB[sb], ((qb / DG_NSB) - 1) & 1);
                        if (a.dbg & 6) { mbar_arrive(&fullB[sb]); continue; }
                        mbar_arrive_expect_tx(&fullB[sb], b_bytes);
                        bulk_copy_g2s(smem + S::off_b + sb * WCHUNK, wdp + (size_t)(ug * 2 + half) * b_bytes, b_bytes, &fullB[sb]);
                    }
                }
            }
        }
        __syncwarp();
    } else {
        if (lane == 0) {
            const uint32_t idesc = make_idesc_bf16(TM, a.H, 0, 1);
            const uint32_t s_base = smem_u32(smem);
            uint32_t qa = 0, qb = 0;
            int i = 0;
            for (int task = blockIdx.x; task < ntask; task += gridDim.x, ++i) {
                const bool two = 2 * (task / a.np) + 1 < a.NT;
                if (i > 0) mbar_wait(acc_free, (uint32_t)((i - 1) & 1));
                tc_fence_after();
                for (int ug = 0; ug < a.UG; ++ug, ++qa) {
                    const uint32_t sa = qa % DG_NSA;
                    mbar_wait(&fullA[sa], (qa / DG_NSA) & 1);
                    for (int half = 0; half < 2; ++half, ++qb) {
                        const uint32_t sb = qb % DG_NSB;
                        mbar_wait(&fullB[sb], (qb / DG_NSB) & 1);
                        tc_fence_after();
                        const uint32_t da = s_base + sa * 2 * WCHUNK + half * 1024, db = s_base + S::off_b + sb * WCHUNK;
#pragma unroll
                        for (int t2 = 0; t2 < 2; ++t2) {
                            if (t2 == 1 && !two) break;
#pragma unroll
                            for (int ks = 0; ks < 4; ++ks)
                                umma_bf16(tmem_base + (uint32_t)(t2 * 256), make_desc(da + t2 * WCHUNK + ks * 256, LBO_BYTES, WRG_BYTES),
                                          make_desc(db + ks * 2 * rgs, rgs, LBO_BYTES), idesc, (ug > 0 || half > 0 || ks > 0) ? 1u : 0u);
                        }
                        umma_commit(&emptyB[sb]);
                    }
                    umma_commit(&emptyA[sa]);
                }
                umma_commit(acc_full);
            }
        }
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 9) tmem_dealloc(tmem_base, 512);
}

// ------------------------------------------------------------------------------------------------ weight gradient
// part[z] = dG^T . [x | h_{t-1}] over the K-chunks (64 rows of one (t, row tile)) of split z; both operands are read MN-major: a
// stage is the 64-row halves (16 KB, contiguous) of two dG chunks (two unit groups = two M tiles) and of the XC or HC activation
// chunks of the x part (nt = 0) or the h part (nt = 1); 4 MMAs (M 128, N 128, K 16) per (M tile, activation chunk).
struct WgradArgs {
    const uint8_t* dG; const uint8_t* xsrc; const uint8_t* hsrc; const uint8_t* zero; float* part;
    int T, NT, UG, XC, HC, H;
    int splits; int64_t kps;
};
constexpr int WG_NST = 3;
constexpr uint32_t WG_STAGE = 2 * WCHUNK, WG_HALF = WCHUNK / 2;
struct WgSmem {
    static constexpr uint32_t off_bar = WG_NST * WG_STAGE;          // full[NST], empty[NST], acc_full
    static constexpr uint32_t off_tmem = off_bar + (2 * WG_NST + 1) * 8, total = off_tmem + 16;
};

__global__ void __launch_bounds__(W2_THREADS, 1) w2_wgrad_kernel(WgradArgs a) {
    using S = WgSmem;
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + S::off_bar);
    uint64_t* empty = full + WG_NST;
    uint64_t* acc_full = empty + WG_NST;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + S::off_tmem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ntask = (a.H / 64) * 2;                          // (M pair, N tile)
    const int z = (int)(blockIdx.x / ntask), task = (int)(blockIdx.x % ntask);
    const int mp = task >> 1, nt = task & 1;
    const int nch = nt == 0 ? a.XC : a.HC;                     // activation chunks (128 columns each) of this N tile
    const int KW = (a.XC + a.HC) * 128, G4 = 4 * a.H;
    const int64_t ktot = (int64_t)a.T * a.NT * 2;
    const int64_t k_beg = (int64_t)z * a.kps, k_end = k_beg + a.kps < ktot ? k_beg + a.kps : ktot;

    if (threadIdx.x == 0) {
        for (int i = 0; i < WG_NST; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        mbar_init(acc_full, 1);
        fence_barrier_init();
    }
    if (warp == 9) tmem_alloc(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp < 8) {
        const int mt = warp >> 2, lq = warp & 3, r = lq * 32 + lane;
        const uint32_t tl = tmem_base + ((uint32_t)(lq * 32) << 16) + (uint32_t)(mt * 256);
        const bool any = k_end > k_beg;
        if (any) {
            mbar_wait(acc_full, 0u);
            tc_fence_after();
        }
        float* dst = a.part + ((size_t)z * G4 + (size_t)(2 * mp + mt) * 128 + r) * KW + (nt == 0 ? 0 : a.XC * 128);
        for (int c0 = 0; c0 < nch * 128; c0 += 32) {
            float v[4][8];
            if (any) {
#pragma unroll
                for (int k = 0; k < 4; ++k) tmem_ld8(tl + (uint32_t)(c0 + k * 8), v[k]);
                tmem_ld_wait();
            } else {
#pragma unroll
                for (int k = 0; k < 4; ++k)
#pragma unroll
                    for (int q = 0; q < 8; ++q) v[k][q] = 0.f;
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) st8(dst + c0 + k * 8, v[k]);
        }
    } else if (warp == 8) {
        if (lane == 0) {
            uint32_t q = 0;
            for (int64_t kq = k_beg; kq < k_end; ++kq, ++q) {
                const uint32_t st = q % WG_NST;
                const int64_t tm = kq >> 1;                        // t * NT + m
                const int half = (int)(kq & 1), t = (int)(tm / a.NT);
                if (q >= WG_NST) mbar_wait(&empty[st], ((q / WG_NST) - 1) & 1);
                uint8_t* sa = smem + st * WG_STAGE;
                uint8_t* sb = sa + WCHUNK;
                mbar_arrive_expect_tx(&full[st], (uint32_t)(2 + nch) * WG_HALF);
                for (int mt = 0; mt < 2; ++mt)
                    bulk_copy_g2s(sa + mt * WG_HALF, a.dG + ((size_t)tm * a.UG + (2 * mp + mt)) * WCHUNK + (size_t)half * WG_HALF, WG_HALF, &full[st]);
                for (int ci = 0; ci < nch; ++ci) {
                    const uint8_t* ch = nt == 0 ? a.xsrc + ((size_t)tm * a.XC + ci) * WCHUNK + (size_t)half * WG_HALF
                                                : (t > 0 ? a.hsrc + ((size_t)(tm - a.NT) * a.HC + ci) * WCHUNK + (size_t)half * WG_HALF : a.zero);
                    bulk_copy_g2s(sb + ci * WG_HALF, ch, WG_HALF, &full[st]);
                }
            }
        }
        __syncwarp();
    } else {
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_bf16(TM, 128, 1, 1);
            const uint32_t s_base = smem_u32(smem);
            uint32_t q = 0;
            for (int64_t kq = k_beg; kq < k_end; ++kq, ++q) {
                const uint32_t st = q % WG_NST;
                mbar_wait(&full[st], (q / WG_NST) & 1);
                tc_fence_after();
                const uint32_t sa = s_base + st * WG_STAGE, sb = sa + WCHUNK;
#pragma unroll
                for (int mt = 0; mt < 2; ++mt)
                    for (int ci = 0; ci < nch; ++ci)
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks)
                            umma_bf16(tmem_base + (uint32_t)(mt * 256 + ci * 128), make_desc(sa + mt * WG_HALF + ks * 2 * WRG_BYTES, WRG_BYTES, LBO_BYTES),
                                      make_desc(sb + ci * WG_HALF + ks * 2 * WRG_BYTES, WRG_BYTES, LBO_BYTES), idesc, (q > 0 || ks > 0) ? 1u : 0u);
                umma_commit(&empty[st]);
            }
            if (k_end > k_beg) umma_commit(acc_full);
        }
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 9) tmem_dealloc(tmem_base, 512);
}

// grads of W_ih [4H][in] and W_hh [4H][H] in torch order = sum over the splits (fixed order), x part scaled by 1/(1-p)
__global__ void __launch_bounds__(256)
w2_wgrad_reduce_kernel(const float* __restrict__ part, int splits, int H, int XC /* K = 64 images of the x part */, int in, float x_scale, float* __restrict__ gw_ih,
                       float* __restrict__ gw_hh) {
    const int G4 = 4 * H, KW = (XC + H / 64) * 64;
    const int64_t total = (int64_t)G4 * (in + H);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int g = (int)(i / (in + H)), k = (int)(i % (in + H));
        const int gate = g / H, u = g % H;
        const int gp = (u >> 5) * 128 + gate * 32 + (u & 31);
        const int col = k < in ? k : XC * 64 + (k - in);
        float s = 0.f;
        for (int z = 0; z < splits; ++z) s += part[((size_t)z * G4 + gp) * KW + col];
        if (k < in) gw_ih[(int64_t)g * in + k] = s * x_scale;
        else gw_hh[(int64_t)g * H + (k - in)] = s;
    }
}


// ------------------------------------------------------------------------------------------------ host side
struct W2Ws {
    uint8_t *himg, *xmask, *x0img, *gimg, *zero, *pack;
    uint32_t* bits;
    uint64_t* tag;
    float *cst, *dc, *part, *bsum;
    uint4 *dx[2], *dh_rec;
    size_t himg_bytes, gimg_bytes, bits_words, pack_stride, slot_floats, total;
    int splits;
};

int wgrad_splits(const Geom& g) {
    const int ntask = (g.H / 64) * 2;
    int s = num_sms() / ntask;
    const int64_t ktot = (int64_t)g.T * g.NT * 2;
    if (s > ktot) s = (int)ktot;
    return s < 1 ? 1 : s;
}

void w2_layout(const Geom& g, uint8_t* p, W2Ws* w) {
    size_t off = 0;
    auto take = [&](size_t bytes) { uint8_t* q = p ? p + off : nullptr; off += al256(bytes); return q; };
    w->slot_floats = (size_t)g.NT * TM * g.H;                       // one [rows x H] fp32 array in the row-slot layout
    const size_t lsz = (size_t)g.T * w->slot_floats * 4;
    w->himg_bytes = (size_t)g.T * g.NT * g.HC * WCHUNK;
    w->gimg_bytes = (size_t)g.T * g.NT * g.UG * WCHUNK;
    w->bits_words = (size_t)g.T * g.N * g.UG;
    w->pack_stride = pack_layout(g, g.L > 1 ? 1 : 0).total;
    w->splits = wgrad_splits(g);
    w->himg = take(w->himg_bytes * g.L);
    w->xmask = take(w->himg_bytes * (g.L > 1 ? g.L - 1 : 0));
    w->x0img = take((size_t)g.T * g.NT * WCHUNK);
    w->gimg = take(w->gimg_bytes * g.L);
    w->bits = (uint32_t*)take(w->bits_words * 4 * (g.L > 1 ? g.L - 1 : 0));
    w->cst = (float*)take(lsz * g.L);
    w->dx[0] = (uint4*)take(g.L > 1 ? lsz / 2 : 0);
    w->dx[1] = (uint4*)take(g.L > 2 ? lsz / 2 : 0);
    w->dh_rec = (uint4*)take(w->slot_floats * 2);
    w->dc = (float*)take(w->slot_floats * 4);
    w->part = (float*)take((size_t)w->splits * 4 * g.H * (2 * g.H) * 4);
    w->bsum = (float*)take((size_t)g.NT * g.UG * 128 * 4);
    w->zero = take(WCHUNK);
    w->tag = (uint64_t*)take(sizeof(W2Tag));
    w->pack = take(w->pack_stride * g.L);
    w->total = off;
}

unsigned ew_grid2(int64_t n) {
    const int64_t b = (n + 255) / 256, cap = (int64_t)num_sms() * 16;
    return (unsigned)(b < cap ? (b < 1 ? 1 : b) : cap);
}

struct LayerP { const float *w_ih, *w_hh, *b_ih, *b_hh; int in; };
LayerP layer_p(const float* params, int F, int H, int l) {
    int64_t off = 0;
    for (int i = 0; i < l; ++i) off += 4LL * H * (i == 0 ? F : H) + 4LL * H * H + 8LL * H;
    LayerP lp;
    lp.in = l == 0 ? F : H;
    lp.w_ih = params + off;
    lp.w_hh = lp.w_ih + 4LL * H * lp.in;
    lp.b_ih = lp.w_hh + 4LL * H * H;
    lp.b_hh = lp.b_ih + 4LL * H;
    return lp;
}

template <typename K>
int opt_in_smem(K kernel, uint32_t bytes, const char* who) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) return cuda_fail(e, who);
    return B200_OK;
}

}  // namespace

static int g_w2_dbg = 0;
void wide2_set_dbg(int v) { g_w2_dbg = v; }
static int g_w2_mc = 0;               // option "wide_mc": forward kernel as two-CTA clusters with multicast activation loads
void wide2_set_mc(int v) { g_w2_mc = v; }

bool wide2_supported(int H, int F) { return (H == 128 || H == 256) && F >= 1 && F <= 63; }

size_t wide2_workspace_bytes(int64_t N, int T, int F, int H, int L) {
    const Geom g = make_geom(N, T, F, H, L);
    W2Ws w;
    w2_layout(g, nullptr, &w);
    return w.total + 256;
}

int wide2_fwd(const float* params, int F, int H, int L, const float* x, int64_t N, int T, float* stash_h, float* stash_c, float p,
              uint64_t seed, int training, void* workspace, size_t ws_bytes, cudaStream_t st) {
    (void)stash_c;                   // the cell state lives in the workspace (row-slot layout); stash_h receives h_T of the top layer only
    const Geom g = make_geom(N, T, F, H, L);
    W2Ws w;
    w2_layout(g, (uint8_t*)workspace, &w);
    B200_CHECK_ARG(workspace != nullptr && w.total <= ws_bytes, "lstm_fwd (wide, bf16): workspace too small (%zu bytes, need %zu)", ws_bytes, w.total);
    const bool drop = training && p > 0.f && L > 1;
    const int64_t NH = N * H, lsz = (int64_t)T * NH;
    int rc = opt_in_smem(w2_fwd_kernel<false>, FwdSmem::total, "w2_fwd_kernel smem");
    if (rc) return rc;
    rc = opt_in_smem(w2_fwd_kernel<true>, FwdSmem::total, "w2_fwd_kernel smem");
    if (rc) return rc;
    // The multicast variant is used only when the workspace (the source of every multicast copy) is ordinary device memory of the
    // GPU this launch runs on; anything else (peer, mapped or managed memory) takes the plain kernel.
    bool mc = g_w2_mc != 0 && !(g_w2_dbg & 2);
    if (mc) {
        cudaPointerAttributes pa;
        int dev = -1;
        mc = cudaGetDevice(&dev) == cudaSuccess && cudaPointerGetAttributes(&pa, workspace) == cudaSuccess &&
             pa.type == cudaMemoryTypeDevice && pa.device == dev;
        cudaGetLastError();
    }
    B200_CUDA(cudaMemsetAsync(w.zero, 0, WCHUNK, st));
    w2_tag_kernel<<<1, 1, 0, st>>>(w.tag, make_tag(seed, drop, p, N, T, L, H));
    B200_LAUNCH_CHECK("w2_tag_kernel");
    { TimedLaunch _t("w2_x0_kernel", st); w2_x0_kernel<<<ew_grid2((int64_t)T * g.NT * TM * 16), 256, 0, st>>>(x, N, T, F, g.NT, w.x0img); }
    B200_LAUNCH_CHECK("w2_x0_kernel");
    for (int l = 0; l < L; ++l) {
        const LayerP lp = layer_p(params, F, H, l);
        const PackLayout pl = pack_layout(g, l);
        uint8_t* pk = w.pack + (size_t)l * w.pack_stride;
        const float in_scale = (l > 0 && drop) ? 1.0f / (1.0f - p) : 1.0f;
        { TimedLaunch _t("w2_pack_kernel", st);
          w2_pack_kernel<<<128, 256, 0, st>>>(lp.w_ih, lp.w_hh, lp.b_ih, lp.b_hh, lp.in, H, 2 * g.xc(l), in_scale, pk + pl.wf, pk + pl.wd_h,
                                              l > 0 ? pk + pl.wd_x : nullptr, (float*)(pk + pl.bias)); }
        B200_LAUNCH_CHECK("w2_pack_kernel");
        if (drop && l > 0) {
            { TimedLaunch _t("w2_bits_kernel", st); w2_bits_kernel<<<ew_grid2((int64_t)T * N * g.UG), 256, 0, st>>>(seed, l, N, T, g.UG, p, w.bits + (size_t)(l - 1) * w.bits_words); }
            B200_LAUNCH_CHECK("w2_bits_kernel");
        }
    }
    int lanes = num_sms() / g.UG;
    if (lanes > g.NT) lanes = g.NT;
    if (lanes < 1) lanes = 1;
    for (int l = 0; l < L; ++l) {
        const PackLayout pl = pack_layout(g, l);
        const uint8_t* pk = w.pack + (size_t)l * w.pack_stride;
        uint8_t* himg = w.himg + (size_t)l * w.himg_bytes;
        const bool mask_out = drop && l < L - 1;
        const uint8_t* xin = l == 0 ? w.x0img : (drop ? w.xmask + (size_t)(l - 1) * w.himg_bytes : w.himg + (size_t)(l - 1) * w.himg_bytes);
        const size_t xstep = (size_t)g.NT * g.xc(l) * WCHUNK, hstep = (size_t)g.NT * g.HC * WCHUNK, gstep = (size_t)g.NT * g.UG * WCHUNK;
        for (int t = 0; t < T; ++t) {
            FwdArgs a;
            a.wf = pk + pl.wf; a.bias = (const float*)(pk + pl.bias);
            a.xin = xin + (size_t)t * xstep;
            a.hprev = t > 0 ? himg + (size_t)(t - 1) * hstep : nullptr;
            a.hout = himg + (size_t)t * hstep;
            a.hmask = mask_out ? w.xmask + (size_t)l * w.himg_bytes + (size_t)t * hstep : nullptr;
            a.keep = mask_out ? w.bits + (size_t)l * w.bits_words + (size_t)t * N * g.UG : nullptr;
            a.c_prev = t > 0 ? w.cst + ((size_t)l * T + (t - 1)) * w.slot_floats : nullptr;
            a.c_out = w.cst + ((size_t)l * T + t) * w.slot_floats;
            a.gates = w.gimg + (size_t)l * w.gimg_bytes + (size_t)t * gstep;
            a.h_last = (l == L - 1 && t == T - 1) ? stash_h + (int64_t)l * lsz + (int64_t)t * NH : nullptr;
            a.N = N; a.NT = g.NT; a.XC = g.xc(l); a.HC = g.HC; a.UG = g.UG; a.GC = g.GC; a.H = H; a.dbg = g_w2_dbg;
            a.trace = ((g_w2_dbg & 32) && l == L - 1 && t == T / 2) ? reinterpret_cast<unsigned long long*>(stash_h) : nullptr;
            if (mc) {
                cudaLaunchConfig_t cfg = {};
                cfg.gridDim = dim3((unsigned)(lanes * g.UG)); cfg.blockDim = dim3(FWD_THREADS); cfg.dynamicSmemBytes = FwdSmem::total; cfg.stream = st;
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeClusterDimension;
                at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
                cfg.attrs = at; cfg.numAttrs = 1;
                TimedLaunch _t("w2_fwd_kernel", st);
                B200_CUDA(cudaLaunchKernelEx(&cfg, w2_fwd_kernel<true>, a));
            } else {
                TimedLaunch _t("w2_fwd_kernel", st);
                w2_fwd_kernel<false><<<(unsigned)(lanes * g.UG), FWD_THREADS, FwdSmem::total, st>>>(a);
            }
            B200_LAUNCH_CHECK("w2_fwd_kernel");
        }
    }
    return B200_OK;
}

int wide2_bwd_layer(int layer, const float* params, int F, int H, int L, const float* x, int64_t N, int T, const float* stash_h,
                    const float* stash_c, const float* dhT, float* grads, float p, uint64_t seed, int training, void* workspace,
                    size_t ws_bytes, cudaStream_t st) {
    (void)x; (void)stash_h; (void)stash_c;
    const Geom g = make_geom(N, T, F, H, L);
    W2Ws w;
    w2_layout(g, (uint8_t*)workspace, &w);
    B200_CHECK_ARG(workspace != nullptr && w.total <= ws_bytes, "lstm_bwd (wide, bf16): workspace too small (%zu bytes, need %zu)", ws_bytes, w.total);
    const bool drop = training && p > 0.f && L > 1;
    const int l = layer;
    const int64_t NH = N * H, lsz = (int64_t)T * NH;
    int rc = opt_in_smem(w2_dgrad_kernel, DgSmem::total, "w2_dgrad_kernel smem");
    if (rc) return rc;
    rc = opt_in_smem(w2_wgrad_kernel, WgSmem::total, "w2_wgrad_kernel smem");
    if (rc) return rc;
    rc = opt_in_smem(w2_cell_bwd_kernel, CellSmem::total, "w2_cell_bwd_kernel smem");
    if (rc) return rc;
    const LayerP lp = layer_p(params, F, H, l);
    const int64_t poff = lp.w_ih - params;
    float* gw_ih = grads + poff;
    float* gw_hh = gw_ih + 4LL * H * lp.in;
    float* gb_ih = gw_hh + 4LL * H * H;
    float* gb_hh = gb_ih + 4LL * H;
    const PackLayout pl = pack_layout(g, l);
    const uint8_t* pk = w.pack + (size_t)l * w.pack_stride;
    uint8_t* gimg = w.gimg + (size_t)l * w.gimg_bytes;
    const float* cs = w.cst + (size_t)l * T * w.slot_floats;
    const bool top = l == L - 1;
    const uint4* d_out = top ? nullptr : w.dx[((L - 1 - l) & 1) ^ 1];        // written by layer l + 1
    uint4* d_in = w.dx[(L - 1 - l) & 1];
    const size_t slot16 = w.slot_floats / 8;                                   // 16-byte slots of one [rows x H] bf16 array
    const size_t gstep = (size_t)g.NT * g.UG * WCHUNK;
    const float in_scale = (l > 0 && drop) ? 1.0f / (1.0f - p) : 1.0f;

    B200_CUDA(cudaMemsetAsync(w.bsum, 0, (size_t)g.NT * g.UG * 128 * 4, st));
    for (int t = T - 1; t >= 0; --t) {
        CellArgs c;
        c.G = gimg + (size_t)t * gstep;
        c.c_prev = t > 0 ? cs + (size_t)(t - 1) * w.slot_floats : nullptr;
        c.dh_rec = t < T - 1 ? w.dh_rec : nullptr;
        c.d_out = d_out ? d_out + (size_t)t * slot16 : nullptr;
        c.dhT = (top && t == T - 1) ? dhT : nullptr;
        c.dc = w.dc; c.bsum = w.bsum; c.first = t == T - 1 ? 1 : 0;
        c.N = N; c.UG = g.UG; c.GC = g.GC; c.H = H;
        c.tag = w.tag; c.want = make_tag(seed, drop, p, N, T, L, H);
        { TimedLaunch _t("w2_cell_bwd_kernel", st); w2_cell_bwd_kernel<<<(unsigned)(g.NT * g.UG), CELL_THREADS, CellSmem::total, st>>>(c); }
        B200_LAUNCH_CHECK("w2_cell_bwd_kernel");
        DgradArgs d;
        d.dG = c.G; d.wd[0] = pk + pl.wd_h; d.wd[1] = l > 0 ? pk + pl.wd_x : nullptr;
        d.out[0] = w.dh_rec; d.out[1] = l > 0 ? d_in + (size_t)t * slot16 : nullptr;
        d.keep = (drop && l > 0) ? w.bits + (size_t)(l - 1) * w.bits_words + (size_t)t * N * g.UG : nullptr;
        d.np = 0;
        if (t > 0) d.part[d.np++] = 0;
        if (l > 0) d.part[d.np++] = 1;
        if (d.np == 1) d.part[1] = d.part[0];
        d.N = N; d.NT = g.NT; d.H = H; d.UG = g.UG; d.dbg = g_w2_dbg;
        if (d.np > 0) {
            int grid = ((g.NT + 1) / 2) * d.np;
            if (grid > num_sms()) grid = num_sms();
            { TimedLaunch _t("w2_dgrad_kernel", st); w2_dgrad_kernel<<<(unsigned)grid, W2_THREADS, DgSmem::total, st>>>(d); }
            B200_LAUNCH_CHECK("w2_dgrad_kernel");
        }
    }
    WgradArgs wa;
    wa.dG = gimg;
    wa.xsrc = l == 0 ? w.x0img : (drop ? w.xmask + (size_t)(l - 1) * w.himg_bytes : w.himg + (size_t)(l - 1) * w.himg_bytes);
    wa.hsrc = w.himg + (size_t)l * w.himg_bytes;
    wa.zero = w.zero; wa.part = w.part;
    wa.T = T; wa.NT = g.NT; wa.UG = g.UG; wa.XC = g.xc(l); wa.HC = g.HC; wa.H = H;
    wa.splits = w.splits;
    const int64_t ktot = (int64_t)T * g.NT * 2;
    wa.kps = (ktot + w.splits - 1) / w.splits;
    { TimedLaunch _t("w2_wgrad_kernel", st); w2_wgrad_kernel<<<(unsigned)((H / 64) * 2 * w.splits), W2_THREADS, WgSmem::total, st>>>(wa); }
    B200_LAUNCH_CHECK("w2_wgrad_kernel");
    { TimedLaunch _t("w2_wgrad_reduce_kernel", st);
      w2_wgrad_reduce_kernel<<<ew_grid2(4LL * H * (lp.in + H)), 256, 0, st>>>(w.part, w.splits, H, 2 * g.xc(l), lp.in, in_scale, gw_ih, gw_hh); }
    B200_LAUNCH_CHECK("w2_wgrad_reduce_kernel");
    { TimedLaunch _t("w2_bsum_reduce_kernel", st); w2_bsum_reduce_kernel<<<(unsigned)((4 * H + 255) / 256), 256, 0, st>>>(w.bsum, g.NT, g.UG, H, gb_ih, gb_hh); }
    B200_LAUNCH_CHECK("w2_bsum_reduce_kernel");
    return B200_OK;
}

int wide2_hidden_states(int layer, int F, int H, int L, int64_t N, int T, const void* workspace, size_t ws_bytes, float* out, cudaStream_t st) {
    const Geom g = make_geom(N, T, F, H, L);
    W2Ws w;
    w2_layout(g, (uint8_t*)workspace, &w);
    B200_CHECK_ARG(workspace != nullptr && w.total <= ws_bytes, "lstm_hidden_states (wide, bf16): workspace too small");
    { TimedLaunch _t("w2_unpack_kernel", st); w2_unpack_kernel<<<ew_grid2((int64_t)T * N * (H / 8)), 256, 0, st>>>(w.himg + (size_t)layer * w.himg_bytes, N, T, H, g.NT, out); }
    B200_LAUNCH_CHECK("w2_unpack_kernel");
    return B200_OK;
}

}  // namespace b200
