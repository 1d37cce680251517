// alpha/beta heads + objective + their gradients, as four small bandwidth-bound kernels.
//
// Replaces (reference file:line)
//   alpha_head / beta_head (addmm)                       src/model.py:107-108
//   decoder r_pred = alpha + beta * r_m, MSE metric      src/model.py:199-206 (torchmetrics MeanSquaredError)
//   inverse_returns_covariance (dense Woodbury)          src/common.py:50-78
//   MultivariateGaussianNLLLoss (matmul, trace, logdet)  src/model.py:54-69
//   combined objective nll + mse_weight * mse            src/model.py:308-319
//   autograd backward of all of the above
//
// NLL closed form.  With p_i = 1/psi_i, d_it = r_it - alpha_i - beta_i*mu, s = sum p_i beta_i^2,
// Dn = 1/sigma^2 + s, u_t = sum_i p_i beta_i d_it, the inverse covariance Psi^-1 - (Psi^-1 b)(Psi^-1 b)^T / Dn
// is diagonal plus rank one, so
//   logdet = sum ln p_i - ln(1 + sigma^2 s)          quad = sum_t [ sum_i p_i d_it^2 - u_t^2 / Dn ]
//   L = 0.5 [ n (K ln 2pi - logdet) + quad ]
//   dL/dalpha_i = -p_i sum_t d_it + p_i beta_i sum_t u_t / Dn
//   dL/dbeta_i  = mu dL/dalpha_i - p_i sum_t u_t d_it / Dn + p_i beta_i sum_t u_t^2 / Dn^2 + n sigma^2 p_i beta_i / (1 + sigma^2 s)
// (validated against autograd through the dense reference path in tests/test_oracle.py).
//
// Kernels: K1 rows -> alpha, beta and per-window partial sums (fp64); K2 windows -> group statistics and
// per-group losses; K2b groups -> losses_out; K3 rows -> dalpha, dbeta, dhT and per-window head-gradient
// partials; K4 partials -> head gradients.  All cross-row reductions are fixed-order (deterministic).
// Algorithmic HBM bytes per sequence: K1 reads hT (4H) + targets (4n); K3 reads hT + targets again and
// writes dhT (4H): 12H + 8n + ~16 = 1,024 B at H=64, n=30.
#include "common.cuh"
#include "tc_common.cuh"

namespace b200 {

constexpr int HEADS_THREADS = 256;
constexpr int HEADS_WARPS = HEADS_THREADS / 32;
constexpr int MAX_N_TARGET = 32;   // one lane per target day
constexpr int PW_FIXED = 8;        // [0] sse [1] sae [2] s [3] sum ln p [4] sum p d^2 [5..7] unused, then u_t

struct HeadsWs {
    float* alpha;    // [N]
    float* beta;     // [N]
    double* wpart;   // [B][PW]
    float* gstat;    // [G][PW]   0 mu, 1 1/Dn, 2 sum u/Dn, 3 sum u^2/Dn^2, 4 n s2/(1+s2 s), 8.. u_t/Dn
    double* gloss;   // [G][4]    nll, sse, sae
    float* hpart;    // [B][2H+2]
};

__host__ __device__ inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

static size_t heads_ws_layout(int64_t B, int S, int n, int H, int64_t G, char* base, HeadsWs* ws) {
    const int PW = PW_FIXED + n;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += align256(bytes); return base ? base + o : (char*)nullptr; };
    char* a = take((size_t)B * S * 4);
    char* b = take((size_t)B * S * 4);
    char* w = take((size_t)B * PW * 8);
    char* g = take((size_t)G * PW * 4);
    char* l = take((size_t)G * 4 * 8);
    char* h = take((size_t)B * (2 * H + 2) * 4);
    if (ws) { ws->alpha = (float*)a; ws->beta = (float*)b; ws->wpart = (double*)w; ws->gstat = (float*)g; ws->gloss = (double*)l; ws->hpart = (float*)h; }
    return off;
}

// K1 ------------------------------------------------------------------------------------------
template <int H>
__global__ void __launch_bounds__(HEADS_THREADS)
heads_stats_kernel(const float* __restrict__ hp, const float* __restrict__ hT, const float* __restrict__ y_stock,
                   const float* __restrict__ y_market, const float* __restrict__ factor, const float* __restrict__ inv_psi,
                   int S, int n, int64_t wpg, bool do_nll, float* __restrict__ alpha_o, float* __restrict__ beta_o,
                   double* __restrict__ wpart) {
    constexpr int HL = H / 32;
    const int PW = PW_FIXED + n;
    const int64_t w = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __shared__ float red[HEADS_WARPS][PW_FIXED + MAX_N_TARGET];

    float wa[HL], wb[HL];
#pragma unroll
    for (int i = 0; i < HL; ++i) { wa[i] = hp[lane + 32 * i]; wb[i] = hp[H + 1 + lane + 32 * i]; }
    const float ba = hp[H], bb = hp[2 * H + 1];
    const float m = (y_market && lane < n) ? y_market[w * n + lane] : 0.f;
    float mu = 0.f;
    if (do_nll) mu = factor[2 * ((w / wpg) * wpg)];

    float sse = 0.f, sae = 0.f, spq = 0.f, uacc = 0.f, s_acc = 0.f, lp_acc = 0.f;
#pragma unroll 5
    for (int r = warp; r < S; r += HEADS_WARPS) {          // unrolled: several rows' loads in flight per warp
        const int64_t row = w * S + r;
        const float* h = hT + row * H;
        float pa = 0.f, pb = 0.f;
#pragma unroll
        for (int i = 0; i < HL; ++i) { float hv = h[lane + 32 * i]; pa += hv * wa[i]; pb += hv * wb[i]; }
        const float al = warp_sum(pa) + ba, be = warp_sum(pb) + bb;
        if (lane == 0) { alpha_o[row] = al; beta_o[row] = be; }
        const float rt = lane < n ? y_stock[row * n + lane] : 0.f;
        if (y_market && lane < n) { float e = al + be * m - rt; sse += e * e; sae += fabsf(e); }
        if (do_nll) {
            const float p = inv_psi[row];
            if (lane < n) { float d = rt - (al + be * mu); spq += p * d * d; uacc += p * be * d; }
            if (lane == 0) { s_acc += p * be * be; lp_acc += logf(p); }
        }
    }
    sse = warp_sum(sse); sae = warp_sum(sae); spq = warp_sum(spq);
    if (lane == 0) { red[warp][0] = sse; red[warp][1] = sae; red[warp][2] = s_acc; red[warp][3] = lp_acc; red[warp][4] = spq; }
    red[warp][PW_FIXED + lane] = uacc;
    __syncthreads();
    for (int c = threadIdx.x; c < PW; c += HEADS_THREADS) {
        double v = 0.0;
        if (c < 5 || c >= PW_FIXED)
            for (int q = 0; q < HEADS_WARPS; ++q) v += (double)red[q][c];
        wpart[w * PW + c] = v;
    }
}

// K2 ------------------------------------------------------------------------------------------
// One CTA per NLL group.  In training the whole batch is ONE group (src/model.py:237-247), so this CTA alone sums
// windows_per_group x PW doubles: 1024 threads = 16 window slices x 64 columns, four independent accumulators per thread
// (a 256-thread version with one serial chain of 1024 dependent adds per thread took as long as the whole heads_grad
// kernel, 70 us at 4096 windows).  Fixed order: the result does not depend on scheduling.
constexpr int GROUP_SLICES = 16;
__global__ void __launch_bounds__(64 * GROUP_SLICES)
heads_group_kernel(const double* __restrict__ wpart, const float* __restrict__ factor, int S, int n, int64_t wpg,
                   bool do_nll, float* __restrict__ gstat, double* __restrict__ gloss) {
    const int PW = PW_FIXED + n;
    const int64_t g = blockIdx.x;
    __shared__ double part[GROUP_SLICES][64];
    __shared__ double tot[64];
    const int c = threadIdx.x & 63, q = threadIdx.x >> 6;
    double v0 = 0.0, v1 = 0.0, v2 = 0.0, v3 = 0.0;
    if (c < PW) {
        const double* base = wpart + g * wpg * PW + c;
        int64_t i = q;
        for (; i + 3 * GROUP_SLICES < wpg; i += 4 * GROUP_SLICES) {
            v0 += base[i * PW];
            v1 += base[(i + GROUP_SLICES) * PW];
            v2 += base[(i + 2 * GROUP_SLICES) * PW];
            v3 += base[(i + 3 * GROUP_SLICES) * PW];
        }
        for (; i < wpg; i += GROUP_SLICES) v0 += base[i * PW];
    }
    part[q][c] = (v0 + v1) + (v2 + v3);
    __syncthreads();
    if (threadIdx.x < 64) {
        double t = 0.0;
#pragma unroll
        for (int k = 0; k < GROUP_SLICES; ++k) t += part[k][c];
        tot[c] = t;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double nll = 0.0;
        if (do_nll) {
            const double mu = factor[2 * g * wpg], s2 = factor[2 * g * wpg + 1];
            const double s = tot[2], lp = tot[3], pq = tot[4];
            const double Dn = 1.0 / s2 + s;
            double su = 0.0, su2 = 0.0;
            for (int t = 0; t < n; ++t) { double u = tot[PW_FIXED + t]; su += u; su2 += u * u; gstat[g * PW + PW_FIXED + t] = (float)(u / Dn); }
            const double K = (double)wpg * S;
            const double logdet = lp - log1p(s2 * s);
            const double quad = pq - su2 / Dn;
            nll = 0.5 * ((double)n * (K * 1.8378770664093453 - logdet) + quad);
            gstat[g * PW + 0] = (float)mu;
            gstat[g * PW + 1] = (float)(1.0 / Dn);
            gstat[g * PW + 2] = (float)(su / Dn);
            gstat[g * PW + 3] = (float)(su2 / (Dn * Dn));
            gstat[g * PW + 4] = (float)((double)n * s2 / (1.0 + s2 * s));
        }
        gloss[g * 4 + 0] = nll;
        gloss[g * 4 + 1] = tot[0];
        gloss[g * 4 + 2] = tot[1];
    }
}

// K2b -----------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
heads_final_kernel(const double* __restrict__ gloss, int64_t G, double n_elems, int loss_kind, float mse_weight,
                   float* __restrict__ losses) {
    __shared__ double red[3][256];
    double a = 0, b = 0, c = 0;
    for (int64_t g = threadIdx.x; g < G; g += 256) { a += gloss[g * 4]; b += gloss[g * 4 + 1]; c += gloss[g * 4 + 2]; }
    red[0][threadIdx.x] = a; red[1][threadIdx.x] = b; red[2][threadIdx.x] = c;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) for (int k = 0; k < 3; ++k) red[k][threadIdx.x] += red[k][threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double nll = red[0][0] / (double)G, mse = red[1][0] / n_elems, mae = red[2][0] / n_elems;
        losses[0] = (float)mse;
        losses[1] = (float)nll;
        losses[2] = (float)(loss_kind == B200_LOSS_MSE ? mse : loss_kind == B200_LOSS_NLL ? nll : nll + (double)mse_weight * mse);
        losses[3] = (float)mae;
    }
}

// K3 ------------------------------------------------------------------------------------------
template <int H>
__global__ void __launch_bounds__(HEADS_THREADS)
heads_grad_kernel(const float* __restrict__ hp, const float* __restrict__ hT, const float* __restrict__ y_stock,
                  const float* __restrict__ y_market, const float* __restrict__ inv_psi,
                  const float* __restrict__ alpha_i, const float* __restrict__ beta_i, const float* __restrict__ gstat,
                  int S, int n, int64_t wpg, float c_mse, float c_nll, float* __restrict__ dhT, float* __restrict__ hpart) {
    constexpr int HL = H / 32;
    const int PW = PW_FIXED + n;
    const int64_t w = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __shared__ float red[HEADS_WARPS][2 * H + 2];

    float wa[HL], wb[HL], ga[HL], gb[HL];
#pragma unroll
    for (int i = 0; i < HL; ++i) { wa[i] = hp[lane + 32 * i]; wb[i] = hp[H + 1 + lane + 32 * i]; ga[i] = 0.f; gb[i] = 0.f; }
    float gba = 0.f, gbb = 0.f;
    const bool do_mse = c_mse != 0.f, do_nll = c_nll != 0.f;
    const float m = (do_mse && lane < n) ? y_market[w * n + lane] : 0.f;
    const float* gs = gstat + (w / wpg) * PW;
    float mu = 0.f, rDn = 0.f, suD = 0.f, su2D = 0.f, c4 = 0.f, uD = 0.f;
    if (do_nll) { mu = gs[0]; rDn = gs[1]; suD = gs[2]; su2D = gs[3]; c4 = gs[4]; uD = lane < n ? gs[PW_FIXED + lane] : 0.f; }
    (void)rDn;

#pragma unroll 5
    for (int r = warp; r < S; r += HEADS_WARPS) {          // unrolled: several rows' loads in flight per warp
        const int64_t row = w * S + r;
        const float al = alpha_i[row], be = beta_i[row];
        const float rt = lane < n ? y_stock[row * n + lane] : 0.f;
        float da = 0.f, db = 0.f;
        if (do_mse) {
            float e = lane < n ? al + be * m - rt : 0.f;
            float se = warp_sum(e), sem = warp_sum(e * m);
            da += c_mse * se;
            db += c_mse * sem;
        }
        if (do_nll) {
            const float p = inv_psi[row];
            float d = lane < n ? rt - (al + be * mu) : 0.f;
            float sd = warp_sum(d), sud = warp_sum(uD * d);
            float dan = -p * sd + p * be * suD;
            float dbn = mu * dan - p * sud + p * be * su2D + c4 * p * be;
            da += c_nll * dan;
            db += c_nll * dbn;
        }
        const float* h = hT + row * H;
        float* o = dhT + row * H;
#pragma unroll
        for (int i = 0; i < HL; ++i) {
            float hv = h[lane + 32 * i];
            ga[i] += da * hv;
            gb[i] += db * hv;
            o[lane + 32 * i] = da * wa[i] + db * wb[i];
        }
        gba += da;
        gbb += db;
    }
#pragma unroll
    for (int i = 0; i < HL; ++i) { red[warp][lane + 32 * i] = ga[i]; red[warp][H + 1 + lane + 32 * i] = gb[i]; }
    if (lane == 0) { red[warp][H] = gba; red[warp][2 * H + 1] = gbb; }
    __syncthreads();
    for (int c = threadIdx.x; c < 2 * H + 2; c += HEADS_THREADS) {
        float v = 0.f;
        for (int q = 0; q < HEADS_WARPS; ++q) v += red[q][c];
        hpart[w * (2 * H + 2) + c] = v;
    }
}

// ---- K1 / K3 for hidden sizes that are multiples of 64: HALF a warp per row ---------------------------------------------------
// A row of hT is H floats: 16 lanes x float4 x (H / 64).  Two rows per warp iteration, 128-bit loads of hT, 64-bit loads of the
// target row (lane l of a half handles days 2l, 2l + 1), 4-step reductions inside the half: half the instructions per byte of the
// one-row-per-warp kernels above (ncu: those ran at 31 / 41 % of DRAM peak with 39 / 46 % of the issue slots busy).
__device__ __forceinline__ float half_sum(float v) {      // sum over the 16 lanes of a half warp
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float2 ld_pair(const float* p, int i, int n) {      // elements i, i + 1 of a row of n floats (0 beyond n)
    float2 r = make_float2(0.f, 0.f);
    if (i + 1 < n && ((n & 1) == 0)) r = *reinterpret_cast<const float2*>(p + i);
    else { if (i < n) r.x = p[i]; if (i + 1 < n) r.y = p[i + 1]; }
    return r;
}

// A window's rows of hT (S x H floats) and of y_stock (S x n floats) are contiguous in global memory: one cp.async.bulk each
// (completion by mbarrier complete_tx) puts the CTA's whole input in flight at once.  The loops below also support fewer CTAs
// than windows with a second stage (next window fetched during the current one); heads_launch uses one window per CTA.
__device__ __forceinline__ void heads_fetch(float* stage, const float* hT, const float* y_stock, int64_t w, int S, int H, int n, uint64_t* bar) {
    const uint32_t hb = (uint32_t)S * H * 4, yb = (uint32_t)S * n * 4;
    tc::mbar_arrive_expect_tx(bar, hb + yb);
    tc::bulk_copy_g2s(stage, hT + w * S * H, hb, bar);
    tc::bulk_copy_g2s(stage + (size_t)S * H, y_stock + w * S * n, yb, bar);
}

template <int H>
__global__ void __launch_bounds__(HEADS_THREADS)
heads_stats2_kernel(const float* __restrict__ hp, const float* __restrict__ hT, const float* __restrict__ y_stock,
                    const float* __restrict__ y_market, const float* __restrict__ factor, const float* __restrict__ inv_psi,
                    int64_t B, int S, int n, int64_t wpg, bool do_nll, float* __restrict__ alpha_o, float* __restrict__ beta_o,
                    double* __restrict__ wpart) {
    constexpr int HQ = H / 64;
    const int PW = PW_FIXED + n;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, half = lane >> 4, l16 = lane & 15;
    __shared__ float red[HEADS_WARPS][PW_FIXED + MAX_N_TARGET];
    __shared__ __align__(8) uint64_t bar[2];
    extern __shared__ __align__(128) float win[];
    const size_t stage_floats = (size_t)S * (H + n);
    if (threadIdx.x == 0) { tc::mbar_init(&bar[0], 1); tc::mbar_init(&bar[1], 1); tc::fence_barrier_init(); }
    __syncthreads();
    if (threadIdx.x == 0 && blockIdx.x < B) heads_fetch(win, hT, y_stock, blockIdx.x, S, H, n, &bar[0]);
    float4 wa[HQ], wb[HQ];
#pragma unroll
    for (int i = 0; i < HQ; ++i) {
        wa[i] = *reinterpret_cast<const float4*>(hp + 64 * i + 4 * l16);
        const float* b = hp + H + 1 + 64 * i + 4 * l16;            // beta head starts at an odd offset: scalar loads
        wb[i] = make_float4(b[0], b[1], b[2], b[3]);
    }
    const float ba = hp[H], bb = hp[2 * H + 1];
    const int d0 = 2 * l16;
    int it = 0;
    for (int64_t w = blockIdx.x; w < B; w += gridDim.x, ++it) {
        const int st = it & 1;
        // every thread left the previous iteration (trailing __syncthreads): the other stage is free for the next window
        if (threadIdx.x == 0 && w + gridDim.x < B) heads_fetch(win + (st ^ 1) * stage_floats, hT, y_stock, w + gridDim.x, S, H, n, &bar[st ^ 1]);
        const float* hs = win + st * stage_floats;
        const float* ys = hs + (size_t)S * H;
        const float2 m = y_market ? ld_pair(y_market + w * n, d0, n) : make_float2(0.f, 0.f);
        float mu = 0.f;
        if (do_nll) mu = factor[2 * ((w / wpg) * wpg)];
        float sse = 0.f, sae = 0.f, spq = 0.f, u0 = 0.f, u1 = 0.f, s_acc = 0.f, lp_acc = 0.f;
        tc::mbar_wait(&bar[st], (uint32_t)((it >> 1) & 1));
#pragma unroll 2
        for (int r0 = 2 * warp; r0 < S; r0 += 2 * HEADS_WARPS) {
            const int r = r0 + half;
            const bool ok = r < S;
            const int rl = ok ? r : 0;
            const int64_t row = w * S + rl;
            float pa = 0.f, pb = 0.f;
#pragma unroll
            for (int i = 0; i < HQ; ++i) {
                const float4 h = ok ? *reinterpret_cast<const float4*>(hs + rl * H + 64 * i + 4 * l16) : make_float4(0.f, 0.f, 0.f, 0.f);
                pa += h.x * wa[i].x + h.y * wa[i].y + h.z * wa[i].z + h.w * wa[i].w;
                pb += h.x * wb[i].x + h.y * wb[i].y + h.z * wb[i].z + h.w * wb[i].w;
            }
            const float2 y = ok ? ld_pair(ys + rl * n, d0, n) : make_float2(0.f, 0.f);
            const float al = half_sum(pa) + ba, be = half_sum(pb) + bb;
            if (ok && l16 == 0) { alpha_o[row] = al; beta_o[row] = be; }
            const bool v0 = ok && d0 < n, v1 = ok && d0 + 1 < n;
            if (y_market) {
                const float e0 = v0 ? al + be * m.x - y.x : 0.f, e1 = v1 ? al + be * m.y - y.y : 0.f;
                sse += e0 * e0 + e1 * e1;
                sae += fabsf(e0) + fabsf(e1);
            }
            if (do_nll && ok) {
                const float p = inv_psi[row];
                const float q0 = v0 ? y.x - (al + be * mu) : 0.f, q1 = v1 ? y.y - (al + be * mu) : 0.f;
                spq += p * (q0 * q0 + q1 * q1);
                u0 += p * be * q0;
                u1 += p * be * q1;
                if (l16 == 0) { s_acc += p * be * be; lp_acc += logf(p); }
            }
        }
        sse = warp_sum(sse); sae = warp_sum(sae); spq = warp_sum(spq); s_acc = warp_sum(s_acc); lp_acc = warp_sum(lp_acc);
        u0 += __shfl_xor_sync(0xffffffffu, u0, 16);
        u1 += __shfl_xor_sync(0xffffffffu, u1, 16);
        if (lane == 0) { red[warp][0] = sse; red[warp][1] = sae; red[warp][2] = s_acc; red[warp][3] = lp_acc; red[warp][4] = spq; }
        if (half == 0) { red[warp][PW_FIXED + d0] = u0; red[warp][PW_FIXED + d0 + 1] = u1; }
        __syncthreads();
        for (int c = threadIdx.x; c < PW; c += HEADS_THREADS) {
            double v = 0.0;
            if (c < 5 || c >= PW_FIXED)
                for (int q = 0; q < HEADS_WARPS; ++q) v += (double)red[q][c];
            wpart[w * PW + c] = v;
        }
        __syncthreads();
    }
}

template <int H>
__global__ void __launch_bounds__(HEADS_THREADS)
heads_grad2_kernel(const float* __restrict__ hp, const float* __restrict__ hT, const float* __restrict__ y_stock,
                   const float* __restrict__ y_market, const float* __restrict__ inv_psi,
                   const float* __restrict__ alpha_i, const float* __restrict__ beta_i, const float* __restrict__ gstat,
                   int64_t B, int S, int n, int64_t wpg, float c_mse, float c_nll, float* __restrict__ dhT, float* __restrict__ hpart) {
    constexpr int HQ = H / 64;
    const int PW = PW_FIXED + n;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, half = lane >> 4, l16 = lane & 15;
    __shared__ float red[HEADS_WARPS][2 * H + 2];
    __shared__ __align__(8) uint64_t bar[2];
    // dL/dhT is written IN PLACE over the hT rows of the stage and leaves with one bulk copy shared -> global; the stage is only
    // refilled after the copy engine has read it (wait_group.read by the issuing thread)
    extern __shared__ __align__(128) float win[];
    const size_t stage_floats = (size_t)S * (H + n);
    if (threadIdx.x == 0) { tc::mbar_init(&bar[0], 1); tc::mbar_init(&bar[1], 1); tc::fence_barrier_init(); }
    __syncthreads();
    if (threadIdx.x == 0 && blockIdx.x < B) heads_fetch(win, hT, y_stock, blockIdx.x, S, H, n, &bar[0]);
    float4 wa[HQ], wb[HQ];
#pragma unroll
    for (int i = 0; i < HQ; ++i) {
        wa[i] = *reinterpret_cast<const float4*>(hp + 64 * i + 4 * l16);
        const float* b = hp + H + 1 + 64 * i + 4 * l16;
        wb[i] = make_float4(b[0], b[1], b[2], b[3]);
    }
    const bool do_mse = c_mse != 0.f, do_nll = c_nll != 0.f;
    const int d0 = 2 * l16;
    int it = 0;
    for (int64_t w = blockIdx.x; w < B; w += gridDim.x, ++it) {
        const int st = it & 1;
        if (threadIdx.x == 0 && w + gridDim.x < B) {
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");       // the other stage's outgoing copy has read its rows
            heads_fetch(win + (st ^ 1) * stage_floats, hT, y_stock, w + gridDim.x, S, H, n, &bar[st ^ 1]);
        }
        float* hs = win + st * stage_floats;
        const float* ys = hs + (size_t)S * H;
        float4 ga[HQ], gb[HQ];
#pragma unroll
        for (int i = 0; i < HQ; ++i) ga[i] = gb[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        float gba = 0.f, gbb = 0.f;
        const float2 m = do_mse ? ld_pair(y_market + w * n, d0, n) : make_float2(0.f, 0.f);
        const float* gs = gstat + (w / wpg) * PW;
        float mu = 0.f, suD = 0.f, su2D = 0.f, c4 = 0.f;
        float2 uD = make_float2(0.f, 0.f);
        if (do_nll) { mu = gs[0]; suD = gs[2]; su2D = gs[3]; c4 = gs[4]; uD = ld_pair(gs + PW_FIXED, d0, n); }
        tc::mbar_wait(&bar[st], (uint32_t)((it >> 1) & 1));
#pragma unroll 2
        for (int r0 = 2 * warp; r0 < S; r0 += 2 * HEADS_WARPS) {
            const int r = r0 + half;
            const bool ok = r < S;
            const int rl = ok ? r : 0;
            const int64_t row = w * S + rl;
            const float al = alpha_i[row], be = beta_i[row];
            const float2 y = ld_pair(ys + rl * n, d0, n);
            const bool v0 = d0 < n, v1 = d0 + 1 < n;
            float da = 0.f, db = 0.f;
            if (do_mse) {
                const float e0 = v0 ? al + be * m.x - y.x : 0.f, e1 = v1 ? al + be * m.y - y.y : 0.f;
                const float se = half_sum(e0 + e1), sem = half_sum(e0 * m.x + e1 * m.y);
                da += c_mse * se;
                db += c_mse * sem;
            }
            if (do_nll) {
                const float p = inv_psi[row];
                const float q0 = v0 ? y.x - (al + be * mu) : 0.f, q1 = v1 ? y.y - (al + be * mu) : 0.f;
                const float sd = half_sum(q0 + q1), sud = half_sum(uD.x * q0 + uD.y * q1);
                const float dan = -p * sd + p * be * suD;
                const float dbn = mu * dan - p * sud + p * be * su2D + c4 * p * be;
                da += c_nll * dan;
                db += c_nll * dbn;
            }
            if (!ok) { da = 0.f; db = 0.f; }
#pragma unroll
            for (int i = 0; i < HQ; ++i) {
                float4* hp4 = reinterpret_cast<float4*>(hs + rl * H + 64 * i + 4 * l16);
                const float4 h = *hp4;
                ga[i].x += da * h.x; ga[i].y += da * h.y; ga[i].z += da * h.z; ga[i].w += da * h.w;
                gb[i].x += db * h.x; gb[i].y += db * h.y; gb[i].z += db * h.z; gb[i].w += db * h.w;
                if (ok)
                    *hp4 = make_float4(da * wa[i].x + db * wb[i].x, da * wa[i].y + db * wb[i].y, da * wa[i].z + db * wb[i].z, da * wa[i].w + db * wb[i].w);
            }
            gba += da;
            gbb += db;
        }
        // the two halves of a warp hold partial sums of the same columns
#pragma unroll
        for (int i = 0; i < HQ; ++i) {
            ga[i].x += __shfl_xor_sync(0xffffffffu, ga[i].x, 16); ga[i].y += __shfl_xor_sync(0xffffffffu, ga[i].y, 16);
            ga[i].z += __shfl_xor_sync(0xffffffffu, ga[i].z, 16); ga[i].w += __shfl_xor_sync(0xffffffffu, ga[i].w, 16);
            gb[i].x += __shfl_xor_sync(0xffffffffu, gb[i].x, 16); gb[i].y += __shfl_xor_sync(0xffffffffu, gb[i].y, 16);
            gb[i].z += __shfl_xor_sync(0xffffffffu, gb[i].z, 16); gb[i].w += __shfl_xor_sync(0xffffffffu, gb[i].w, 16);
        }
        gba += __shfl_xor_sync(0xffffffffu, gba, 16);
        gbb += __shfl_xor_sync(0xffffffffu, gbb, 16);
        tc::fence_proxy_async();            // the in-place dL/dhT rows (generic stores) -> visible to the bulk copy engine
        if (half == 0) {
#pragma unroll
            for (int i = 0; i < HQ; ++i) {
                float* a = &red[warp][64 * i + 4 * l16];
                float* b = &red[warp][H + 1 + 64 * i + 4 * l16];
                a[0] = ga[i].x; a[1] = ga[i].y; a[2] = ga[i].z; a[3] = ga[i].w;
                b[0] = gb[i].x; b[1] = gb[i].y; b[2] = gb[i].z; b[3] = gb[i].w;
            }
            if (l16 == 0) { red[warp][H] = gba; red[warp][2 * H + 1] = gbb; }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            tc::bulk_copy_s2g(dhT + w * S * H, hs, (uint32_t)S * H * 4);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        for (int c = threadIdx.x; c < 2 * H + 2; c += HEADS_THREADS) {
            float v = 0.f;
            for (int q = 0; q < HEADS_WARPS; ++q) v += red[q][c];
            hpart[w * (2 * H + 2) + c] = v;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");    // shared memory stays valid until it is read
}

// K4 ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
heads_reduce_kernel(const float* __restrict__ hpart, int64_t B, int NC, float* __restrict__ out) {
    __shared__ double red[256];
    const int c = blockIdx.x;
    double v = 0.0;
    for (int64_t w = threadIdx.x; w < B; w += 256) v += (double)hpart[w * NC + c];
    red[threadIdx.x] = v;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[c] = (float)red[0];
}

template <int H>
static int heads_launch(int loss_kind, float mse_weight, const float* hp, const float* hT, const float* y_stock,
                        const float* y_market, const float* factor, const float* inv_psi, int64_t B, int S, int n,
                        int64_t wpg, float grad_scale, float* alpha_out, float* beta_out, float* losses_out, float* dhT,
                        float* head_grads, const HeadsWs& ws, cudaStream_t st) {
    const bool do_nll = loss_kind != B200_LOSS_MSE;
    const int64_t G = B / wpg;
    float* al = alpha_out ? alpha_out : ws.alpha;
    float* be = beta_out ? beta_out : ws.beta;
    // bulk-copy kernels: hidden size a multiple of 64, a window's target rows a multiple of 16 bytes, everything 16-byte aligned,
    // and the window (S x (H + n) floats) fits shared memory
    // One window per CTA, one stage (grid = B): measured at 4096 windows 44 us (stats) / 67 us (grad) with 256 threads; 128 threads
    // 53 / 78 us, 512 threads 64 / 85 us.  The kernels also run persistently with two stages (grid < B: the next window is fetched
    // while the current one is processed), but with the 2 CTAs per SM that two stages leave room for that measured 60 / 80 us:
    // the per-window arithmetic (half-warp reductions), not the memory latency, is what a CTA spends its time on.
    const size_t win_bytes = (size_t)S * (H + n) * 4;
    int pgrid = 1;
    bool bulk = false;
    if constexpr (H % 64 == 0) {
        bulk = ((size_t)S * n * 4) % 16 == 0 && win_bytes <= 200 * 1024 &&
               ((((uintptr_t)hp | (uintptr_t)hT | (uintptr_t)y_stock | (uintptr_t)(dhT ? dhT : hT)) & 15) == 0);
        if (bulk) {
            B200_CUDA(cudaFuncSetAttribute(heads_stats2_kernel<H>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)win_bytes));
            B200_CUDA(cudaFuncSetAttribute(heads_grad2_kernel<H>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)win_bytes));
            pgrid = (int)B;
        }
    }
    {
        TimedLaunch _t("heads_stats_kernel", st);
        bool done = false;
        if constexpr (H % 64 == 0) {
            if (bulk) { heads_stats2_kernel<H><<<pgrid, HEADS_THREADS, win_bytes, st>>>(hp, hT, y_stock, y_market, factor, inv_psi, B, S, n, wpg, do_nll, al, be, ws.wpart); done = true; }
        }
        if (!done) heads_stats_kernel<H><<<(unsigned)B, HEADS_THREADS, 0, st>>>(hp, hT, y_stock, y_market, factor, inv_psi, S, n, wpg, do_nll, al, be, ws.wpart);
    }
    B200_LAUNCH_CHECK("heads_stats_kernel");
    { TimedLaunch _t("heads_group_kernel", st); heads_group_kernel<<<(unsigned)G, 64 * GROUP_SLICES, 0, st>>>(ws.wpart, factor, S, n, wpg, do_nll, ws.gstat, ws.gloss); }
    B200_LAUNCH_CHECK("heads_group_kernel");
    const double n_elems = (double)B * S * n;
    { TimedLaunch _t("heads_final_kernel", st); heads_final_kernel<<<1, 256, 0, st>>>(ws.gloss, G, n_elems, loss_kind, mse_weight, losses_out); }
    B200_LAUNCH_CHECK("heads_final_kernel");
    if (dhT) {
        const float w_mse = loss_kind == B200_LOSS_MSE ? 1.f : loss_kind == B200_LOSS_NLL ? 0.f : mse_weight;
        const float c_mse = (float)((double)grad_scale * w_mse * 2.0 / n_elems);
        const float c_nll = do_nll ? (float)((double)grad_scale / (double)G) : 0.f;
        {
            TimedLaunch _t("heads_grad_kernel", st);
            bool done = false;
            if constexpr (H % 64 == 0) {
                if (bulk) { heads_grad2_kernel<H><<<pgrid, HEADS_THREADS, win_bytes, st>>>(hp, hT, y_stock, y_market, inv_psi, al, be, ws.gstat, B, S, n, wpg, c_mse, c_nll, dhT, ws.hpart); done = true; }
            }
            if (!done) heads_grad_kernel<H><<<(unsigned)B, HEADS_THREADS, 0, st>>>(hp, hT, y_stock, y_market, inv_psi, al, be, ws.gstat, S, n, wpg, c_mse, c_nll, dhT, ws.hpart);
        }
        B200_LAUNCH_CHECK("heads_grad_kernel");
        if (head_grads) {
            { TimedLaunch _t("heads_reduce_kernel", st); heads_reduce_kernel<<<2 * H + 2, 256, 0, st>>>(ws.hpart, B, 2 * H + 2, head_grads); }
            B200_LAUNCH_CHECK("heads_reduce_kernel");
        }
    }
    return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" size_t b200_heads_workspace_bytes(int64_t n_windows, int n_stocks, int n_target, int hidden) {
    if (n_windows <= 0 || n_stocks <= 0 || n_target <= 0 || hidden <= 0) return 0;
    // sized for the worst case windows_per_group = 1 (one group per window)
    return heads_ws_layout(n_windows, n_stocks, n_target, hidden, n_windows, nullptr, nullptr);
}

extern "C" int b200_heads_loss_fwd_bwd(int loss_kind, float mse_weight, const float* head_params, int hidden,
                                       const float* hT, const float* y_stock, const float* y_market, const float* factor,
                                       const float* inv_psi, int64_t n_windows, int n_stocks, int n_target,
                                       int64_t windows_per_group, float grad_scale, float* alpha_out, float* beta_out,
                                       float* losses_out, float* dhT, float* head_grads, void* workspace,
                                       size_t workspace_bytes, void* stream) {
    B200_CHECK_ARG(loss_kind >= 0 && loss_kind <= 2, "heads: unknown loss kind %d", loss_kind);
    B200_CHECK_ARG(n_windows > 0 && n_stocks > 0, "heads: empty batch (n_windows=%lld, n_stocks=%d)", (long long)n_windows, n_stocks);
    B200_CHECK_ARG(n_target >= 2 && n_target <= MAX_N_TARGET, "heads: target window must be 2..%d days, got %d", MAX_N_TARGET, n_target);
    B200_CHECK_ARG(head_params && hT && y_stock && losses_out && workspace, "heads: null pointer");
    B200_CHECK_ARG(windows_per_group > 0 && n_windows % windows_per_group == 0,
                   "heads: n_windows=%lld not divisible by windows_per_group=%lld", (long long)n_windows, (long long)windows_per_group);
    if (loss_kind != B200_LOSS_MSE) B200_CHECK_ARG(factor && inv_psi, "heads: NLL needs factor and inv_psi");
    if (loss_kind != B200_LOSS_NLL) B200_CHECK_ARG(y_market, "heads: MSE needs y_market");
    B200_CHECK_ARG(!head_grads || dhT, "heads: head_grads requires dhT");
    HeadsWs ws;
    size_t need = heads_ws_layout(n_windows, n_stocks, n_target, hidden, n_windows / windows_per_group, (char*)workspace, &ws);
    B200_CHECK_ARG(workspace_bytes >= need, "heads: workspace too small (%zu < %zu)", workspace_bytes, need);
    cudaStream_t st = (cudaStream_t)stream;
    switch (hidden) {
        case 32: return heads_launch<32>(loss_kind, mse_weight, head_params, hT, y_stock, y_market, factor, inv_psi, n_windows, n_stocks, n_target, windows_per_group, grad_scale, alpha_out, beta_out, losses_out, dhT, head_grads, ws, st);
        case 64: return heads_launch<64>(loss_kind, mse_weight, head_params, hT, y_stock, y_market, factor, inv_psi, n_windows, n_stocks, n_target, windows_per_group, grad_scale, alpha_out, beta_out, losses_out, dhT, head_grads, ws, st);
        case 128: return heads_launch<128>(loss_kind, mse_weight, head_params, hT, y_stock, y_market, factor, inv_psi, n_windows, n_stocks, n_target, windows_per_group, grad_scale, alpha_out, beta_out, losses_out, dhT, head_grads, ws, st);
        case 256: return heads_launch<256>(loss_kind, mse_weight, head_params, hT, y_stock, y_market, factor, inv_psi, n_windows, n_stocks, n_target, windows_per_group, grad_scale, alpha_out, beta_out, losses_out, dhT, head_grads, ws, st);
        default: set_error("heads: hidden size %d not supported (32, 64, 128, 256)", hidden); return B200_ERR_UNSUPPORTED;
    }
}

extern "C" int b200_heads_mse_fwd_bwd(const float* head_params, int hidden, const float* hT, const float* y_stock,
                                      const float* y_market, int64_t n_windows, int n_stocks, int n_target,
                                      float grad_scale, float* losses_out, float* dhT, float* head_grads, void* workspace,
                                      size_t workspace_bytes, void* stream) {
    return b200_heads_loss_fwd_bwd(B200_LOSS_MSE, 0.f, head_params, hidden, hT, y_stock, y_market, nullptr, nullptr, n_windows,
                                   n_stocks, n_target, n_windows > 0 ? n_windows : 1, grad_scale, nullptr, nullptr, losses_out,
                                   dhT, head_grads, workspace, workspace_bytes, stream);
}

extern "C" int b200_heads_nll_fwd_bwd(const float* head_params, int hidden, const float* hT, const float* y_stock,
                                      const float* factor, const float* inv_psi, int64_t n_windows, int n_stocks,
                                      int n_target, int64_t windows_per_group, float grad_scale, float* losses_out,
                                      float* dhT, float* head_grads, void* workspace, size_t workspace_bytes, void* stream) {
    return b200_heads_loss_fwd_bwd(B200_LOSS_NLL, 0.f, head_params, hidden, hT, y_stock, nullptr, factor, inv_psi, n_windows,
                                   n_stocks, n_target, windows_per_group, grad_scale, nullptr, nullptr, losses_out, dhT,
                                   head_grads, workspace, workspace_bytes, stream);
}
