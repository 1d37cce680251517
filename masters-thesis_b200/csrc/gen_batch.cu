// On-device batch generation: Philox4x32-10 -> Box-Muller -> Student-t(5) single-factor DGP,
// fused with windowing, the interaction feature and the per-window OLS targets.
//
// Replaces (reference file:line)
//   SyntheticLogReturns.generate          src/data.py:49-59   (distribution parameters 36-39)
//   lookback_target_split                 src/common.py:81-112
//   add_quadratic_features                src/common.py:115-130
//   ols_features / _batched_ols           src/common.py:132-149, 33-47
//   DataLoader collate + pinned H2D       src/data.py:236-238
// The generator is a pure function of (seed, stock, absolute day); the written specification
// is oracle/lstm_oracle.py ("generator spec"), which the tests compare against.
//
// One CTA per window: the (S x 90) block of returns is generated into shared memory once, then
// written out time-major ([T][N][F], coalesced 4-byte stores over stock*F) and reduced per stock
// for the OLS targets.  HBM-bound in principle (writes (T*F + n + 1) * 4 B per sequence), in
// practice Philox/Box-Muller ALU bound; either way <1% of a training step.
#include "common.cuh"

namespace b200 {

constexpr uint32_t ID_MARKET = 0xFFFFFFFFu;
constexpr uint32_t ID_COEFF = 0xFFFFFFFEu;

__device__ __forceinline__ float u01(uint32_t x) { return ((float)(x >> 8) + 0.5f) * 5.9604644775390625e-08f; }

__device__ __forceinline__ void box_muller(uint32_t xa, uint32_t xb, float& z0, float& z1) {
    float r = sqrtf(-2.0f * logf(u01(xa)));
    // cos / sin of 2 pi u through the MUFU units: the argument is moved from (0, 2 pi) to (-pi, pi), where sin.approx / cos.approx
    // are good to ~5e-7 absolute (cos(a) = -cos(a - pi), sin(a) = -sin(a - pi)); the precise sincosf was a quarter of this ALU-bound
    // kernel's instructions.  The written specification (oracle: numpy cos / sin of the same fp32 angle) is met to < 1e-6.
    float ang = 6.283185307179586f * u01(xb) - 3.14159265358979f;
    float s, c;
    __sincosf(ang, &s, &c);
    z0 = -(r * c);
    z1 = -(r * s);
}

// Student-t(5) = Z0 / sqrt((Z1^2+...+Z5^2)/5), exact for integer df; two Philox blocks per draw.
__device__ __forceinline__ float student_t5(uint64_t day, uint32_t ident, uint64_t seed) {
    uint32_t lo = (uint32_t)day, hi = (uint32_t)(day >> 32);
    U4 a = philox4x32_10(lo, hi, ident, 0u, (uint32_t)seed, (uint32_t)(seed >> 32));
    U4 b = philox4x32_10(lo, hi, ident, 1u, (uint32_t)seed, (uint32_t)(seed >> 32));
    float z0, z1, z2, z3, z4, z5;
    box_muller(a.x, a.y, z0, z1);
    box_muller(a.z, a.w, z2, z3);
    box_muller(b.x, b.y, z4, z5);
    float chi = z1 * z1 + z2 * z2 + z3 * z3 + z4 * z4 + z5 * z5;
    return z0 / sqrtf(chi * 0.2f);
}

__device__ __forceinline__ void stock_coeff(uint64_t seed, uint32_t i, float& alpha, float& beta) {
    U4 x = philox4x32_10(i, 0u, ID_COEFF, 0u, (uint32_t)seed, (uint32_t)(seed >> 32));
    float z0, z1;
    box_muller(x.x, x.y, z0, z1);
    alpha = 0.0098f + 0.1271f * z0;   // src/data.py:38
    beta = 0.9444f + 0.3521f * z1;    // src/data.py:39
}

// Source of the window's returns: the Philox generator (stocks == nullptr) or a resident series
// (stocks [S][n_samples], market [n_samples], e.g. the reference's data_dir/stocks.pt + market.pt).
// prediction != 0: window = T + n days, target = the n days after the look-back (src/common.py:104-106);
// prediction == 0: window = T days, target = the last n days of the look-back itself (src/common.py:107-110).
__global__ void __launch_bounds__(256)
gen_batch_kernel(uint64_t seed, const float* __restrict__ stocks, const float* __restrict__ market, int64_t n_samples,
                 const int64_t* __restrict__ window_idx, int64_t first_window, int64_t n_windows,
                 int S, int T, int n, int64_t stride, int F, int prediction,
                 float* __restrict__ x, float* __restrict__ y_stock, float* __restrict__ y_market,
                 float* __restrict__ factor, float* __restrict__ inv_psi,
                 float* __restrict__ ols_alpha, float* __restrict__ ols_beta) {
    extern __shared__ float sm[];
    const int W = prediction ? T + n : T;       // days per window
    const int Y0 = prediction ? T : T - n;      // first target day inside the window
    float* rs = sm;               // [S][W]
    float* rm = rs + S * W;       // [W]
    float* coef = rm + W;         // [S][2]
    float* stat = coef + 2 * S;   // [2]: market mean, centred sum of squares

    const int64_t N = n_windows * S;
    for (int64_t wl = blockIdx.x; wl < n_windows; wl += gridDim.x) {
        const int64_t w = window_idx ? window_idx[wl] : first_window + wl;
        const uint64_t day0 = (uint64_t)w * (uint64_t)stride;
        const int64_t row0 = wl * S;

        if (stocks) {
            for (int d = threadIdx.x; d < W; d += blockDim.x) rm[d] = market[day0 + d];
            for (int i = threadIdx.x; i < S * W; i += blockDim.x) {
                int s = i / W, d = i - s * W;
                rs[i] = stocks[(int64_t)s * n_samples + (int64_t)day0 + d];
            }
        } else {
            for (int d = threadIdx.x; d < W; d += blockDim.x)
                rm[d] = 0.0678f + 0.5099f * student_t5(day0 + d, ID_MARKET, seed);   // src/data.py:36
            for (int s = threadIdx.x; s < S; s += blockDim.x) stock_coeff(seed, (uint32_t)s, coef[2 * s], coef[2 * s + 1]);
            __syncthreads();
            for (int i = threadIdx.x; i < S * W; i += blockDim.x) {
                int s = i / W, d = i - s * W;
                float eps = 0.3140f * student_t5(day0 + d, (uint32_t)s, seed);          // src/data.py:37
                rs[i] = (coef[2 * s] + coef[2 * s + 1] * rm[d]) + eps;
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            float mean = 0.f;
            for (int j = 0; j < n; ++j) mean += rm[Y0 + j];
            mean /= (float)n;
            float ss = 0.f;
            for (int j = 0; j < n; ++j) { float d = rm[Y0 + j] - mean; ss += d * d; }
            stat[0] = mean;
            stat[1] = ss;
            factor[2 * wl] = mean;
            factor[2 * wl + 1] = ss / (float)(n - 1);      // torch.var: unbiased (src/common.py:141-144)
        }
        __syncthreads();

        // look-back features, time-major
        const int SF = S * F;
        for (int i = threadIdx.x; i < T * SF; i += blockDim.x) {
            int t = i / SF, r = i - t * SF;
            int s = r / F, f = r - s * F;
            float a = rs[s * W + t], m = rm[t];
            float v = f == 0 ? a : f == 1 ? m : f == 2 ? a * m : f == 3 ? a * a : m * m;
            x[((int64_t)t * N + row0) * F + r] = v;
        }
        for (int i = threadIdx.x; i < S * n; i += blockDim.x) {
            int s = i / n, j = i - s * n;
            y_stock[row0 * n + i] = rs[s * W + Y0 + j];
        }
        for (int j = threadIdx.x; j < n; j += blockDim.x) y_market[wl * n + j] = rm[Y0 + j];

        // per-stock OLS on the target window (closed form of pinv(X^T X) X^T y for [1, m])
        const float mbar = stat[0], smm = stat[1];
        for (int s = threadIdx.x; s < S; s += blockDim.x) {
            const float* r = rs + s * W + Y0;
            float rbar = 0.f;
            for (int j = 0; j < n; ++j) rbar += r[j];
            rbar /= (float)n;
            float smr = 0.f;
            for (int j = 0; j < n; ++j) smr += (rm[Y0 + j] - mbar) * (r[j] - rbar);
            float b = smr / smm, a = rbar - b * mbar;
            float ebar = 0.f;
            for (int j = 0; j < n; ++j) ebar += r[j] - (a + b * rm[Y0 + j]);
            ebar /= (float)n;
            float see = 0.f;
            for (int j = 0; j < n; ++j) { float e = r[j] - (a + b * rm[Y0 + j]) - ebar; see += e * e; }
            inv_psi[row0 + s] = (float)(n - 1) / see;      // 1 / unbiased residual variance
            if (ols_alpha) ols_alpha[row0 + s] = a;         // y channels 2, 3 when no true coefficients exist (src/data.py:205-207)
            if (ols_beta) ols_beta[row0 + s] = b;
        }
        __syncthreads();
    }
}

__global__ void gen_coeff_kernel(uint64_t seed, int S, float* alpha, float* beta) {
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < S) stock_coeff(seed, (uint32_t)s, alpha[s], beta[s]);
}

__global__ void philox_kernel(uint64_t seed, const uint32_t* __restrict__ ctr, int64_t n, uint32_t* __restrict__ out) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    U4 r = philox4x32_10(ctr[4 * i], ctr[4 * i + 1], ctr[4 * i + 2], ctr[4 * i + 3], (uint32_t)seed, (uint32_t)(seed >> 32));
    out[4 * i] = r.x; out[4 * i + 1] = r.y; out[4 * i + 2] = r.z; out[4 * i + 3] = r.w;
}

__global__ void dropout_mask_kernel(uint64_t seed, int layer, int64_t N, int T, int H, float p, float* __restrict__ out) {
    const int HB = H / 8;
    const int64_t total = (int64_t)T * N * HB;
    const uint32_t thr = dropout_threshold(p);
    const float scale = 1.0f / (1.0f - p);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int ub = (int)(i % HB);
        int64_t rt = i / HB;
        int64_t row = rt % N;
        int t = (int)(rt / N);
        U4 b = dropout_block(seed, layer, row, t, ub);
        float* o = out + ((int64_t)t * N + row) * H + ub * 8;
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = dropout_lane16(b, j) >= thr ? scale : 0.0f;
    }
}

// Reference-layout batch (device copies of what the reference's DataLoader yields, src/data.py:216, 236-238:
// X [B][S][T][F], y [B][S][n][4], factor [B][2], inv_psi [B][S]) -> the kernels' layout: x time-major [T][N][F], y_stock [N][n]
// (channel 0), y_market [B][n] (channel 1 of stock 0).  One pass; reads and writes are both coalesced over the F-wide rows
// through a shared-memory transpose of a (32 sequences x 32 steps) tile.
__global__ void __launch_bounds__(256)
collate_reference_kernel(const float* __restrict__ X, const float* __restrict__ y, int64_t B, int S, int T, int n, int F,
                         float* __restrict__ x, float* __restrict__ y_stock, float* __restrict__ y_market) {
    extern __shared__ float tile[];                       // [32 seq][32 steps * F + 1]
    const int64_t N = B * S;
    const int TF = 32 * F + 1;
    const int64_t tiles_n = (N + 31) / 32;
    const int tiles_t = (T + 31) / 32;
    for (int64_t tl = blockIdx.x; tl < tiles_n * tiles_t; tl += gridDim.x) {
        const int64_t n0 = (tl / tiles_t) * 32;
        const int t0 = (int)(tl % tiles_t) * 32;
        const int tw = T - t0 < 32 ? T - t0 : 32;
        // read: for each sequence, tw * F contiguous floats of X[(n, t0.., :)]
        for (int i = threadIdx.x; i < 32 * tw * F; i += blockDim.x) {
            const int r = i / (tw * F), c = i - r * (tw * F);
            if (n0 + r < N) tile[r * TF + c] = X[((n0 + r) * T + t0) * F + c];
        }
        __syncthreads();
        // write: for each step, 32 * F contiguous floats of x[(t, n0.., :)]
        for (int i = threadIdx.x; i < tw * 32 * F; i += blockDim.x) {
            const int tt = i / (32 * F), c = i - tt * (32 * F);
            const int r = c / F, f = c - r * F;
            if (n0 + r < N) x[((int64_t)(t0 + tt) * N + n0 + r) * F + f] = tile[r * TF + tt * F + f];
        }
        __syncthreads();
    }
    const int64_t total = N * n;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        y_stock[i] = y[i * 4];
        const int64_t row = i / n;
        if (row % S == 0) y_market[(row / S) * n + (i - row * n)] = y[i * 4 + 1];
    }
}

// the generator's series itself, for consumers that need days rather than windows (reconstruction mode, test.py)
__global__ void __launch_bounds__(256)
gen_series_kernel(uint64_t seed, int64_t first_day, int64_t n_days, int S, float* __restrict__ stocks, float* __restrict__ market) {
    const int64_t total = (int64_t)(S + 1) * n_days;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t s = i / n_days, d = i - s * n_days;
        const uint64_t day = (uint64_t)(first_day + d);
        const float m = 0.0678f + 0.5099f * student_t5(day, ID_MARKET, seed);
        if (s == S) { market[d] = m; continue; }
        float a, b;
        stock_coeff(seed, (uint32_t)s, a, b);
        stocks[s * n_days + d] = (a + b * m) + 0.3140f * student_t5(day, (uint32_t)s, seed);
    }
}

static int launch_batch(const char* name, uint64_t seed, const float* stocks, const float* market, int64_t n_samples,
                        const int64_t* window_idx, int64_t first_window, int64_t n_windows, int n_stocks, int lookback,
                        int target, int64_t stride, int n_features, int prediction, float* x, float* y_stock, float* y_market,
                        float* factor, float* inv_psi, float* ols_alpha, float* ols_beta, void* stream) {
    const int W = prediction ? lookback + target : lookback;
    size_t smem = ((size_t)n_stocks * W + W + 2 * n_stocks + 2) * sizeof(float);
    B200_CHECK_ARG(smem <= 200 * 1024, "%s: window of %d stocks x %d days does not fit shared memory", name, n_stocks, W);
    B200_CUDA(cudaFuncSetAttribute(gen_batch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t grid = n_windows < (int64_t)num_sms() * 16 ? n_windows : (int64_t)num_sms() * 16;
    {
        TimedLaunch _t(stocks ? "series_batch_kernel" : "gen_batch_kernel", (cudaStream_t)stream);
        gen_batch_kernel<<<(unsigned)grid, 256, smem, (cudaStream_t)stream>>>(seed, stocks, market, n_samples, window_idx, first_window,
                                                                         n_windows, n_stocks, lookback, target, stride, n_features,
                                                                         prediction, x, y_stock, y_market, factor, inv_psi, ols_alpha,
                                                                         ols_beta);
    }
    B200_LAUNCH_CHECK("gen_batch_kernel");
    return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_gen_batch(uint64_t seed, const int64_t* window_idx, int64_t first_window, int64_t n_windows,
                              int n_stocks, int lookback, int target, int64_t stride, int n_features,
                              float* x, float* y_stock, float* y_market, float* factor, float* inv_psi,
                              void* stream) {
    B200_CHECK_ARG(n_windows >= 0 && n_stocks > 0 && lookback > 0 && target > 1 && stride > 0,
                   "gen_batch: bad sizes (n_windows=%lld S=%d T=%d n=%d)", (long long)n_windows, n_stocks, lookback, target);
    B200_CHECK_ARG(n_features == 3 || n_features == 5, "gen_batch: n_features must be 3 or 5, got %d", n_features);
    if (n_windows == 0) return B200_OK;                 // empty batch: nothing to write (outputs may be NULL)
    B200_CHECK_ARG(x && y_stock && y_market && factor && inv_psi, "gen_batch: null output pointer");
    return launch_batch("gen_batch", seed, nullptr, nullptr, 0, window_idx, first_window, n_windows, n_stocks, lookback, target,
                        stride, n_features, 1, x, y_stock, y_market, factor, inv_psi, nullptr, nullptr, stream);
}

extern "C" int b200_series_batch(const float* stocks, const float* market, int64_t n_samples, const int64_t* window_idx,
                                 int64_t first_window, int64_t n_windows, int n_stocks, int lookback, int target,
                                 int64_t stride, int n_features, int prediction_task, float* x, float* y_stock,
                                 float* y_market, float* factor, float* inv_psi, float* ols_alpha, float* ols_beta,
                                 void* stream) {
    B200_CHECK_ARG(stocks && market && n_samples > 0, "series_batch: null series");
    B200_CHECK_ARG(n_windows >= 0 && n_stocks > 0 && lookback > 0 && target > 1 && stride > 0,
                   "series_batch: bad sizes (n_windows=%lld S=%d T=%d n=%d)", (long long)n_windows, n_stocks, lookback, target);
    B200_CHECK_ARG(prediction_task || target <= lookback, "series_batch: target window must be <= lookback window for the reconstruction task");
    B200_CHECK_ARG(n_features == 3 || n_features == 5, "series_batch: n_features must be 3 or 5, got %d", n_features);
    const int64_t W = prediction_task ? lookback + target : lookback;
    B200_CHECK_ARG(n_samples >= W, "series_batch: series of %lld days is shorter than one window (%lld)", (long long)n_samples, (long long)W);
    // contiguous ranges are checked here; explicit index lists are the caller's contract (indices < (n_samples - W) / stride + 1)
    const int64_t n_win_total = (n_samples - W) / stride + 1;
    B200_CHECK_ARG(window_idx || (first_window >= 0 && first_window + n_windows <= n_win_total),
                   "series_batch: windows [%lld, %lld) exceed the %lld windows of the series", (long long)first_window,
                   (long long)(first_window + n_windows), (long long)n_win_total);
    if (n_windows == 0) return B200_OK;                 // empty batch: nothing to write (outputs may be NULL)
    B200_CHECK_ARG(x && y_stock && y_market && factor && inv_psi, "series_batch: null output pointer");
    return launch_batch("series_batch", 0, stocks, market, n_samples, window_idx, first_window, n_windows, n_stocks, lookback, target,
                        stride, n_features, prediction_task ? 1 : 0, x, y_stock, y_market, factor, inv_psi, ols_alpha, ols_beta, stream);
}

extern "C" int b200_collate_reference(const float* X, const float* y, int64_t n_windows, int n_stocks, int lookback, int target,
                                      int n_features, float* x, float* y_stock, float* y_market, void* stream) {
    B200_CHECK_ARG(n_windows >= 0 && n_stocks > 0 && lookback > 0 && target > 0 && n_features > 0 && n_features <= 8,
                   "collate_reference: bad sizes");
    if (n_windows == 0) return B200_OK;
    B200_CHECK_ARG(X && y && x && y_stock && y_market, "collate_reference: null pointer");
    const int64_t N = n_windows * n_stocks;
    const int64_t tiles = ((N + 31) / 32) * ((lookback + 31) / 32);
    int64_t grid = tiles < (int64_t)num_sms() * 8 ? tiles : (int64_t)num_sms() * 8;
    const size_t smem = (size_t)32 * (32 * n_features + 1) * sizeof(float);
    { TimedLaunch _t("collate_reference_kernel", (cudaStream_t)stream); collate_reference_kernel<<<(unsigned)grid, 256, smem, (cudaStream_t)stream>>>(X, y, n_windows, n_stocks, lookback, target, n_features, x, y_stock, y_market); }
    B200_LAUNCH_CHECK("collate_reference_kernel");
    return B200_OK;
}

extern "C" int b200_gen_series(uint64_t seed, int64_t first_day, int64_t n_days, int n_stocks, float* stocks, float* market,
                               void* stream) {
    B200_CHECK_ARG(first_day >= 0 && n_days >= 0 && n_stocks > 0, "gen_series: bad sizes");
    B200_CHECK_ARG(stocks && market, "gen_series: null output pointer");
    if (n_days == 0) return B200_OK;
    const int64_t total = (int64_t)(n_stocks + 1) * n_days;
    int64_t grid = (total + 255) / 256;
    if (grid > (int64_t)num_sms() * 32) grid = (int64_t)num_sms() * 32;
    { TimedLaunch _t("gen_series_kernel", (cudaStream_t)stream); gen_series_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(seed, first_day, n_days, n_stocks, stocks, market); }
    B200_LAUNCH_CHECK("gen_series_kernel");
    return B200_OK;
}

extern "C" int b200_gen_coefficients(uint64_t seed, int n_stocks, float* alpha, float* beta, void* stream) {
    B200_CHECK_ARG(n_stocks > 0 && alpha && beta, "gen_coefficients: bad arguments");
    { TimedLaunch _t("gen_coeff_kernel", (cudaStream_t)stream); gen_coeff_kernel<<<(n_stocks + 127) / 128, 128, 0, (cudaStream_t)stream>>>(seed, n_stocks, alpha, beta); }
    B200_LAUNCH_CHECK("gen_coeff_kernel");
    return B200_OK;
}

extern "C" int b200_philox4x32(uint64_t seed, const uint32_t* counters, int64_t n, uint32_t* out, void* stream) {
    B200_CHECK_ARG(n >= 0 && (n == 0 || (counters && out)), "philox4x32: bad arguments");
    if (n == 0) return B200_OK;
    { TimedLaunch _t("philox_kernel", (cudaStream_t)stream); philox_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(seed, counters, n, out); }
    B200_LAUNCH_CHECK("philox_kernel");
    return B200_OK;
}

extern "C" int b200_dropout_mask(uint64_t seed, int layer, int64_t n_seq, int T, int hidden, float p, float* out,
                                 void* stream) {
    B200_CHECK_ARG(n_seq > 0 && T > 0 && hidden > 0 && hidden % 8 == 0 && out, "dropout_mask: bad sizes");
    B200_CHECK_ARG(p >= 0.f && p < 1.f, "dropout_mask: p must be in [0,1), got %f", p);
    B200_CHECK_ARG(layer >= 1 && layer < 65536 && T < 65536, "dropout_mask: layer/T out of range");
    { TimedLaunch _t("dropout_mask_kernel", (cudaStream_t)stream); dropout_mask_kernel<<<num_sms() * 8, 256, 0, (cudaStream_t)stream>>>(seed, layer, n_seq, T, hidden, p, out); }
    B200_LAUNCH_CHECK("dropout_mask_kernel");
    return B200_OK;
}
