// fp32 CUDA-core LSTM engine (B200_MODE_FFMA): forward, BPTT and weight gradients of one layer.
//
// Replaces (reference file:line)
//   self.lstm(lstm_input): nn.LSTM forward -> aten::mkldnn_rnn_layer / _cudnn_rnn     src/model.py:104 (module 88-94)
//   its autograd backward (mkldnn_rnn_layer_backward): dgrad, wgrad, bias grad        Lightning loss.backward()
//   inter-layer dropout (aten::bernoulli_ inside nn.LSTM)                              src/model.py:92
//
// This engine exists to be exactly fp32 (every product and sum in FFMA, accurate expf/tanhf): it is
// the on-device reference the tcgen05 engines (lstm_tc.cu) are validated against at sizes the CPU
// oracle cannot reach, and the engine the strict parity gate (loss 1e-4, grads 1e-3) is pinned on.
// Roofline: CUDA-core FFMA (not the headline tensor-pipe bound), see DESIGN.md.
//
// Design.  A warp owns 8 sequences for all T steps; a CTA is 8 independent warps that share one
// shared-memory copy of the layer's transposed weights Wt[k][j] (k over [x | h], j over the 4H gate
// columns, row stride 257 floats so that both the "row k, 32 consecutive j" reads of the gate GEMM
// and the "32 consecutive k, column j" reads of the data-gradient GEMM are bank-conflict free).
// Lane l of a warp owns hidden units l and l+32, i.e. gate columns c*32+l for c = gate*2 + half, so
// the LSTM cell update, the state stash ([T][N][H], 128-byte coalesced per half row) and the recurrent
// dh all stay in the lane's registers; the warp's rows of the operand tile U are private to the warp,
// so the time loop needs only __syncwarp().
// Backward is two kernels per layer: `bwd_recurrent` (gate recompute from the stashed h, c; da;
// dx/dh_{t-1} = da * W) which streams da to a workspace, and `wgrad` (dW = da^T [x | h_{t-1}] as a
// split-K SGEMM with per-CTA partials reduced in fixed order, so gradients are run-to-run identical).
#include "common.cuh"

namespace b200 {

constexpr int LH = 64;            // hidden size of this engine
constexpr int LG = 4 * LH;        // gate columns
constexpr int WT_STRIDE = LG + 1; // 257
constexpr int ROWS_PER_WARP = 8;
constexpr int L_WARPS = 8;
constexpr int L_THREADS = 32 * L_WARPS;
constexpr int L_ROWS = ROWS_PER_WARP * L_WARPS;   // 64 sequences per CTA tile
constexpr int WG_ROWS = 32;       // rows per K-block of the wgrad kernel
constexpr int WG_LD = 132;        // smem row stride of the wgrad tiles

struct LayerGeom {
    int layer, kin, kp, ktot;     // input width, padded input width, kp + H
    const float *w_ih, *w_hh, *b_ih, *b_hh;
};

__host__ __device__ inline int pad4(int v) { return (v + 3) & ~3; }

// Wt[k][j]: k < kin -> W_ih[j][k]; kin <= k < kp -> 0; k >= kp -> W_hh[j][k-kp].  bsum[j] = b_ih + b_hh.
__device__ void load_weights(float* Wt, float* bsum, const LayerGeom& g) {
    for (int i = threadIdx.x; i < LG * g.kin; i += blockDim.x) {
        int j = i / g.kin, k = i - j * g.kin;
        Wt[k * WT_STRIDE + j] = g.w_ih[i];
    }
    for (int i = threadIdx.x; i < LG * (g.kp - g.kin); i += blockDim.x) {
        int j = i / (g.kp - g.kin), k = g.kin + i - j * (g.kp - g.kin);
        Wt[k * WT_STRIDE + j] = 0.f;
    }
    for (int i = threadIdx.x; i < LG * LH; i += blockDim.x) {
        int j = i / LH, k = i - j * LH;
        Wt[(g.kp + k) * WT_STRIDE + j] = g.w_hh[i];
    }
    for (int j = threadIdx.x; j < LG; j += blockDim.x) bsum[j] = g.b_ih[j] + g.b_hh[j];
}

// scale (0 or 1/(1-p)) of hidden unit `unit` of (row, t) at the input of layer `layer`
__device__ __forceinline__ float drop_scale(uint64_t seed, int layer, int64_t row, int t, int unit, uint32_t thr, float inv_keep) {
    U4 b = dropout_block(seed, layer, row, t, unit >> 3);
    return dropout_lane16(b, unit & 7) >= thr ? inv_keep : 0.f;
}

// Loads the x part of the warp's 8 operand rows for step t.  LAYER0: raw features [T][N][F];
// else the lower layer's hidden states [T][N][H] with the dropout mask applied.
template <bool LAYER0>
__device__ __forceinline__ void load_x_rows(float* Uw, int us, const float* __restrict__ xin, int64_t N, int64_t row_end,
                                            int64_t row0, int t, int kin, int kp, int lane, bool drop, uint64_t seed,
                                            int layer, uint32_t thr, float inv_keep) {
#pragma unroll
    for (int i = 0; i < ROWS_PER_WARP; ++i) {
        const int64_t row = row0 + i;
        if (LAYER0) {
            if (lane < kp) Uw[i * us + lane] = (lane < kin && row < row_end) ? xin[((int64_t)t * N + row) * kin + lane] : 0.f;
        } else {
#pragma unroll
            for (int hf = 0; hf < 2; ++hf) {
                const int u = lane + 32 * hf;
                float v = row < row_end ? xin[((int64_t)t * N + row) * LH + u] : 0.f;
                if (drop) v *= drop_scale(seed, layer, row, t, u, thr, inv_keep);
                Uw[i * us + u] = v;
            }
        }
    }
}

// acc[i][c] = bsum[c*32+lane] + sum_k U[i][k] * Wt[k][c*32+lane]
__device__ __forceinline__ void gate_gemm(float (&acc)[ROWS_PER_WARP][8], const float* Uw, int us, const float* Wt,
                                          const float* bsum, int ktot, int lane) {
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const float b = bsum[c * 32 + lane];
#pragma unroll
        for (int i = 0; i < ROWS_PER_WARP; ++i) acc[i][c] = b;
    }
    for (int k0 = 0; k0 < ktot; k0 += 4) {
        float4 a[ROWS_PER_WARP];
#pragma unroll
        for (int i = 0; i < ROWS_PER_WARP; ++i) a[i] = *reinterpret_cast<const float4*>(Uw + i * us + k0);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            float b[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) b[c] = Wt[(k0 + kk) * WT_STRIDE + c * 32 + lane];
#pragma unroll
            for (int i = 0; i < ROWS_PER_WARP; ++i) {
                const float av = kk == 0 ? a[i].x : kk == 1 ? a[i].y : kk == 2 ? a[i].z : a[i].w;
#pragma unroll
                for (int c = 0; c < 8; ++c) acc[i][c] = fmaf(av, b[c], acc[i][c]);
            }
        }
    }
}

__device__ __forceinline__ float sigm(float x) { return 1.0f / (1.0f + expf(-x)); }

// ------------------------------------------------------------------------------------------ forward
template <bool LAYER0>
__global__ void __launch_bounds__(L_THREADS, 1)
lstm_ffma_fwd_kernel(LayerGeom g, const float* __restrict__ xin, int64_t N, int T, float* __restrict__ h_out,
                     float* __restrict__ c_out, bool drop, uint64_t seed, float p) {
    extern __shared__ __align__(16) float smem[];
    float* Wt = smem;                                   // [ktot][257]
    float* bsum = Wt + g.ktot * WT_STRIDE;              // [256]
    float* U = bsum + LG;                               // [64][ktot+4]
    const int us = g.ktot + 4;
    load_weights(Wt, bsum, g);
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float* Uw = U + warp * ROWS_PER_WARP * us;
    const uint32_t thr = dropout_threshold(p);
    const float inv_keep = 1.0f / (1.0f - p);
    const int64_t ntiles = (N + L_ROWS - 1) / L_ROWS;

    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t row0 = tile * L_ROWS + warp * ROWS_PER_WARP;
        if (row0 >= N) continue;
        float cst[ROWS_PER_WARP][2];
#pragma unroll
        for (int i = 0; i < ROWS_PER_WARP; ++i) {
            cst[i][0] = cst[i][1] = 0.f;
            Uw[i * us + g.kp + lane] = 0.f;            // h_{-1} = 0
            Uw[i * us + g.kp + 32 + lane] = 0.f;
        }
        for (int t = 0; t < T; ++t) {
            load_x_rows<LAYER0>(Uw, us, xin, N, N, row0, t, g.kin, g.kp, lane, drop, seed, g.layer, thr, inv_keep);
            __syncwarp();
            float acc[ROWS_PER_WARP][8];
            gate_gemm(acc, Uw, us, Wt, bsum, g.ktot, lane);
            __syncwarp();
#pragma unroll
            for (int i = 0; i < ROWS_PER_WARP; ++i) {
                const int64_t row = row0 + i;
#pragma unroll
                for (int hf = 0; hf < 2; ++hf) {
                    const float ig = sigm(acc[i][0 + hf]), fg = sigm(acc[i][2 + hf]);
                    const float gg = tanhf(acc[i][4 + hf]), og = sigm(acc[i][6 + hf]);
                    const float c = fg * cst[i][hf] + ig * gg;
                    const float h = og * tanhf(c);
                    cst[i][hf] = c;
                    Uw[i * us + g.kp + hf * 32 + lane] = h;
                    if (row < N) {
                        const int64_t o = ((int64_t)t * N + row) * LH + hf * 32 + lane;
                        h_out[o] = h;
                        c_out[o] = c;
                    }
                }
            }
        }
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------ backward, recurrent part
// d_out: [T][N][H] gradient w.r.t. this layer's outputs from the layer above (already masked), or
// nullptr for the top layer, which only receives dhT at t = T-1.
template <bool LAYER0>
__global__ void __launch_bounds__(L_THREADS, 1)
lstm_ffma_bwd_kernel(LayerGeom g, const float* __restrict__ xin, int64_t N, int T, const float* __restrict__ h_st,
                     const float* __restrict__ c_st, const float* __restrict__ d_out, const float* __restrict__ dhT,
                     float* __restrict__ d_in, float* __restrict__ da_buf, int64_t chunk0, int64_t chunkN,
                     float* __restrict__ bpart, int accumulate, bool drop, uint64_t seed, float p) {
    extern __shared__ __align__(16) float smem[];
    float* Wt = smem;
    float* bsum = Wt + g.ktot * WT_STRIDE;
    float* U = bsum + LG;                               // [64][260]: [x | h_{t-1}] then da
    constexpr int us = LG + 4;
    load_weights(Wt, bsum, g);
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float* Uw = U + warp * ROWS_PER_WARP * us;
    const uint32_t thr = dropout_threshold(p);
    const float inv_keep = 1.0f / (1.0f - p);
    const int64_t ntiles = (chunkN + L_ROWS - 1) / L_ROWS;
    constexpr int NQ = LAYER0 ? 2 : 4;
    const int kbase = LAYER0 ? g.kp : 0;

    float dbias[8];
    float* bp = bpart + ((int64_t)blockIdx.x * L_WARPS + warp) * LG;
#pragma unroll
    for (int c = 0; c < 8; ++c) dbias[c] = accumulate ? bp[c * 32 + lane] : 0.f;

    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t rc0 = tile * L_ROWS + warp * ROWS_PER_WARP;     // row within the chunk
        if (rc0 >= chunkN) continue;
        const int64_t row0 = chunk0 + rc0;
        const int64_t Nend = chunk0 + chunkN;                         // rows >= Nend are padding
        float dc[ROWS_PER_WARP][2], dhr[ROWS_PER_WARP][2];
#pragma unroll
        for (int i = 0; i < ROWS_PER_WARP; ++i) dc[i][0] = dc[i][1] = dhr[i][0] = dhr[i][1] = 0.f;

        for (int t = T - 1; t >= 0; --t) {
            load_x_rows<LAYER0>(Uw, us, xin, N, Nend, row0, t, g.kin, g.kp, lane, drop, seed, g.layer, thr, inv_keep);
#pragma unroll
            for (int i = 0; i < ROWS_PER_WARP; ++i) {
                const int64_t row = row0 + i;
#pragma unroll
                for (int hf = 0; hf < 2; ++hf)
                    Uw[i * us + g.kp + hf * 32 + lane] =
                        (t > 0 && row < Nend) ? h_st[((int64_t)(t - 1) * N + row) * LH + hf * 32 + lane] : 0.f;
            }
            __syncwarp();
            float acc[ROWS_PER_WARP][8];
            gate_gemm(acc, Uw, us, Wt, bsum, g.ktot, lane);
            __syncwarp();
#pragma unroll
            for (int i = 0; i < ROWS_PER_WARP; ++i) {
                const int64_t row = row0 + i;
                const bool live = row < Nend;
#pragma unroll
                for (int hf = 0; hf < 2; ++hf) {
                    const int u = hf * 32 + lane;
                    const int64_t o = ((int64_t)t * N + row) * LH + u;
                    const float ig = sigm(acc[i][0 + hf]), fg = sigm(acc[i][2 + hf]);
                    const float gg = tanhf(acc[i][4 + hf]), og = sigm(acc[i][6 + hf]);
                    const float ct = live ? c_st[o] : 0.f;
                    const float cp = (t > 0 && live) ? c_st[o - N * LH] : 0.f;
                    const float tc = tanhf(ct);
                    float dh = dhr[i][hf];
                    if (live) {
                        if (d_out) dh += d_out[o];
                        else if (t == T - 1) dh += dhT[row * LH + u];
                    }
                    const float d_o = dh * tc;
                    const float dcc = dc[i][hf] + dh * og * (1.f - tc * tc);
                    const float d_i = dcc * gg, d_g = dcc * ig, d_f = dcc * cp;
                    dc[i][hf] = dcc * fg;
                    const float a_i = d_i * ig * (1.f - ig), a_f = d_f * fg * (1.f - fg);
                    const float a_g = d_g * (1.f - gg * gg), a_o = d_o * og * (1.f - og);
                    acc[i][0 + hf] = a_i; acc[i][2 + hf] = a_f; acc[i][4 + hf] = a_g; acc[i][6 + hf] = a_o;
                }
            }
            // da -> private rows of U (for the dgrad GEMM) and -> workspace (for the wgrad kernel)
#pragma unroll
            for (int i = 0; i < ROWS_PER_WARP; ++i) {
                const int64_t rc = rc0 + i;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const float v = (rc < chunkN) ? acc[i][c] : 0.f;
                    dbias[c] += v;
                    Uw[i * us + c * 32 + lane] = v;
                    if (rc < chunkN) da_buf[((int64_t)t * chunkN + rc) * LG + c * 32 + lane] = v;
                }
            }
            __syncwarp();
            // [dx | dh_{t-1}][i][k] = sum_j da[i][j] * Wt[k][j],  k = kbase + lane + 32 q
            float out[ROWS_PER_WARP][NQ];
#pragma unroll
            for (int i = 0; i < ROWS_PER_WARP; ++i)
#pragma unroll
                for (int q = 0; q < NQ; ++q) out[i][q] = 0.f;
            for (int j0 = 0; j0 < LG; j0 += 4) {
                float4 a[ROWS_PER_WARP];
#pragma unroll
                for (int i = 0; i < ROWS_PER_WARP; ++i) a[i] = *reinterpret_cast<const float4*>(Uw + i * us + j0);
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) {
                    float b[NQ];
#pragma unroll
                    for (int q = 0; q < NQ; ++q) b[q] = Wt[(kbase + lane + 32 * q) * WT_STRIDE + j0 + jj];
#pragma unroll
                    for (int i = 0; i < ROWS_PER_WARP; ++i) {
                        const float av = jj == 0 ? a[i].x : jj == 1 ? a[i].y : jj == 2 ? a[i].z : a[i].w;
#pragma unroll
                        for (int q = 0; q < NQ; ++q) out[i][q] = fmaf(av, b[q], out[i][q]);
                    }
                }
            }
            __syncwarp();
#pragma unroll
            for (int i = 0; i < ROWS_PER_WARP; ++i) {
                dhr[i][0] = out[i][NQ - 2];
                dhr[i][1] = out[i][NQ - 1];
                if (!LAYER0) {
                    const int64_t row = row0 + i;
                    if (row < Nend) {
#pragma unroll
                        for (int hf = 0; hf < 2; ++hf) {
                            const int u = hf * 32 + lane;
                            float v = out[i][hf];
                            if (drop) v *= drop_scale(seed, g.layer, row, t, u, thr, inv_keep);
                            d_in[((int64_t)t * N + row) * LH + u] = v;
                        }
                    }
                }
            }
        }
    }
#pragma unroll
    for (int c = 0; c < 8; ++c) bp[c * 32 + lane] = dbias[c];
}

// ------------------------------------------------------------------------------------------ weight gradients
// part[split][j][k] (+)= sum over this split's rows r=(t, rc) of da[r][j] * xh[r][k],
// xh = [x_t (kp wide, zero padded) | h_{t-1} (H wide)], zero beyond ktot.  grid = (2 j-halves, nsplit).
template <bool LAYER0>
__global__ void __launch_bounds__(256)
lstm_ffma_wgrad_kernel(LayerGeom g, const float* __restrict__ xin, int64_t N, int T, const float* __restrict__ h_st,
                       const float* __restrict__ da_buf, int64_t chunk0, int64_t chunkN, float* __restrict__ part,
                       int accumulate, bool drop, uint64_t seed, float p) {
    __shared__ __align__(16) float Da[WG_ROWS][WG_LD];
    __shared__ __align__(16) float Xh[WG_ROWS][WG_LD];
    const int jh = blockIdx.x, split = blockIdx.y, nsplit = gridDim.y;
    const int64_t R = (int64_t)T * chunkN;
    const int64_t nblk = (R + WG_ROWS - 1) / WG_ROWS;
    const int64_t b_lo = nblk * split / nsplit, b_hi = nblk * (split + 1) / nsplit;
    const uint32_t thr = dropout_threshold(p);
    const float inv_keep = 1.0f / (1.0f - p);

    const int lr = threadIdx.x >> 3, seg = threadIdx.x & 7;      // loader: row lr, 16 floats at seg*16
    const int tj = threadIdx.x >> 4, tk = threadIdx.x & 15;      // compute: j in {tj*4.., 64+tj*4..}, k likewise
    float acc[8][8];
#pragma unroll
    for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 8; ++b) acc[a][b] = 0.f;

    for (int64_t blk = b_lo; blk < b_hi; ++blk) {
        const int64_t r = blk * WG_ROWS + lr;
        const bool live = r < R;
        const int t = live ? (int)(r / chunkN) : 0;
        const int64_t rc = live ? r - (int64_t)t * chunkN : 0;
        const int64_t row = chunk0 + rc;
        __syncthreads();
#pragma unroll
        for (int v = 0; v < 4; ++v) {
            float4 d = make_float4(0.f, 0.f, 0.f, 0.f);
            if (live) d = *reinterpret_cast<const float4*>(da_buf + r * LG + jh * 128 + seg * 16 + v * 4);
            *reinterpret_cast<float4*>(&Da[lr][seg * 16 + v * 4]) = d;
        }
#pragma unroll
        for (int v = 0; v < 16; ++v) {
            const int k = seg * 16 + v;
            float val = 0.f;
            if (live) {
                if (k < g.kp) {
                    if (LAYER0) { if (k < g.kin) val = xin[((int64_t)t * N + row) * g.kin + k]; }
                    else {
                        val = xin[((int64_t)t * N + row) * LH + k];
                        if (drop) val *= drop_scale(seed, g.layer, row, t, k, thr, inv_keep);
                    }
                } else if (k < g.ktot) {
                    if (t > 0) val = h_st[((int64_t)(t - 1) * N + row) * LH + (k - g.kp)];
                }
            }
            Xh[lr][k] = val;
        }
        __syncthreads();
#pragma unroll 4
        for (int rr = 0; rr < WG_ROWS; ++rr) {
            const float4 a0 = *reinterpret_cast<const float4*>(&Da[rr][tj * 4]);
            const float4 a1 = *reinterpret_cast<const float4*>(&Da[rr][64 + tj * 4]);
            const float4 b0 = *reinterpret_cast<const float4*>(&Xh[rr][tk * 4]);
            const float4 b1 = *reinterpret_cast<const float4*>(&Xh[rr][64 + tk * 4]);
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int a = 0; a < 8; ++a)
#pragma unroll
                for (int b = 0; b < 8; ++b) acc[a][b] = fmaf(av[a], bv[b], acc[a][b]);
        }
    }
    float* out = part + (int64_t)split * LG * 128;
#pragma unroll
    for (int a = 0; a < 8; ++a) {
        const int j = jh * 128 + (a < 4 ? tj * 4 + a : 64 + tj * 4 + (a - 4));
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            const int k = b < 4 ? tk * 4 + b : 64 + tk * 4 + (b - 4);
            float* o = out + j * 128 + k;
            *o = accumulate ? *o + acc[a][b] : acc[a][b];
        }
    }
}

// partials -> the layer's slice of the flat gradient vector (fixed summation order)
__global__ void __launch_bounds__(256)
lstm_grad_reduce_kernel(const float* __restrict__ wpart, int nsplit, const float* __restrict__ bpart, int nbpart, int kin,
                        int kp, float ih_scale, float* __restrict__ gw_ih, float* __restrict__ gw_hh, float* __restrict__ gb_ih,
                        float* __restrict__ gb_hh) {
    const int n_ih = LG * kin, n_hh = LG * LH;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx < n_ih + n_hh) {
        int j, k;
        float* dst;
        if (idx < n_ih) { j = idx / kin; k = idx - j * kin; dst = gw_ih + idx; }
        else { int i = idx - n_ih; j = i / LH; k = kp + (i - j * LH); dst = gw_hh + i; }
        float s = 0.f;
        for (int sp = 0; sp < nsplit; ++sp) s += wpart[((int64_t)sp * LG + j) * 128 + k];
        *dst = idx < n_ih ? s * ih_scale : s;        // ih_scale: a factor folded into the packed W_ih by the caller (tcgen05 dropout), else 1
    } else if (idx < n_ih + n_hh + LG) {
        const int j = idx - n_ih - n_hh;
        float s = 0.f;
        for (int w = 0; w < nbpart; ++w) s += bpart[(int64_t)w * LG + j];
        gb_ih[j] = s;
        gb_hh[j] = s;
    }
}

// ------------------------------------------------------------------------------------------ host side
static LayerGeom make_geom(const float* params, int F, int L, int layer) {
    LayerGeom g;
    g.layer = layer;
    g.kin = layer == 0 ? F : LH;
    g.kp = pad4(g.kin);
    g.ktot = g.kp + LH;
    const float* base = params + b200_param_offset(F, LH, L, layer);
    g.w_ih = base;
    g.w_hh = g.w_ih + (int64_t)LG * g.kin;
    g.b_ih = g.w_hh + (int64_t)LG * LH;
    g.b_hh = g.b_ih + LG;
    return g;
}

struct FfmaWs {
    float* dx[2];     // [T][N][H] inter-layer gradients (ping-pong), only when L > 1
    float* da;        // [T][chunk][4H]
    float* wpart;     // [nsplit][4H][128]
    float* bpart;     // [grid*8][4H]
    int64_t chunk;
    int nsplit, bwd_grid;
};

static size_t ffma_ws_layout(int64_t N, int T, int L, size_t avail, char* base, FfmaWs* ws) {
    const int sms = num_sms();
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t)255; return base ? base + o : (char*)nullptr; };
    char* dx0 = nullptr; char* dx1 = nullptr;
    if (L > 1) { dx0 = take((size_t)T * N * LH * 4); dx1 = take((size_t)T * N * LH * 4); }
    const int nsplit = sms;                               // 2 CTAs (j halves) per split
    char* wp = take((size_t)nsplit * LG * 128 * 4);
    char* bp = take((size_t)sms * L_WARPS * LG * 4);
    // da chunk: as many sequences as fit (multiple of the CTA tile), capped at 64 Ki sequences
    int64_t chunk = N < 65536 ? N : 65536;
    if (base) {
        const size_t per_seq = (size_t)T * LG * 4;
        const size_t room = avail > off ? avail - off : 0;
        int64_t fit = (int64_t)(room / per_seq);
        if (fit < chunk) chunk = fit / L_ROWS * L_ROWS;
    }
    char* da = take((size_t)T * (chunk > 0 ? chunk : 0) * LG * 4);
    if (ws) {
        ws->dx[0] = (float*)dx0; ws->dx[1] = (float*)dx1; ws->da = (float*)da; ws->wpart = (float*)wp; ws->bpart = (float*)bp;
        ws->chunk = chunk; ws->nsplit = nsplit; ws->bwd_grid = sms;
    }
    return off;
}

size_t ffma_workspace_bytes(int64_t N, int T, int L) { return ffma_ws_layout(N, T, L, 0, nullptr, nullptr); }

// Shared with the tcgen05 engine (lstm_tc.cu): the dx ping-pong buffers and the partial-gradient areas of the same
// workspace layout, and the fixed-order reduction of the partials into the layer's gradient slice.
int ffma_ws_pointers(int64_t N, int T, int L, void* workspace, size_t ws_bytes, float** dx0, float** dx1, float** wpart,
                     float** bpart, int* nsplit) {
    FfmaWs ws;
    const size_t need = ffma_ws_layout(N, T, L, ws_bytes, (char*)workspace, &ws);
    (void)need;
    *dx0 = ws.dx[0]; *dx1 = ws.dx[1]; *wpart = ws.wpart; *bpart = ws.bpart; *nsplit = ws.nsplit;
    // everything up to (and including) bpart must fit; the da chunk that follows is unused by the tcgen05 engine
    const size_t fixed = (size_t)((char*)ws.bpart - (char*)workspace) + (size_t)ws.bwd_grid * L_WARPS * LG * 4;
    if (fixed > ws_bytes) { set_error("lstm_bwd: workspace too small (%zu < %zu)", ws_bytes, fixed); return B200_ERR_INVALID; }
    return B200_OK;
}

int launch_grad_reduce(const float* wpart, int nsplit, const float* bpart, int nbpart, int kin, int kp, float ih_scale, float* gw_ih,
                       float* gw_hh, float* gb_ih, float* gb_hh, cudaStream_t st) {
    const int total = LG * kin + LG * LH + LG;
    { TimedLaunch _t("lstm_grad_reduce_kernel", st); lstm_grad_reduce_kernel<<<(total + 255) / 256, 256, 0, st>>>(wpart, nsplit, bpart, nbpart, kin, kp, ih_scale, gw_ih, gw_hh, gb_ih, gb_hh); }
    B200_LAUNCH_CHECK("lstm_grad_reduce_kernel");
    return B200_OK;
}

static size_t fwd_smem(const LayerGeom& g) { return ((size_t)g.ktot * WT_STRIDE + LG + (size_t)L_ROWS * (g.ktot + 4)) * 4; }
static size_t bwd_smem(const LayerGeom& g) { return ((size_t)g.ktot * WT_STRIDE + LG + (size_t)L_ROWS * (LG + 4)) * 4; }

int ffma_fwd(const float* params, int F, int H, int L, const float* x, int64_t N, int T, float* stash_h, float* stash_c,
             float p, uint64_t seed, int training, cudaStream_t st) {
    if (H != LH) { set_error("FFMA engine supports hidden=64 only, got %d", H); return B200_ERR_UNSUPPORTED; }
    const int64_t ntiles = (N + L_ROWS - 1) / L_ROWS;
    const int grid = (int)(ntiles < num_sms() ? ntiles : num_sms());
    const bool drop = training && p > 0.f;
    const int64_t lsz = (int64_t)T * N * LH;
    for (int l = 0; l < L; ++l) {
        LayerGeom g = make_geom(params, F, L, l);
        const size_t smem = fwd_smem(g);
        if (l == 0) {
            B200_CUDA(cudaFuncSetAttribute(lstm_ffma_fwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            TimedLaunch _t("lstm_ffma_fwd_kernel", st);
            lstm_ffma_fwd_kernel<true><<<grid, L_THREADS, smem, st>>>(g, x, N, T, stash_h, stash_c, false, seed, p);
        } else {
            B200_CUDA(cudaFuncSetAttribute(lstm_ffma_fwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            TimedLaunch _t("lstm_ffma_fwd_kernel", st);
            lstm_ffma_fwd_kernel<false><<<grid, L_THREADS, smem, st>>>(g, stash_h + (l - 1) * lsz, N, T, stash_h + l * lsz,
                                                                      stash_c + l * lsz, drop, seed, p);
        }
        B200_LAUNCH_CHECK("lstm_ffma_fwd_kernel");
    }
    return B200_OK;
}

int ffma_bwd_layer(int layer, const float* params, int F, int H, int L, const float* x, int64_t N, int T,
                   const float* stash_h, const float* stash_c, const float* dhT, float* grads, float p, uint64_t seed,
                   int training, void* workspace, size_t ws_bytes, cudaStream_t st) {
    if (H != LH) { set_error("FFMA engine supports hidden=64 only, got %d", H); return B200_ERR_UNSUPPORTED; }
    FfmaWs ws;
    const size_t need = ffma_ws_layout(N, T, L, ws_bytes, (char*)workspace, &ws);
    if (ws.chunk <= 0 || need > ws_bytes) { set_error("lstm_bwd: workspace too small (%zu bytes given)", ws_bytes); return B200_ERR_INVALID; }
    const bool drop = training && p > 0.f;
    const int64_t lsz = (int64_t)T * N * LH;
    LayerGeom g = make_geom(params, F, L, layer);
    const float* xin = layer == 0 ? x : stash_h + (layer - 1) * lsz;
    const float* h_st = stash_h + layer * lsz;
    const float* c_st = stash_c + layer * lsz;
    // gradient w.r.t. this layer's outputs: written by layer+1 into dx[(layer+1)&1]; this layer writes dx[layer&1]
    const float* d_out = layer == L - 1 ? nullptr : ws.dx[(layer + 1) & 1];
    float* d_in = layer == 0 ? nullptr : ws.dx[layer & 1];
    const size_t smem = bwd_smem(g);
    int acc = 0;
    for (int64_t c0 = 0; c0 < N; c0 += ws.chunk, acc = 1) {
        const int64_t cn = N - c0 < ws.chunk ? N - c0 : ws.chunk;
        const int64_t ntiles = (cn + L_ROWS - 1) / L_ROWS;
        // the grid must not change between chunks: bias partial slots are indexed by blockIdx
        const int grid = ws.bwd_grid;
        (void)ntiles;
        if (layer == 0) {
            B200_CUDA(cudaFuncSetAttribute(lstm_ffma_bwd_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            TimedLaunch _t("lstm_ffma_bwd_kernel", st);
            lstm_ffma_bwd_kernel<true><<<grid, L_THREADS, smem, st>>>(g, xin, N, T, h_st, c_st, d_out, dhT, d_in, ws.da, c0, cn,
                                                                      ws.bpart, acc, false, seed, p);
        } else {
            B200_CUDA(cudaFuncSetAttribute(lstm_ffma_bwd_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            TimedLaunch _t("lstm_ffma_bwd_kernel", st);
            lstm_ffma_bwd_kernel<false><<<grid, L_THREADS, smem, st>>>(g, xin, N, T, h_st, c_st, d_out, dhT, d_in, ws.da, c0, cn,
                                                                       ws.bpart, acc, drop, seed, p);
        }
        B200_LAUNCH_CHECK("lstm_ffma_bwd_kernel");
        dim3 wg(2, ws.nsplit);
        if (layer == 0)
            { TimedLaunch _t("lstm_ffma_wgrad_kernel", st); lstm_ffma_wgrad_kernel<true><<<wg, 256, 0, st>>>(g, xin, N, T, h_st, ws.da, c0, cn, ws.wpart, acc, false, seed, p); }
        else
            { TimedLaunch _t("lstm_ffma_wgrad_kernel", st); lstm_ffma_wgrad_kernel<false><<<wg, 256, 0, st>>>(g, xin, N, T, h_st, ws.da, c0, cn, ws.wpart, acc, drop, seed, p); }
            B200_LAUNCH_CHECK("lstm_ffma_wgrad_kernel");
    }
    float* gbase = grads + b200_param_offset(F, LH, L, layer);
    float* gw_ih = gbase;
    float* gw_hh = gw_ih + (int64_t)LG * g.kin;
    float* gb_ih = gw_hh + (int64_t)LG * LH;
    float* gb_hh = gb_ih + LG;
    const int total = LG * g.kin + LG * LH + LG;
    { TimedLaunch _t("lstm_grad_reduce_kernel", st); lstm_grad_reduce_kernel<<<(total + 255) / 256, 256, 0, st>>>(ws.wpart, ws.nsplit, ws.bpart, ws.bwd_grid * L_WARPS, g.kin, g.kp,
                                                                 1.0f, gw_ih, gw_hh, gb_ih, gb_hh); }
    B200_LAUNCH_CHECK("lstm_grad_reduce_kernel");
    return B200_OK;
}

}  // namespace b200
