// Global gradient norm (for clipping) and the fused clip + Adam update over ONE flat parameter vector.
//
// Replaces (reference file:line)
//   Trainer(gradient_clip_val=...) -> clip_grad_norm_          train.py:171 (configs/trainer/*.yaml:2)
//   torch.optim.Adam(lr, weight_decay=1e-5).step()             src/model.py:150-154
//   optimizer.zero_grad()                                       Lightning automatic optimisation
//
// Both kernels are HBM-bound: sqnorm reads 4 B/param, adam_step reads p,g,m,v and writes p,m,v(,g)
// = 28 (+4) B/param with 128-bit accesses.  At the reference's 51,074..250,754 parameters they are
// launch-latency bound (<= 8 MB moved); the roofline claim is made on multi-million parameter vectors
// (DESIGN.md).
#include "common.cuh"

namespace b200 {

constexpr int SQ_THREADS = 256;
constexpr int SQ_MAX_BLOCKS = 1024;

__global__ void __launch_bounds__(SQ_THREADS)
sqnorm_partial_kernel(const float* __restrict__ g, int64_t count, double* __restrict__ part) {
    __shared__ double red[SQ_THREADS / 32];
    double acc = 0.0;
    const int64_t n4 = count >> 2;
    const float4* g4 = reinterpret_cast<const float4*>(g);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        float4 v = g4[i];
        float s = v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
        acc += (double)s;
    }
    if (blockIdx.x == 0)
        for (int64_t i = (n4 << 2) + threadIdx.x; i < count; i += blockDim.x) acc += (double)g[i] * (double)g[i];
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < SQ_THREADS / 32; ++w) t += red[w];
        part[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(256)
sqnorm_final_kernel(const double* __restrict__ part, int nblocks, float* __restrict__ out) {
    __shared__ double red[256];
    double v = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += 256) v += part[i];
    red[threadIdx.x] = v;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[0] = (float)red[0];
}

struct AdamArgs {
    float lr, beta1, beta2, eps, wd, clip_val, step_size, rsqrt_bc2;
};

__device__ __forceinline__ void adam_elem(float& p, float& g, float& m, float& v, float coef, const AdamArgs& a) {
    float gg = g * coef + a.wd * p;                       // clip, then L2-coupled weight decay (torch: grad.add(param, alpha=wd))
    m = m + (gg - m) * (1.0f - a.beta1);                  // exp_avg.lerp_(grad, 1 - beta1)
    v = v * a.beta2 + (1.0f - a.beta2) * gg * gg;         // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
    float denom = sqrtf(v) * a.rsqrt_bc2 + a.eps;         // (sqrt(v) / sqrt(bias_correction2)).add_(eps)
    p = p - a.step_size * (m / denom);                    // param.addcdiv_(exp_avg, denom, value=-lr/bias_correction1)
}

__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, int64_t count,
            AdamArgs a, const float* __restrict__ sqnorm, int zero_grad) {
    float coef = 1.0f;
    if (sqnorm != nullptr && a.clip_val > 0.f) {
        float nrm = sqrtf(*sqnorm);
        coef = fminf(1.0f, a.clip_val / (nrm + 1e-6f));    // clip_grad_norm_: clamp(max_norm / (total_norm + 1e-6), max=1)
    }
    const int64_t n4 = count >> 2;
    float4* p4 = reinterpret_cast<float4*>(p);
    float4* g4 = reinterpret_cast<float4*>(g);
    float4* m4 = reinterpret_cast<float4*>(m);
    float4* v4 = reinterpret_cast<float4*>(v);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        float4 pp = p4[i], gg = g4[i], mm = m4[i], vv = v4[i];
        adam_elem(pp.x, gg.x, mm.x, vv.x, coef, a);
        adam_elem(pp.y, gg.y, mm.y, vv.y, coef, a);
        adam_elem(pp.z, gg.z, mm.z, vv.z, coef, a);
        adam_elem(pp.w, gg.w, mm.w, vv.w, coef, a);
        p4[i] = pp; m4[i] = mm; v4[i] = vv;
        if (zero_grad) g4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (blockIdx.x == 0)
        for (int64_t i = (n4 << 2) + threadIdx.x; i < count; i += blockDim.x) {
            float pp = p[i], gg = g[i], mm = m[i], vv = v[i];
            adam_elem(pp, gg, mm, vv, coef, a);
            p[i] = pp; m[i] = mm; v[i] = vv;
            if (zero_grad) g[i] = 0.f;
        }
}

// Fused global-norm clip + Adam in ONE launch for small parameter vectors: every CTA first reduces ||g||^2 over the WHOLE vector
// (same order in every CTA, fp64 accumulation: all CTAs arrive at the identical clip coefficient; the vector is L2-resident,
// 200 KB for the small model), then updates its own slice.  g is only read, so no CTA can disturb another CTA's norm.
__global__ void __launch_bounds__(256)
clip_adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, int64_t count,
                 AdamArgs a, float* __restrict__ sqnorm_out) {
    __shared__ double red[8];
    __shared__ float coef_s;
    const int64_t n4 = count >> 2;
    const float4* g4 = reinterpret_cast<const float4*>(g);
    double acc = 0.0;
    for (int64_t i = threadIdx.x; i < n4; i += blockDim.x) {
        const float4 x = g4[i];
        acc += (double)(x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w);
    }
    for (int64_t i = (n4 << 2) + threadIdx.x; i < count; i += blockDim.x) acc += (double)g[i] * (double)g[i];
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += red[w];
        const float sq = (float)t;
        if (blockIdx.x == 0 && sqnorm_out) sqnorm_out[0] = sq;
        coef_s = a.clip_val > 0.f ? fminf(1.0f, a.clip_val / (sqrtf(sq) + 1e-6f)) : 1.0f;
    }
    __syncthreads();
    const float coef = coef_s;
    float4* p4 = reinterpret_cast<float4*>(p);
    float4* m4 = reinterpret_cast<float4*>(m);
    float4* v4 = reinterpret_cast<float4*>(v);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        float4 pp = p4[i], gg = g4[i], mm = m4[i], vv = v4[i];
        adam_elem(pp.x, gg.x, mm.x, vv.x, coef, a);
        adam_elem(pp.y, gg.y, mm.y, vv.y, coef, a);
        adam_elem(pp.z, gg.z, mm.z, vv.z, coef, a);
        adam_elem(pp.w, gg.w, mm.w, vv.w, coef, a);
        p4[i] = pp; m4[i] = mm; v4[i] = vv;
    }
    if (blockIdx.x == 0)
        for (int64_t i = (n4 << 2) + threadIdx.x; i < count; i += blockDim.x) {
            float pp = p[i], gg = g[i], mm = m[i], vv = v[i];
            adam_elem(pp, gg, mm, vv, coef, a);
            p[i] = pp; m[i] = mm; v[i] = vv;
        }
}

// Scratch of b200_grad_sqnorm_static: one fp64 partial per 4,096-gradient tile, so 1,024 slots cover
// B200_SQNORM_STATIC_MAX gradients.  Lives in every CUDA context that loads the library until exit (8 KB).
// In host code this name is only a handle; the device address comes from cudaGetSymbolAddress.
static __device__ __align__(256) double g_sqnorm_static_scratch[SQ_MAX_BLOCKS];

static_assert((int64_t)B200_SQNORM_STATIC_MAX == (int64_t)SQ_MAX_BLOCKS * SQ_THREADS * 16,
              "B200_SQNORM_STATIC_MAX (header) must equal the gradients covered by SQ_MAX_BLOCKS scratch slots");

// dst += scale * src (micro-batch gradient accumulation; 8 B read + 4 B written per element)
__global__ void __launch_bounds__(256)
grad_accumulate_kernel(float* __restrict__ dst, const float* __restrict__ src, int64_t count, float scale) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x)
        dst[i] = fmaf(scale, src[i], dst[i]);
}

static int sq_blocks(int64_t count) {
    int64_t b = (count / 4 + SQ_THREADS * 4 - 1) / (SQ_THREADS * 4);
    if (b < 1) b = 1;
    if (b > SQ_MAX_BLOCKS) b = SQ_MAX_BLOCKS;
    return (int)b;
}

}  // namespace b200

using namespace b200;

extern "C" size_t b200_sqnorm_workspace_bytes(int64_t count) { (void)count; return SQ_MAX_BLOCKS * sizeof(double); }

extern "C" int b200_grad_sqnorm(const float* grads, int64_t count, float* sqnorm_out, void* workspace,
                                size_t workspace_bytes, void* stream) {
    B200_CHECK_ARG(count > 0 && grads && sqnorm_out && workspace, "grad_sqnorm: bad arguments");
    B200_CHECK_ARG(((uintptr_t)grads & 15) == 0, "grad_sqnorm: grads must be 16-byte aligned");
    B200_CHECK_ARG(workspace_bytes >= SQ_MAX_BLOCKS * sizeof(double), "grad_sqnorm: workspace too small");
    int nb = sq_blocks(count);
    { TimedLaunch _t("sqnorm_partial_kernel", (cudaStream_t)stream); sqnorm_partial_kernel<<<nb, SQ_THREADS, 0, (cudaStream_t)stream>>>(grads, count, (double*)workspace); }
    B200_LAUNCH_CHECK("sqnorm_partial_kernel");
    { TimedLaunch _t("sqnorm_final_kernel", (cudaStream_t)stream); sqnorm_final_kernel<<<1, 256, 0, (cudaStream_t)stream>>>((const double*)workspace, nb, sqnorm_out); }
    B200_LAUNCH_CHECK("sqnorm_final_kernel");
    return B200_OK;
}

extern "C" int b200_adam_step(float* params, float* grads, float* exp_avg, float* exp_avg_sq, int64_t count, float lr,
                              float beta1, float beta2, float eps, float weight_decay, int64_t step, float clip_val,
                              const float* sqnorm, int zero_grad, void* stream) {
    B200_CHECK_ARG(count > 0 && params && grads && exp_avg && exp_avg_sq, "adam_step: bad arguments");
    B200_CHECK_ARG(step >= 1, "adam_step: step is 1-based, got %lld", (long long)step);
    B200_CHECK_ARG((((uintptr_t)params | (uintptr_t)grads | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq) & 15) == 0,
                   "adam_step: buffers must be 16-byte aligned");
    AdamArgs a;
    a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.wd = weight_decay; a.clip_val = clip_val;
    // bias corrections in double on the host, exactly as torch's python scalars
    double bc1 = 1.0 - pow((double)beta1, (double)step);
    double bc2 = 1.0 - pow((double)beta2, (double)step);
    a.step_size = (float)((double)lr / bc1);
    a.rsqrt_bc2 = (float)(1.0 / sqrt(bc2));
    int64_t nb = (count / 4 + 255) / 256;
    int64_t cap = (int64_t)num_sms() * 8;
    if (nb > cap) nb = cap;
    if (nb < 1) nb = 1;
    { TimedLaunch _t("adam_kernel", (cudaStream_t)stream); adam_kernel<<<(unsigned)nb, 256, 0, (cudaStream_t)stream>>>(params, grads, exp_avg, exp_avg_sq, count, a, sqnorm, zero_grad); }
    B200_LAUNCH_CHECK("adam_kernel");
    return B200_OK;
}

extern "C" int b200_grad_accumulate(float* dst, const float* src, int64_t count, float scale, void* stream) {
    B200_CHECK_ARG(count > 0 && dst && src, "grad_accumulate: bad arguments");
    int64_t nb = (count + 255) / 256;
    const int64_t cap = (int64_t)num_sms() * 8;
    if (nb > cap) nb = cap;
    { TimedLaunch _t("grad_accumulate_kernel", (cudaStream_t)stream); grad_accumulate_kernel<<<(unsigned)nb, 256, 0, (cudaStream_t)stream>>>(dst, src, count, scale); }
    B200_LAUNCH_CHECK("grad_accumulate_kernel");
    return B200_OK;
}

extern "C" int b200_clip_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t count, float lr,
                                   float beta1, float beta2, float eps, float weight_decay, int64_t step, float clip_val,
                                   float* sqnorm_out, void* stream) {
    B200_CHECK_ARG(count > 0 && params && grads && exp_avg && exp_avg_sq, "clip_adam_step: bad arguments");
    B200_CHECK_ARG(step >= 1, "clip_adam_step: step is 1-based, got %lld", (long long)step);
    B200_CHECK_ARG((((uintptr_t)params | (uintptr_t)grads | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq) & 15) == 0,
                   "clip_adam_step: buffers must be 16-byte aligned");
    if (count > (int64_t)B200_CLIP_ADAM_FUSED_MAX) {
        set_error("clip_adam_step: %lld parameters exceed the fused kernel's limit (%lld): use b200_grad_sqnorm + b200_adam_step",
                  (long long)count, (long long)B200_CLIP_ADAM_FUSED_MAX);
        return B200_ERR_UNSUPPORTED;
    }
    AdamArgs a;
    a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.wd = weight_decay; a.clip_val = clip_val;
    const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
    a.step_size = (float)((double)lr / bc1);
    a.rsqrt_bc2 = (float)(1.0 / sqrt(bc2));
    int64_t nb = (count / 4 + 255) / 256;
    if (nb < 1) nb = 1;
    { TimedLaunch _t("clip_adam_kernel", (cudaStream_t)stream); clip_adam_kernel<<<(unsigned)nb, 256, 0, (cudaStream_t)stream>>>(params, grads, exp_avg, exp_avg_sq, count, a, sqnorm_out); }
    B200_LAUNCH_CHECK("clip_adam_kernel");
    return B200_OK;
}

// Allocation-free variant of b200_grad_sqnorm (see include/b200_lstm.h).
extern "C" int b200_grad_sqnorm_static(const float* grads, int64_t count, float* sqnorm_out, void* stream) {
    B200_CHECK_ARG(count > 0 && grads && sqnorm_out, "grad_sqnorm_static: bad arguments");
    B200_CHECK_ARG(count <= (int64_t)B200_SQNORM_STATIC_MAX,
                   "grad_sqnorm_static: %lld gradients exceed the built-in scratch (max %lld); use b200_grad_sqnorm with a caller workspace",
                   (long long)count, (long long)B200_SQNORM_STATIC_MAX);
    B200_CHECK_ARG(((uintptr_t)grads & 15) == 0, "grad_sqnorm_static: grads must be 16-byte aligned");
    // The array's device address is looked up on every call (the device is current): after the first call on a device
    // this is a table lookup inside the runtime, and it stays correct across cudaDeviceReset / context re-creation,
    // which a cached pointer would not.  The first call per device loads the module: make it before any stream capture.
    double* scratch = nullptr;
    B200_CUDA(cudaGetSymbolAddress((void**)&scratch, g_sqnorm_static_scratch));
    const int nb = sq_blocks(count);       // <= SQ_MAX_BLOCKS; every slot read below is written first, nothing to clear
    { TimedLaunch _t("sqnorm_partial_kernel", (cudaStream_t)stream); sqnorm_partial_kernel<<<nb, SQ_THREADS, 0, (cudaStream_t)stream>>>(grads, count, scratch); }
    B200_LAUNCH_CHECK("sqnorm_partial_kernel");
    { TimedLaunch _t("sqnorm_final_kernel", (cudaStream_t)stream); sqnorm_final_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(scratch, nb, sqnorm_out); }
    B200_LAUNCH_CHECK("sqnorm_final_kernel");
    return B200_OK;
}
