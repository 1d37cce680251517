// C-ABI glue: error reporting, parameter-vector geometry and the LSTM engine dispatch.
#include <stdarg.h>

#include <atomic>
#include <mutex>

#include "common.cuh"

namespace b200 {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
    set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
    return -1000 - (int)e;
}

// ---- per-kernel timing registry ---------------------------------------------------------------
// Process-wide and OFF unless b200_timing_enable(1) was called (bench.py's roofline leg): the only mutable state of the
// library besides b200_set_option's two tuning values.  A mutex serialises the registry, so concurrent host threads are
// safe; while it is off, timing_begin/timing_end are one relaxed flag test.
struct TimingSlot {
    const char* name;
    cudaEvent_t ev[2 * 16384];
    int used;       // events recorded (pairs * 2)
    int allocated;  // events created
};
static TimingSlot g_slots[32];
static int g_nslots = 0;
static std::atomic<bool> g_timing{false};
static std::mutex g_timing_mu;
static thread_local TimingSlot* g_open = nullptr;      // a launch is bracketed by the thread that issues it

void timing_begin(const char* name, cudaStream_t st) {
    if (!g_timing.load(std::memory_order_relaxed)) return;
    std::lock_guard<std::mutex> lock(g_timing_mu);
    TimingSlot* s = nullptr;
    for (int i = 0; i < g_nslots; ++i)
        if (g_slots[i].name == name || strcmp(g_slots[i].name, name) == 0) { s = &g_slots[i]; break; }
    if (!s) {
        if (g_nslots == 32) return;
        s = &g_slots[g_nslots++];
        s->name = name; s->used = 0; s->allocated = 0;
    }
    if (s->used + 2 > 2 * 16384) return;
    while (s->allocated < s->used + 2) cudaEventCreate(&s->ev[s->allocated++]);
    cudaEventRecord(s->ev[s->used], st);
    g_open = s;
}

void timing_end(cudaStream_t st) {
    if (!g_open) return;
    std::lock_guard<std::mutex> lock(g_timing_mu);
    cudaEventRecord(g_open->ev[g_open->used + 1], st);
    g_open->used += 2;
    g_open = nullptr;
}

// engines (lstm_ffma.cu, lstm_tc.cu)
size_t ffma_workspace_bytes(int64_t N, int T, int L);
int ffma_fwd(const float* params, int F, int H, int L, const float* x, int64_t N, int T, float* stash_h, float* stash_c,
             float p, uint64_t seed, int training, cudaStream_t st);
int ffma_bwd_layer(int layer, const float* params, int F, int H, int L, const float* x, int64_t N, int T,
                   const float* stash_h, const float* stash_c, const float* dhT, float* grads, float p, uint64_t seed,
                   int training, void* workspace, size_t ws_bytes, cudaStream_t st);
bool tc_available();
void tc_set_prefetch(int mask);
void tc_set_fwd_eh(int eh);
void tc_set_wgrad(int v);
void tc_set_dgrad(int v);
void wide_set_fuse(int on);
size_t tc_workspace_bytes(int mode, int64_t N, int T, int F, int H, int L);
int tc_fwd(int mode, const float* params, int F, int H, int L, const float* x, int64_t N, int T, float* stash_h,
           float* stash_c, float p, uint64_t seed, int training, void* workspace, size_t ws_bytes, cudaStream_t st);
int tc_bwd_layer(int mode, int layer, const float* params, int F, int H, int L, const float* x, int64_t N, int T,
                 const float* stash_h, const float* stash_c, const float* dhT, float* grads, float p, uint64_t seed,
                 int training, void* workspace, size_t ws_bytes, cudaStream_t st);
int tc_hidden_states(int mode, int layer, int L, int64_t N, int T, const void* workspace, size_t ws_bytes, float* out, cudaStream_t st);
// wide engine (lstm_wide.cu): hidden sizes other than 64 in the tensor-core modes
size_t wide_workspace_bytes(int mode, int64_t N, int T, int F, int H, int L);
void wide_set_images(int on);
void wide2_set_dbg(int v);
void wide2_set_mc(int v);
bool wide_uses_images(int mode, int H, int F);
int wide2_hidden_states(int layer, int F, int H, int L, int64_t N, int T, const void* workspace, size_t ws_bytes, float* out, cudaStream_t st);
int wide_fwd(int mode, const float* params, int F, int H, int L, const float* x, int64_t N, int T, float* stash_h,
             float* stash_c, float p, uint64_t seed, int training, void* workspace, size_t ws_bytes, cudaStream_t st);
int wide_bwd_layer(int mode, int layer, const float* params, int F, int H, int L, const float* x, int64_t N, int T,
                   const float* stash_h, const float* stash_c, const float* dhT, float* grads, float p, uint64_t seed,
                   int training, void* workspace, size_t ws_bytes, cudaStream_t st);

}  // namespace b200

using namespace b200;

extern "C" int b200_version(void) { return 200; }
#ifndef B200_SRC_HASH
#define B200_SRC_HASH "unknown"
#endif
extern "C" const char* b200_build_hash(void) { return B200_SRC_HASH; }
extern "C" const char* b200_last_error(void) { return g_err; }
extern "C" int b200_mode_available(int mode) {
    if (mode == B200_MODE_FFMA) return 1;
    if (mode == B200_MODE_TC_PRECISE || mode == B200_MODE_TC_FAST) return tc_available() ? 1 : 0;
    return 0;
}

extern "C" int b200_timing_enable(int on) {
    std::lock_guard<std::mutex> lock(g_timing_mu);
    g_timing.store(on != 0);
    if (on) for (int i = 0; i < g_nslots; ++i) g_slots[i].used = 0;
    return B200_OK;
}

extern "C" int b200_set_option(const char* name, int value) {
    B200_CHECK_ARG(name != nullptr, "set_option: null name");
    if (strcmp(name, "tc_prefetch") == 0) { tc_set_prefetch(value); return B200_OK; }
    if (strcmp(name, "tc_dgrad") == 0) { tc_set_dgrad(value); return B200_OK; }
    if (strcmp(name, "tc_wgrad") == 0) { tc_set_wgrad(value); return B200_OK; }
    if (strcmp(name, "tc_fwd_eh") == 0) { tc_set_fwd_eh(value); return B200_OK; }
    if (strcmp(name, "wide_fuse") == 0) { wide_set_fuse(value); return B200_OK; }
    if (strcmp(name, "wide_images") == 0) { wide_set_images(value); return B200_OK; }
    if (strcmp(name, "wide_dbg") == 0) { wide2_set_dbg(value); return B200_OK; }
    if (strcmp(name, "wide_mc") == 0) { wide2_set_mc(value); return B200_OK; }
    set_error("set_option: unknown option '%s' (known: tc_prefetch, tc_fwd_eh, tc_wgrad, tc_dgrad, wide_fuse, wide_images)", name);
    return B200_ERR_INVALID;
}

extern "C" int b200_timing_report(char* buf, size_t buf_bytes) {
    B200_CHECK_ARG(buf && buf_bytes > 0, "timing_report: null buffer");
    std::lock_guard<std::mutex> lock(g_timing_mu);
    size_t off = 0;
    buf[0] = 0;
    for (int i = 0; i < g_nslots; ++i) {
        TimingSlot& s = g_slots[i];
        if (s.used == 0) continue;
        B200_CUDA(cudaEventSynchronize(s.ev[s.used - 1]));
        double total = 0.0;
        for (int k = 0; k < s.used; k += 2) {
            float ms = 0.f;
            B200_CUDA(cudaEventElapsedTime(&ms, s.ev[k], s.ev[k + 1]));
            total += ms;
        }
        int n = snprintf(buf + off, buf_bytes - off, "%s %d %.6f\n", s.name, s.used / 2, total);
        if (n < 0 || (size_t)n >= buf_bytes - off) break;
        off += (size_t)n;
    }
    return B200_OK;
}

extern "C" int64_t b200_param_offset(int F, int H, int L, int l) {
    if (F <= 0 || H <= 0 || L <= 0 || l < 0 || l > L) return -1;
    int64_t off = 0;
    for (int i = 0; i < l; ++i) {
        const int64_t in = i == 0 ? F : H;
        off += 4LL * H * in + 4LL * H * H + 8LL * H;
    }
    return off;
}

extern "C" int64_t b200_param_count(int F, int H, int L) {
    if (F <= 0 || H <= 0 || L <= 0) return -1;
    return b200_param_offset(F, H, L, L) + 2LL * H + 2;
}

static int check_lstm_args(const char* who, int mode, const float* params, int F, int H, int L, const float* x,
                           int64_t N, int T, const void* sh, const void* sc, float p) {
    B200_CHECK_ARG(mode >= 0 && mode <= 2, "%s: unknown mode %d", who, mode);
    B200_CHECK_ARG(params && x && sh && sc, "%s: null pointer", who);
    B200_CHECK_ARG(F > 0 && F <= 8 && H > 0 && L > 0 && L < 65536, "%s: bad model dims F=%d H=%d L=%d", who, F, H, L);
    B200_CHECK_ARG(N > 0 && T > 0 && T < 65536, "%s: bad batch dims N=%lld T=%d", who, (long long)N, T);
    B200_CHECK_ARG(p >= 0.f && p < 1.f, "%s: dropout must be in [0,1), got %f", who, p);
    return B200_OK;
}

extern "C" size_t b200_lstm_workspace_bytes(int mode, int64_t n_seq, int T, int n_features, int hidden, int layers) {
    if (n_seq <= 0 || T <= 0 || layers <= 0) return 0;
    // generous default: room for a 64 Ki-sequence da chunk; callers may pass less (smaller chunks)
    if (hidden != 64) return mode == B200_MODE_FFMA ? 0 : wide_workspace_bytes(mode, n_seq, T, n_features, hidden, layers);
    size_t ffma = ffma_workspace_bytes(n_seq, T, layers);
    if (mode == B200_MODE_FFMA) return ffma;
    size_t tc = tc_workspace_bytes(mode, n_seq, T, n_features, hidden, layers);
    return tc > ffma ? tc : ffma;
}

extern "C" int b200_lstm_fwd(int mode, const float* params, int n_features, int hidden, int layers, const float* x,
                             int64_t n_seq, int T, float* stash_h, float* stash_c, float dropout_p,
                             uint64_t dropout_seed, int training, void* workspace, size_t workspace_bytes, void* stream) {
    int rc = check_lstm_args("lstm_fwd", mode, params, n_features, hidden, layers, x, n_seq, T, stash_h, stash_c, dropout_p);
    if (rc) return rc;
    if (hidden != 64 && mode != B200_MODE_FFMA)
        return wide_fwd(mode, params, n_features, hidden, layers, x, n_seq, T, stash_h, stash_c, dropout_p, dropout_seed, training,
                        workspace, workspace_bytes, (cudaStream_t)stream);
    if (mode == B200_MODE_FFMA)
        return ffma_fwd(params, n_features, hidden, layers, x, n_seq, T, stash_h, stash_c, dropout_p, dropout_seed, training,
                        (cudaStream_t)stream);
    return tc_fwd(mode, params, n_features, hidden, layers, x, n_seq, T, stash_h, stash_c, dropout_p, dropout_seed, training,
                  workspace, workspace_bytes, (cudaStream_t)stream);
}

extern "C" int b200_lstm_bwd_layer(int mode, int layer, const float* params, int n_features, int hidden, int layers,
                                   const float* x, int64_t n_seq, int T, const float* stash_h, const float* stash_c,
                                   const float* dhT, float* grads, float dropout_p, uint64_t dropout_seed, int training,
                                   void* workspace, size_t workspace_bytes, void* stream) {
    int rc = check_lstm_args("lstm_bwd", mode, params, n_features, hidden, layers, x, n_seq, T, stash_h, stash_c, dropout_p);
    if (rc) return rc;
    B200_CHECK_ARG(layer >= 0 && layer < layers, "lstm_bwd: layer %d out of range", layer);
    B200_CHECK_ARG(dhT && grads && workspace, "lstm_bwd: null pointer");
    if (hidden != 64 && mode != B200_MODE_FFMA)
        return wide_bwd_layer(mode, layer, params, n_features, hidden, layers, x, n_seq, T, stash_h, stash_c, dhT, grads, dropout_p,
                              dropout_seed, training, workspace, workspace_bytes, (cudaStream_t)stream);
    if (mode == B200_MODE_FFMA)
        return ffma_bwd_layer(layer, params, n_features, hidden, layers, x, n_seq, T, stash_h, stash_c, dhT, grads, dropout_p,
                              dropout_seed, training, workspace, workspace_bytes, (cudaStream_t)stream);
    return tc_bwd_layer(mode, layer, params, n_features, hidden, layers, x, n_seq, T, stash_h, stash_c, dhT, grads, dropout_p,
                        dropout_seed, training, workspace, workspace_bytes, (cudaStream_t)stream);
}

extern "C" int b200_lstm_bwd(int mode, const float* params, int n_features, int hidden, int layers, const float* x,
                             int64_t n_seq, int T, const float* stash_h, const float* stash_c, const float* dhT,
                             float* grads, float dropout_p, uint64_t dropout_seed, int training, void* workspace,
                             size_t workspace_bytes, void* stream) {
    for (int l = layers - 1; l >= 0; --l) {
        int rc = b200_lstm_bwd_layer(mode, l, params, n_features, hidden, layers, x, n_seq, T, stash_h, stash_c, dhT, grads,
                                     dropout_p, dropout_seed, training, workspace, workspace_bytes, stream);
        if (rc) return rc;
    }
    return B200_OK;
}

extern "C" int b200_lstm_hidden_states(int mode, int layer, int n_features, int hidden, int layers, int64_t n_seq, int T,
                                       const float* stash_h, const void* workspace, size_t workspace_bytes, float* out, void* stream) {
    B200_CHECK_ARG(mode >= 0 && mode <= 2 && layer >= 0 && layer < layers && n_seq > 0 && T > 0 && out, "lstm_hidden_states: bad arguments");
    if (hidden == 64 && mode != B200_MODE_FFMA) {
        B200_CHECK_ARG(workspace != nullptr, "lstm_hidden_states: the tcgen05 engines keep the hidden states in the workspace");
        return tc_hidden_states(mode, layer, layers, n_seq, T, workspace, workspace_bytes, out, (cudaStream_t)stream);
    }
    if (hidden != 64 && wide_uses_images(mode, hidden, n_features)) {
        B200_CHECK_ARG(workspace != nullptr, "lstm_hidden_states: the bf16 wide engine keeps the hidden states in the workspace");
        return wide2_hidden_states(layer, n_features, hidden, layers, n_seq, T, workspace, workspace_bytes, out, (cudaStream_t)stream);
    }
    B200_CHECK_ARG(stash_h != nullptr, "lstm_hidden_states: null stash");
    const size_t lsz = (size_t)T * n_seq * hidden;
    B200_CUDA(cudaMemcpyAsync(out, stash_h + (size_t)layer * lsz, lsz * sizeof(float), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return B200_OK;
}
