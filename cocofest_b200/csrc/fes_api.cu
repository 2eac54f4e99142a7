// fes_api.cu -- the C ABI of include/fes_b200.h: argument checks, workspace carving, table build,
// dispatch to the (kind, fatigue, direction-set) kernel instantiations, host-buffer variants, probes.
#include <atomic>
#include <cstring>
#include <vector>

#include "fes_abi_common.hpp"

namespace fes {
#define FES_EXTERN(K, F)                                                                                   \
    extern template int launch_ivp_kf<double, K, F>(int, const IvpArgs<double> &, cudaStream_t);           \
    extern template int launch_shoot_kf<double, K, F>(int, const ShootArgs<double> &, cudaStream_t);       \
    extern template int launch_ivp_kf<float, K, F>(int, const IvpArgs<float> &, cudaStream_t);             \
    extern template int launch_shoot_kf<float, K, F>(int, const ShootArgs<float> &, cudaStream_t);
FES_EXTERN(0, false) FES_EXTERN(0, true) FES_EXTERN(1, false) FES_EXTERN(1, true) FES_EXTERN(2, false) FES_EXTERN(2, true)
#undef FES_EXTERN
#define FES_EXTERN_ID(K)                                                                      \
    extern template int launch_ident_k<double, K>(const IdentArgs<double> &, cudaStream_t);   \
    extern template int launch_ident_k<float, K>(const IdentArgs<float> &, cudaStream_t);
FES_EXTERN_ID(0) FES_EXTERN_ID(1) FES_EXTERN_ID(2)
#undef FES_EXTERN_ID
}  // namespace fes

namespace {

using namespace fes;
using namespace fes::abi;

std::atomic<int64_t> g_launches{0};

template <typename T>
int dispatch_ivp(int kind, int fat, int groups, const IvpArgs<T> &a, cudaStream_t s) {
    switch (kind * 2 + fat) {
        case 0: return launch_ivp_kf<T, 0, false>(groups, a, s);
        case 1: return launch_ivp_kf<T, 0, true>(groups, a, s);
        case 2: return launch_ivp_kf<T, 1, false>(groups, a, s);
        case 3: return launch_ivp_kf<T, 1, true>(groups, a, s);
        case 4: return launch_ivp_kf<T, 2, false>(groups, a, s);
        case 5: return launch_ivp_kf<T, 2, true>(groups, a, s);
    }
    return E_MODEL;
}
template <typename T>
int dispatch_shoot(int kind, int fat, int groups, const ShootArgs<T> &a, cudaStream_t s) {
    switch (kind * 2 + fat) {
        case 0: return launch_shoot_kf<T, 0, false>(groups, a, s);
        case 1: return launch_shoot_kf<T, 0, true>(groups, a, s);
        case 2: return launch_shoot_kf<T, 1, false>(groups, a, s);
        case 3: return launch_shoot_kf<T, 1, true>(groups, a, s);
        case 4: return launch_shoot_kf<T, 2, false>(groups, a, s);
        case 5: return launch_shoot_kf<T, 2, true>(groups, a, s);
    }
    return E_MODEL;
}

template <typename T>
int build_tables(const FesDesc *d, const double *win, void *ws, int64_t ws_bytes, const Carve &c, cudaStream_t s) {
    if (ws == nullptr || ws_bytes < (int64_t)c.total) return E_WORKSPACE;
    if (d->flags & FES_TABLES_READY) return E_OK;
    if (win == nullptr) return E_BUF;
    T *eps = reinterpret_cast<T *>((char *)ws + c.eps_off);
    T *tab = reinterpret_cast<T *>((char *)ws + c.tab_off);
    const int64_t total = std::max<int64_t>((int64_t)d->n_shooting * c.tB, stage_table_len(d->n_steps));
    const unsigned grid = (unsigned)((total + 127) / 128);
    const int64_t wB = (d->flags & FES_WIN_SHARED) ? 1 : d->batch, wS = (d->flags & FES_WIN_SHARED) ? 0 : 1;
    fes_tables_kernel<T><<<grid, 128, 0, s>>>(win, wB, wS, d->n_trunc, d->n_shooting, c.row, d->dt, d->t_start, d->tauc,
                                              tab, c.tB, d->n_steps, d->dt / d->n_steps, d->scheme, eps);
    ++g_launches;
    return (int)cudaGetLastError();
}

template <typename T>
int ivp_impl(const FesDesc *d, const FesIvpBuffers *b, void *stream) {
    int rc = check_desc(d, true);
    if (rc) return rc;
    if (b == nullptr || b->params == nullptr || b->x0 == nullptr || b->x_final == nullptr) return E_BUF;
    const int nu = num_controls(d->model_kind, d->n_trunc);
    if (nu > 0 && b->ctrl == nullptr) return E_BUF;
    if (d->trace_every != 0 && d->trace_batch > 0 && b->trace == nullptr && b->sens_trace == nullptr) return E_TRACE;
    cudaStream_t s = (cudaStream_t)stream;
    const Carve c = carve(d, sizeof(T));
    rc = build_tables<T>(d, b->stim_win, b->workspace, b->workspace_bytes, c, s);
    if (rc) return rc;
    IvpArgs<T> a;
    fill_ivp_args<T>(d, b, c, a);
    rc = dispatch_ivp<T>(d->model_kind, d->fatigue, d->dirs, a, s);
    if (rc == 0) ++g_launches;
    return rc;
}

template <typename T>
int shoot_impl(const FesDesc *d, const FesShootBuffers *b, void *stream) {
    int rc = check_desc(d, false);
    if (rc) return rc;
    if (b == nullptr || b->params == nullptr || b->X == nullptr) return E_BUF;
    const int nu = num_controls(d->model_kind, d->n_trunc);
    if (nu > 0 && b->ctrl == nullptr) return E_BUF;
    int groups = d->dirs;
    if (d->model_kind == FES_DING2003) groups &= ~FES_DIR_U;
    const int nd = num_dirs(d->model_kind, d->fatigue, groups);
    if (nd > 0 && b->J != nullptr && (b->j_col0 < 0 || b->j_col0 + nd > b->j_cols)) return E_SIZE;
    cudaStream_t s = (cudaStream_t)stream;
    const Carve c = carve(d, sizeof(T));
    rc = build_tables<T>(d, b->stim_win, b->workspace, b->workspace_bytes, c, s);
    if (rc) return rc;
    ShootArgs<T> a;
    fill_shoot_args<T>(d, b, c, a);
    rc = dispatch_shoot<T>(d->model_kind, d->fatigue, groups, a, s);
    if (rc == 0) ++g_launches;
    return rc;
}

template <typename T>
int ident_impl(const FesDesc *d, const FesIdentBuffers *b, const FesIdentOptions *o, void *stream) {
    int rc = check_ident(d, b, o);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const Carve c = carve(d, sizeof(T));
    rc = build_tables<T>(d, b->stim_win, b->workspace, b->workspace_bytes, c, s);
    if (rc) return rc;
    IdentArgs<T> a;
    fill_ident_args<T>(d, b, o, c, a);
    switch (d->model_kind) {
        case 0: rc = launch_ident_k<T, 0>(a, s); break;
        case 1: rc = launch_ident_k<T, 1>(a, s); break;
        default: rc = launch_ident_k<T, 2>(a, s); break;
    }
    if (rc == 0) ++g_launches;
    return rc;
}

// ---- host-buffer variants: stream-ordered temporaries, H2D, kernels, D2H --------------------
struct DevPool {
    cudaStream_t s;
    std::vector<void *> ptrs;
    int err = 0;
    explicit DevPool(cudaStream_t st) : s(st) {}
    void *alloc(size_t bytes) {
        if (err || bytes == 0) return nullptr;
        void *p = nullptr;
        err = (int)cudaMallocAsync(&p, bytes, s);
        if (!err) ptrs.push_back(p);
        return p;
    }
    void *upload(const void *h, size_t bytes) {
        if (h == nullptr) return nullptr;
        void *p = alloc(bytes);
        if (p && !err) err = (int)cudaMemcpyAsync(p, h, bytes, cudaMemcpyHostToDevice, s);
        return p;
    }
    void download(void *h, const void *d, size_t bytes) {
        if (h && d && !err) err = (int)cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, s);
    }
    ~DevPool() {
        for (void *p : ptrs) cudaFreeAsync(p, s);
    }
};

template <typename T>
int ivp_host_impl(const FesDesc *d, const FesIvpBuffers *h, void *stream) {
    int rc = check_desc(d, true);
    if (rc) return rc;
    if (h == nullptr || h->params == nullptr || h->x0 == nullptr || h->x_final == nullptr || h->stim_win == nullptr) return E_BUF;
    cudaStream_t s = (cudaStream_t)stream;
    const int nx = d->fatigue ? 5 : 2, nu = num_controls(d->model_kind, d->n_trunc);
    const int nd = num_dirs(d->model_kind, d->fatigue, d->dirs);
    const size_t B = (size_t)d->batch, ns = (size_t)d->n_shooting, N = (size_t)d->n_trunc, tB = (size_t)d->trace_batch;
    const size_t npts = d->trace_every == 1 ? ns * d->n_steps + 1 : ns + 1;
    FesDesc dd = *d;
    dd.flags &= ~FES_TABLES_READY;
    FesIvpBuffers g;
    std::memset(&g, 0, sizeof(g));
    {
        DevPool pool(s);
        g.params = pool.upload(h->params, sizeof(T) * FES_NP * ((d->flags & FES_PARAMS_SHARED) ? 1 : B));
        g.x0 = pool.upload(h->x0, sizeof(T) * nx * B);
        g.stim_win = (const double *)pool.upload(h->stim_win, sizeof(double) * ns * N * ((d->flags & FES_WIN_SHARED) ? 1 : B));
        if (nu) g.ctrl = pool.upload(h->ctrl, sizeof(T) * ns * nu * ((d->flags & FES_CTRL_SHARED) ? 1 : B));
        g.x_final = pool.alloc(sizeof(T) * nx * B);
        const bool tr = d->trace_every != 0 && tB > 0;
        if (tr && h->trace) g.trace = pool.alloc(sizeof(T) * npts * nx * tB);
        if (nd > 0 && h->sens_final) g.sens_final = pool.alloc(sizeof(T) * nx * nd * B);
        if (nd > 0 && tr && h->sens_trace) g.sens_trace = pool.alloc(sizeof(T) * (ns + 1) * nx * nd * tB);
        g.workspace_bytes = fes_workspace_bytes(&dd, sizeof(T) == 4);
        g.workspace = pool.alloc((size_t)g.workspace_bytes);
        if (pool.err) return pool.err;
        rc = ivp_impl<T>(&dd, &g, stream);
        if (rc) return rc;
        pool.download(h->x_final, g.x_final, sizeof(T) * nx * B);
        if (g.trace) pool.download(h->trace, g.trace, sizeof(T) * npts * nx * tB);
        if (g.sens_final) pool.download(h->sens_final, g.sens_final, sizeof(T) * nx * nd * B);
        if (g.sens_trace) pool.download(h->sens_trace, g.sens_trace, sizeof(T) * (ns + 1) * nx * nd * tB);
        if (pool.err) return pool.err;
    }
    return (int)cudaStreamSynchronize(s);
}

template <typename T>
int shoot_host_impl(const FesDesc *d, const FesShootBuffers *h, void *stream) {
    int rc = check_desc(d, false);
    if (rc) return rc;
    if (h == nullptr || h->params == nullptr || h->X == nullptr || h->stim_win == nullptr) return E_BUF;
    cudaStream_t s = (cudaStream_t)stream;
    const int nx = d->fatigue ? 5 : 2, nu = num_controls(d->model_kind, d->n_trunc);
    const size_t B = (size_t)d->batch, ns = (size_t)d->n_shooting, N = (size_t)d->n_trunc;
    FesDesc dd = *d;
    dd.flags &= ~FES_TABLES_READY;
    FesShootBuffers g = *h;
    {
        DevPool pool(s);
        g.params = pool.upload(h->params, sizeof(T) * FES_NP * ((d->flags & FES_PARAMS_SHARED) ? 1 : B));
        g.X = pool.upload(h->X, sizeof(T) * (ns + 1) * nx * B);
        g.stim_win = (const double *)pool.upload(h->stim_win, sizeof(double) * ns * N * ((d->flags & FES_WIN_SHARED) ? 1 : B));
        g.ctrl = nu ? pool.upload(h->ctrl, sizeof(T) * ns * nu * ((d->flags & FES_CTRL_SHARED) ? 1 : B)) : nullptr;
        g.G = h->G ? pool.alloc(sizeof(T) * ns * nx * B) : nullptr;
        g.Phi = h->Phi ? pool.alloc(sizeof(T) * ns * nx * B) : nullptr;
        const size_t jbytes = sizeof(T) * ns * nx * (size_t)h->j_cols * B;
        // a multi-pass caller fills J column ranges one call at a time: keep what is already there
        g.J = (h->J && d->dirs) ? pool.upload(h->J, jbytes) : nullptr;
        g.workspace_bytes = fes_workspace_bytes(&dd, sizeof(T) == 4);
        g.workspace = pool.alloc((size_t)g.workspace_bytes);
        if (pool.err) return pool.err;
        rc = shoot_impl<T>(&dd, &g, stream);
        if (rc) return rc;
        if (g.G) pool.download(h->G, g.G, sizeof(T) * ns * nx * B);
        if (g.Phi) pool.download(h->Phi, g.Phi, sizeof(T) * ns * nx * B);
        if (g.J) pool.download(h->J, g.J, jbytes);
        if (pool.err) return pool.err;
    }
    return (int)cudaStreamSynchronize(s);
}

template <typename T>
int ident_host_impl(const FesDesc *d, const FesIdentBuffers *h, const FesIdentOptions *o, void *stream) {
    int rc = check_ident(d, h, o);
    if (rc) return rc;
    if (h->stim_win == nullptr) return E_BUF;
    cudaStream_t s = (cudaStream_t)stream;
    const int nu = num_controls(d->model_kind, d->n_trunc), np = ident_num_params(d->model_kind);
    const size_t B = (size_t)d->batch, ns = (size_t)d->n_shooting, N = (size_t)d->n_trunc;
    FesDesc dd = *d;
    dd.flags &= ~FES_TABLES_READY;
    FesIdentBuffers g;
    std::memset(&g, 0, sizeof(g));
    {
        DevPool pool(s);
        g.params = pool.upload(h->params, sizeof(T) * FES_NP * ((d->flags & FES_PARAMS_SHARED) ? 1 : B));
        g.x0 = pool.upload(h->x0, sizeof(T) * 2 * B);
        g.stim_win = (const double *)pool.upload(h->stim_win, sizeof(double) * ns * N * ((d->flags & FES_WIN_SHARED) ? 1 : B));
        if (nu) g.ctrl = pool.upload(h->ctrl, sizeof(T) * ns * nu * ((d->flags & FES_CTRL_SHARED) ? 1 : B));
        g.target = pool.upload(h->target, sizeof(T) * (ns + 1) * ((d->flags & FES_TARGET_SHARED) ? 1 : B));
        g.weight = pool.upload(h->weight, sizeof(T) * (ns + 1));
        g.lower = pool.upload(h->lower, sizeof(T) * FES_NP);
        g.upper = pool.upload(h->upper, sizeof(T) * FES_NP);
        g.params_out = pool.alloc(sizeof(T) * FES_NP * B);
        g.cost = pool.alloc(sizeof(T) * B);
        g.grad = pool.alloc(sizeof(T) * np * B);
        g.hess = pool.alloc(sizeof(T) * (np * (np + 1) / 2) * B);
        if (h->iters) g.iters = (int32_t *)pool.alloc(sizeof(int32_t) * B);
        g.workspace_bytes = fes_workspace_bytes(&dd, sizeof(T) == 4);
        g.workspace = pool.alloc((size_t)g.workspace_bytes);
        if (pool.err) return pool.err;
        rc = ident_impl<T>(&dd, &g, o, stream);
        if (rc) return rc;
        pool.download(h->params_out, g.params_out, sizeof(T) * FES_NP * B);
        pool.download(h->cost, g.cost, sizeof(T) * B);
        pool.download(h->grad, g.grad, sizeof(T) * np * B);
        pool.download(h->hess, g.hess, sizeof(T) * (np * (np + 1) / 2) * B);
        if (g.iters) pool.download(h->iters, g.iters, sizeof(int32_t) * B);
        if (pool.err) return pool.err;
    }
    return (int)cudaStreamSynchronize(s);
}

// ---- FMA peak probe ------------------------------------------------------------------------------
// Register-only dependent FMA chains; several (chains per thread, unroll, block size) shapes are timed and the
// best rate is reported, so that the roofline denominator is the pipe's rate and not one loop's overhead.
template <typename T, int CHAINS, int UNROLL>
__global__ void fes_fma_probe_kernel(T *out, int iters, T m, T c) {
    T acc[CHAINS];
#pragma unroll
    for (int j = 0; j < CHAINS; ++j) acc[j] = (T)(threadIdx.x + j) * (T)1e-3;
    for (int it = 0; it < iters; it += UNROLL) {
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
#pragma unroll
            for (int j = 0; j < CHAINS; ++j) acc[j] = xfma(acc[j], m, c);
        }
    }
    T sum = 0;
#pragma unroll
    for (int j = 0; j < CHAINS; ++j) sum += acc[j];
    if (sum == (T)123456789) out[0] = sum;  // keeps the chains alive, practically never taken
}

template <typename T, int CHAINS, int UNROLL>
int probe_shape(int launches, int threads, int blocks_per_sm, int sms, T *out, cudaEvent_t e0, cudaEvent_t e1,
                cudaStream_t s, double *best) {
    const int iters = (sizeof(T) == 8 ? 65536 : 131072) / CHAINS;
    const unsigned grid = (unsigned)(sms * blocks_per_sm);
    int rc = 0;
    for (int l = 0; l < launches + 1; ++l) {
        cudaEventRecord(e0, s);
        fes_fma_probe_kernel<T, CHAINS, UNROLL><<<grid, threads, 0, s>>>(out, iters, (T)0.999999, (T)1e-7);
        cudaEventRecord(e1, s);
        rc = (int)cudaEventSynchronize(e1);
        ++g_launches;
        if (rc) return rc;
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flops = 2.0 * CHAINS * (double)iters * (double)threads * (double)grid;
        if (l > 0 && ms > 0.f) *best = std::max(*best, flops / (ms * 1e-3) / 1e12);
    }
    return (int)cudaGetLastError();
}

template <typename T>
int probe_impl(int launches, cudaStream_t s, double *tflops) {
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    T *out = nullptr;
    int rc = (int)cudaMalloc(&out, sizeof(T));
    if (rc) return rc;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    double best = 0.0;
    rc = probe_shape<T, 16, 1>(launches, 256, 8, sms, out, e0, e1, s, &best);
    if (!rc) rc = probe_shape<T, 16, 8>(launches, 256, 8, sms, out, e0, e1, s, &best);
    if (!rc) rc = probe_shape<T, 8, 16>(launches, 128, 16, sms, out, e0, e1, s, &best);
    if (!rc) rc = probe_shape<T, 32, 4>(launches, 128, 8, sms, out, e0, e1, s, &best);
    if (!rc) rc = probe_shape<T, 8, 16>(launches, 1024, 2, sms, out, e0, e1, s, &best);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    *tflops = best;
    return rc;
}

}  // namespace

extern "C" {

int fes_version(void) { return FES_ABI_VERSION; }

const char *fes_strerror(int code) {
    if (const char *m = strerror_abi(code)) return m;
    if (code > 0) return cudaGetErrorString((cudaError_t)code);
    return "unknown error";
}

int fes_num_states(int fatigue) { return fatigue ? 5 : 2; }
int fes_num_controls(int model_kind, int n_trunc) { return num_controls(model_kind, n_trunc); }
int fes_num_dirs(int model_kind, int fatigue, int dirs) {
    if (model_kind < 0 || model_kind > 2) return -1;
    if (model_kind == FES_DING2003 && dirs == (FES_DIR_X0 | FES_DIR_U)) dirs = FES_DIR_X0;
    return num_dirs(model_kind, fatigue ? 1 : 0, dirs);
}

int64_t fes_workspace_bytes(const FesDesc *desc, int is_f32) {
    if (desc == nullptr || desc->n_steps < 1 || desc->n_shooting < 1 || desc->n_trunc < 1 || desc->batch < 1) return -1;
    return (int64_t)carve(desc, is_f32 ? 4 : 8).total;
}

int fes_ivp_f64(const FesDesc *d, const FesIvpBuffers *b, void *s) { return ivp_impl<double>(d, b, s); }
int fes_ivp_f32(const FesDesc *d, const FesIvpBuffers *b, void *s) { return ivp_impl<float>(d, b, s); }
int fes_shoot_f64(const FesDesc *d, const FesShootBuffers *b, void *s) { return shoot_impl<double>(d, b, s); }
int fes_shoot_f32(const FesDesc *d, const FesShootBuffers *b, void *s) { return shoot_impl<float>(d, b, s); }
int fes_ivp_host_f64(const FesDesc *d, const FesIvpBuffers *b, void *s) { return ivp_host_impl<double>(d, b, s); }
int fes_ivp_host_f32(const FesDesc *d, const FesIvpBuffers *b, void *s) { return ivp_host_impl<float>(d, b, s); }
int fes_shoot_host_f64(const FesDesc *d, const FesShootBuffers *b, void *s) { return shoot_host_impl<double>(d, b, s); }
int fes_shoot_host_f32(const FesDesc *d, const FesShootBuffers *b, void *s) { return shoot_host_impl<float>(d, b, s); }

int fes_identify_f64(const FesDesc *d, const FesIdentBuffers *b, const FesIdentOptions *o, void *s) { return ident_impl<double>(d, b, o, s); }
int fes_identify_f32(const FesDesc *d, const FesIdentBuffers *b, const FesIdentOptions *o, void *s) { return ident_impl<float>(d, b, o, s); }
int fes_identify_host_f64(const FesDesc *d, const FesIdentBuffers *b, const FesIdentOptions *o, void *s) { return ident_host_impl<double>(d, b, o, s); }
int fes_identify_host_f32(const FesDesc *d, const FesIdentBuffers *b, const FesIdentOptions *o, void *s) { return ident_host_impl<float>(d, b, o, s); }

int fes_probe_fma_peak(int is_f32, int launches, void *stream, double *tflops_out) {
    if (tflops_out == nullptr || launches < 1) return E_BUF;
    return is_f32 ? probe_impl<float>(launches, (cudaStream_t)stream, tflops_out)
                  : probe_impl<double>(launches, (cudaStream_t)stream, tflops_out);
}

int64_t fes_launch_count(void) { return g_launches.load(); }

}  // extern "C"
