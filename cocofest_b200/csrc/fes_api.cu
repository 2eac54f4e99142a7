// fes_api.cu -- the C ABI of include/fes_b200.h: argument checks, workspace carving, table build,
// dispatch to the (kind, fatigue, direction-set) kernel instantiations, host-buffer variants, probes.
#include <atomic>
#include <chrono>
#include <cstring>
#include <mutex>
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
    const int nu = num_controls(d->model_kind, d->n_trunc) + ((d->dirs & FES_DIR_C) ? 1 : 0);
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

// ---- fes_shoot_hess_f64: second-order tangents of the interval map (fes_hess.cuh) ---------------------------------------
int hess_impl(const FesDesc *d, const FesHessBuffers *b, void *stream) {
    int rc = check_hess(d, b);
    if (rc) return rc;
    HessArgs a;
    fill_hess_args(d, b, a);
    rc = launch_hess(d->model_kind, d->fatigue, a, (cudaStream_t)stream);
    if (rc == 0) ++g_launches;
    return rc;
}

// ---- fes_shoot_static_f64: the interval kernel with its tables in library-owned static device memory ----------------------
// For callers that cannot allocate (stream capture, serving loops).  The scratch is a file-scope device array: every GPU that
// loads the library holds FES_STATIC_SCRATCH_BYTES of it until the process exits.  In host code the name `g_static_scratch` is
// only a handle for the runtime, NOT a device address: the address is asked for once per device, with that device current
// (cudaGetSymbolAddress), and that pointer is what the kernels get.  All calls on a device share the array.
static __device__ __align__(256) unsigned char g_static_scratch[FES_STATIC_SCRATCH_BYTES];

std::mutex g_scratch_mu;
void *g_scratch_ptr[64] = {};

int static_scratch(void **out) {
    int dev = 0;
    int rc = (int)cudaGetDevice(&dev);
    if (rc) return rc;
    if (dev < 0 || dev >= 64) return (int)cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_scratch_mu);
    if (g_scratch_ptr[dev] == nullptr) {
        void *ptr = nullptr;
        rc = (int)cudaGetSymbolAddress(&ptr, g_static_scratch);
        if (rc) return rc;
        if (ptr == nullptr) return (int)cudaErrorInvalidSymbol;
        g_scratch_ptr[dev] = ptr;
    }
    *out = g_scratch_ptr[dev];
    return E_OK;
}

int shoot_static_impl(const FesDesc *d, const FesShootBuffers *b, void *stream) {
    int rc = check_desc(d, false);
    if (rc) return rc;
    if (b == nullptr) return E_BUF;
    const Carve c = carve(d, sizeof(double));
    if ((int64_t)c.total > (int64_t)FES_STATIC_SCRATCH_BYTES) return E_STATIC_SCRATCH;
    void *ptr = nullptr;
    rc = static_scratch(&ptr);
    if (rc) return rc;
    FesShootBuffers g = *b;  // the caller's workspace fields are ignored
    g.workspace = ptr;
    g.workspace_bytes = FES_STATIC_SCRATCH_BYTES;
    return shoot_impl<double>(d, &g, stream);
}

// ---- host-buffer variants: stream-ordered temporaries, H2D, kernels, D2H --------------------
// The staging buffers come from a memory pool the LIBRARY owns (one per device, created on first use) whose release
// threshold is unbounded: memory freed by one call stays mapped for the next one.  With the device's default pool
// (threshold 0) every cudaStreamSynchronize handed the ~0.8 GB of a config-2 call back to the driver and the next
// call had to map it again -- round 1 measured 0.56 s of that per call on one box (e2e = 0.75 of the device rate).
// fes_host_pool_trim() gives the memory back.
constexpr int MAX_DEV = 64;
cudaMemPool_t g_pool[MAX_DEV] = {};
std::mutex g_pool_mu;

int host_pool(cudaMemPool_t *out) {
    int dev = 0;
    int rc = (int)cudaGetDevice(&dev);
    if (rc) return rc;
    if (dev < 0 || dev >= MAX_DEV) return (int)cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(g_pool_mu);
    if (g_pool[dev] == nullptr) {
        cudaMemPoolProps props;
        std::memset(&props, 0, sizeof(props));
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = dev;
        cudaMemPool_t pool = nullptr;
        rc = (int)cudaMemPoolCreate(&pool, &props);
        if (rc) return rc;
        uint64_t keep = UINT64_MAX;
        rc = (int)cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        if (rc) {
            cudaMemPoolDestroy(pool);
            return rc;
        }
        g_pool[dev] = pool;
    }
    *out = g_pool[dev];
    return 0;
}

// phases of the last host-buffer call of this thread, in milliseconds (fes_last_host_timing)
struct HostTiming {
    double alloc = 0, h2d = 0, kernels = 0, d2h = 0, total = 0;
};
thread_local HostTiming t_last;

struct DevPool {
    cudaStream_t s;
    cudaMemPool_t pool = nullptr;
    std::vector<void *> ptrs;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    std::chrono::steady_clock::time_point t0;
    int err = 0;
    explicit DevPool(cudaStream_t st) : s(st), t0(std::chrono::steady_clock::now()) {
        err = host_pool(&pool);
        for (int i = 0; i < 4 && !err; ++i) err = (int)cudaEventCreate(&ev[i]);
    }
    void *alloc(size_t bytes) {
        if (err || bytes == 0) return nullptr;
        void *p = nullptr;
        err = (int)cudaMallocFromPoolAsync(&p, bytes, pool, s);
        if (!err) ptrs.push_back(p);
        return p;
    }
    void upload_to(void *d, const void *h, size_t bytes) {
        if (d && h && !err) err = (int)cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, s);
    }
    void download(void *h, const void *d, size_t bytes) {
        if (h && d && !err) err = (int)cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, s);
    }
    void mark(int i) {
        if (!err) err = (int)cudaEventRecord(ev[i], s);
    }
    // frees (stream-ordered), waits for the stream, and records the phase times of this call
    int finish() {
        for (void *p : ptrs) cudaFreeAsync(p, s);
        ptrs.clear();
        int rc = err ? err : (int)cudaStreamSynchronize(s);
        if (!rc) {
            float a = 0.f, b = 0.f, c = 0.f;
            cudaEventElapsedTime(&a, ev[0], ev[1]);
            cudaEventElapsedTime(&b, ev[1], ev[2]);
            cudaEventElapsedTime(&c, ev[2], ev[3]);
            t_last.h2d = a;
            t_last.kernels = b;
            t_last.d2h = c;
            t_last.total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        }
        return rc;
    }
    ~DevPool() {
        for (void *p : ptrs) cudaFreeAsync(p, s);
        for (cudaEvent_t e : ev)
            if (e) cudaEventDestroy(e);
    }
};

// one staged array: allocation first (all of them, timed as `alloc`), then the copies
struct Staged {
    void *dev = nullptr;
    const void *host = nullptr;
    size_t bytes = 0;
};

template <typename T>
int ivp_host_impl(const FesDesc *d, const FesIvpBuffers *h, void *stream) {
    int rc = check_desc(d, true);
    if (rc) return rc;
    if (h == nullptr || h->params == nullptr || h->x0 == nullptr || h->x_final == nullptr || h->stim_win == nullptr) return E_BUF;
    cudaStream_t s = (cudaStream_t)stream;
    const int nx = d->fatigue ? 5 : 2, nu = num_controls(d->model_kind, d->n_trunc);
    if (nu > 0 && h->ctrl == nullptr) return E_BUF;
    const int nd = num_dirs(d->model_kind, d->fatigue, d->dirs);
    const size_t B = (size_t)d->batch, ns = (size_t)d->n_shooting, N = (size_t)d->n_trunc, tB = (size_t)d->trace_batch;
    const size_t npts = d->trace_every == 1 ? ns * d->n_steps + 1 : ns + 1;
    FesDesc dd = *d;
    dd.flags &= ~FES_TABLES_READY;
    FesIvpBuffers g;
    std::memset(&g, 0, sizeof(g));
    DevPool pool(s);
    const size_t b_par = sizeof(T) * FES_NP * ((d->flags & FES_PARAMS_SHARED) ? 1 : B), b_x = sizeof(T) * nx * B;
    const size_t b_win = sizeof(double) * ns * N * ((d->flags & FES_WIN_SHARED) ? 1 : B);
    const size_t b_ctl = nu ? sizeof(T) * ns * nu * ((d->flags & FES_CTRL_SHARED) ? 1 : B) : 0;
    const bool tr = d->trace_every != 0 && tB > 0;
    const size_t b_tr = (tr && h->trace) ? sizeof(T) * npts * nx * tB : 0;
    const size_t b_sf = (nd > 0 && h->sens_final) ? sizeof(T) * nx * nd * B : 0;
    const size_t b_st = (nd > 0 && tr && h->sens_trace) ? sizeof(T) * (ns + 1) * nx * nd * tB : 0;
    void *params = pool.alloc(b_par), *x0 = pool.alloc(b_x), *win = pool.alloc(b_win), *ctl = pool.alloc(b_ctl);
    g.x_final = pool.alloc(b_x);
    g.trace = pool.alloc(b_tr);
    g.sens_final = pool.alloc(b_sf);
    g.sens_trace = pool.alloc(b_st);
    g.workspace_bytes = fes_workspace_bytes(&dd, sizeof(T) == 4);
    g.workspace = pool.alloc((size_t)g.workspace_bytes);
    t_last = HostTiming();
    t_last.alloc = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - pool.t0).count();
    pool.mark(0);
    pool.upload_to(params, h->params, b_par);
    pool.upload_to(x0, h->x0, b_x);
    pool.upload_to(win, h->stim_win, b_win);
    pool.upload_to(ctl, h->ctrl, b_ctl);
    pool.mark(1);
    if (pool.err) return pool.err;
    g.params = params;
    g.x0 = x0;
    g.stim_win = (const double *)win;
    g.ctrl = ctl;
    rc = ivp_impl<T>(&dd, &g, stream);
    if (rc) return rc;
    pool.mark(2);
    pool.download(h->x_final, g.x_final, b_x);
    pool.download(h->trace, g.trace, b_tr);
    pool.download(h->sens_final, g.sens_final, b_sf);
    pool.download(h->sens_trace, g.sens_trace, b_st);
    pool.mark(3);
    return pool.finish();
}

template <typename T>
int shoot_host_impl(const FesDesc *d, const FesShootBuffers *h, void *stream) {
    int rc = check_desc(d, false);
    if (rc) return rc;
    if (h == nullptr || h->params == nullptr || h->X == nullptr || h->stim_win == nullptr) return E_BUF;
    cudaStream_t s = (cudaStream_t)stream;
    const int nx = d->fatigue ? 5 : 2, nu = num_controls(d->model_kind, d->n_trunc) + ((d->dirs & FES_DIR_C) ? 1 : 0);
    if (nu > 0 && h->ctrl == nullptr) return E_BUF;
    const size_t B = (size_t)d->batch, ns = (size_t)d->n_shooting, N = (size_t)d->n_trunc;
    int groups = d->dirs;
    if (d->model_kind == FES_DING2003) groups &= ~FES_DIR_U;
    const int nd = num_dirs(d->model_kind, d->fatigue, groups);
    FesDesc dd = *d;
    dd.flags &= ~FES_TABLES_READY;
    FesShootBuffers g = *h;
    DevPool pool(s);
    const size_t b_par = sizeof(T) * FES_NP * ((d->flags & FES_PARAMS_SHARED) ? 1 : B), b_X = sizeof(T) * (ns + 1) * nx * B;
    const size_t b_win = sizeof(double) * ns * N * ((d->flags & FES_WIN_SHARED) ? 1 : B);
    const size_t b_ctl = nu ? sizeof(T) * ns * nu * ((d->flags & FES_CTRL_SHARED) ? 1 : B) : 0;
    const size_t b_g = sizeof(T) * ns * nx * B;
    const size_t b_J = (h->J && nd > 0) ? sizeof(T) * ns * nx * (size_t)h->j_cols * B : 0;
    void *params = pool.alloc(b_par), *X = pool.alloc(b_X), *win = pool.alloc(b_win), *ctl = pool.alloc(b_ctl);
    g.G = h->G ? pool.alloc(b_g) : nullptr;
    g.Phi = h->Phi ? pool.alloc(b_g) : nullptr;
    g.J = pool.alloc(b_J);
    g.workspace_bytes = fes_workspace_bytes(&dd, sizeof(T) == 4);
    g.workspace = pool.alloc((size_t)g.workspace_bytes);
    t_last = HostTiming();
    t_last.alloc = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - pool.t0).count();
    pool.mark(0);
    pool.upload_to(params, h->params, b_par);
    pool.upload_to(X, h->X, b_X);
    pool.upload_to(win, h->stim_win, b_win);
    pool.upload_to(ctl, h->ctrl, b_ctl);
    // a multi-pass caller fills J column ranges one call at a time: what the other passes wrote is kept.  A call that
    // fills every column of J has nothing to keep and uploads nothing.  (cocofest_b200.fes.shoot stages host tensors on
    // the device ONCE and runs all passes there; this path serves single-pass C callers.)
    if (b_J && nd < h->j_cols) pool.upload_to(g.J, h->J, b_J);
    pool.mark(1);
    if (pool.err) return pool.err;
    g.params = params;
    g.X = X;
    g.stim_win = (const double *)win;
    g.ctrl = ctl;
    rc = shoot_impl<T>(&dd, &g, stream);
    if (rc) return rc;
    pool.mark(2);
    pool.download(h->G, g.G, b_g);
    pool.download(h->Phi, g.Phi, b_g);
    pool.download(h->J, g.J, b_J);
    pool.mark(3);
    return pool.finish();
}

int hess_host_impl(const FesDesc *d, const FesHessBuffers *h, void *stream) {
    int rc = check_hess(d, h);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const size_t nx = d->fatigue ? 5 : 2, nur = num_controls(d->model_kind, d->n_trunc) + ((d->dirs & FES_DIR_C) ? 1 : 0);
    const size_t B = (size_t)d->batch, ns = (size_t)d->n_shooting, N = (size_t)d->n_trunc;
    const size_t nz = (size_t)hess_num_dirs(d), npairs = nz * (nz + 1) / 2;
    DevPool pool(s);
    Staged in[5] = {
        {nullptr, h->params, sizeof(double) * FES_NP * ((d->flags & FES_PARAMS_SHARED) ? 1 : B)},
        {nullptr, h->X, sizeof(double) * (ns + 1) * nx * B},
        {nullptr, h->stim_win, sizeof(double) * ns * N * ((d->flags & FES_WIN_SHARED) ? 1 : B)},
        {nullptr, h->ctrl, nur ? sizeof(double) * ns * nur * ((d->flags & FES_CTRL_SHARED) ? 1 : B) : 0},
        {nullptr, h->multipliers, (h->H && h->multipliers) ? sizeof(double) * ns * nx * B : 0},
    };
    for (Staged &st : in) st.dev = pool.alloc(st.bytes);
    const size_t b_d2 = h->D2 ? sizeof(double) * ns * nx * npairs * B : 0, b_h = h->H ? sizeof(double) * ns * npairs * B : 0;
    FesHessBuffers g;
    std::memset(&g, 0, sizeof(g));
    g.D2 = pool.alloc(b_d2);
    g.H = pool.alloc(b_h);
    t_last = HostTiming();
    t_last.alloc = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - pool.t0).count();
    pool.mark(0);
    for (Staged &st : in) pool.upload_to(st.dev, st.host, st.bytes);
    pool.mark(1);
    if (pool.err) return pool.err;
    g.params = in[0].dev;
    g.X = in[1].dev;
    g.stim_win = (const double *)in[2].dev;
    g.ctrl = in[3].dev;
    g.multipliers = in[4].dev;
    rc = hess_impl(d, &g, stream);
    if (rc) return rc;
    pool.mark(2);
    pool.download(h->D2, g.D2, b_d2);
    pool.download(h->H, g.H, b_h);
    pool.mark(3);
    return pool.finish();
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
    DevPool pool(s);
    Staged in[8] = {
        {nullptr, h->params, sizeof(T) * FES_NP * ((d->flags & FES_PARAMS_SHARED) ? 1 : B)},
        {nullptr, h->x0, h->x0 ? sizeof(T) * 2 * B : 0},
        {nullptr, h->stim_win, sizeof(double) * ns * N * ((d->flags & FES_WIN_SHARED) ? 1 : B)},
        {nullptr, h->ctrl, nu ? sizeof(T) * ns * nu * ((d->flags & FES_CTRL_SHARED) ? 1 : B) : 0},
        {nullptr, h->target, sizeof(T) * (ns + 1) * ((d->flags & FES_TARGET_SHARED) ? 1 : B)},
        {nullptr, h->weight, h->weight ? sizeof(T) * (ns + 1) : 0},
        {nullptr, h->lower, h->lower ? sizeof(T) * FES_NP : 0},
        {nullptr, h->upper, h->upper ? sizeof(T) * FES_NP : 0},
    };
    for (Staged &st : in) st.dev = pool.alloc(st.bytes);
    const size_t b_p = sizeof(T) * FES_NP * B, b_c = sizeof(T) * B, b_g = sizeof(T) * np * B;
    const size_t b_h = sizeof(T) * (np * (np + 1) / 2) * B, b_i = h->iters ? sizeof(int32_t) * B : 0;
    g.params_out = pool.alloc(b_p);
    g.cost = pool.alloc(b_c);
    g.grad = pool.alloc(b_g);
    g.hess = pool.alloc(b_h);
    g.iters = (int32_t *)pool.alloc(b_i);
    g.workspace_bytes = fes_workspace_bytes(&dd, sizeof(T) == 4);
    g.workspace = pool.alloc((size_t)g.workspace_bytes);
    t_last = HostTiming();
    t_last.alloc = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - pool.t0).count();
    pool.mark(0);
    for (Staged &st : in) pool.upload_to(st.dev, st.host, st.bytes);
    pool.mark(1);
    if (pool.err) return pool.err;
    g.params = in[0].dev;
    g.x0 = in[1].dev;
    g.stim_win = (const double *)in[2].dev;
    g.ctrl = in[3].dev;
    g.target = in[4].dev;
    g.weight = in[5].dev;
    g.lower = in[6].dev;
    g.upper = in[7].dev;
    rc = ident_impl<T>(&dd, &g, o, stream);
    if (rc) return rc;
    pool.mark(2);
    pool.download(h->params_out, g.params_out, b_p);
    pool.download(h->cost, g.cost, b_c);
    pool.download(h->grad, g.grad, b_g);
    pool.download(h->hess, g.hess, b_h);
    pool.download(h->iters, g.iters, b_i);
    pool.mark(3);
    return pool.finish();
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
int fes_shoot_static_f64(const FesDesc *d, const FesShootBuffers *b, void *s) { return shoot_static_impl(d, b, s); }
int fes_shoot_hess_num_dirs(int model_kind, int fatigue, int n_trunc, int dirs) {
    if (model_kind < 0 || model_kind > 2 || n_trunc < 1) return -1;
    return (fatigue ? 5 : 2) + num_controls(model_kind, n_trunc) + ((dirs & FES_DIR_C) ? 1 : 0);
}
int fes_shoot_hess_f64(const FesDesc *d, const FesHessBuffers *b, void *s) { return hess_impl(d, b, s); }
int fes_shoot_hess_host_f64(const FesDesc *d, const FesHessBuffers *b, void *s) { return hess_host_impl(d, b, s); }
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

int fes_last_host_timing(double *ms5) {
    if (ms5 == nullptr) return E_BUF;
    ms5[0] = t_last.alloc;
    ms5[1] = t_last.h2d;
    ms5[2] = t_last.kernels;
    ms5[3] = t_last.d2h;
    ms5[4] = t_last.total;
    return E_OK;
}

int fes_host_pool_trim(void) {
    int dev = 0;
    int rc = (int)cudaGetDevice(&dev);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(g_pool_mu);
    if (dev < 0 || dev >= MAX_DEV || g_pool[dev] == nullptr) return E_OK;
    return (int)cudaMemPoolTrimTo(g_pool[dev], 0);
}

}  // extern "C"
