// fes_emul.cpp -- HOST BUILD OF THE KERNEL ARITHMETIC, TEST INFRASTRUCTURE ONLY.
//
// The build container has no GPU.  To let the CPU-only CI exercise the same templates the CUDA kernels
// are made of (cocofest_b200/csrc/fes_core.cuh: interval_setup, rk_step, the tangent columns) and the
// same argument checks / workspace carving (fes_abi_common.hpp), this file compiles them with g++ and
// drives them with plain loops over (trajectory, node), exporting the fes_* symbols of
// include/fes_b200.h with HOST pointers.  Nothing under cocofest_b200/ references this library; only
// tests/ load it, explicitly, through the `emul_lib` fixture.  It is not a product CPU path.
#include <cmath>
#include <cstdint>
#include <vector>

#include "fes_abi_common.hpp"

using namespace fes;
using namespace fes::abi;

namespace {

template <typename T>
void tables_cpu(const FesDesc *d, const double *win, const Carve &c, void *ws) {
    T *eps = reinterpret_cast<T *>((char *)ws + c.eps_off);
    T *tab = reinterpret_cast<T *>((char *)ws + c.tab_off);
    const int S = d->n_steps, N = d->n_trunc, ns = d->n_shooting, row = c.row;
    const double h = d->dt / S, tauc = d->tauc;
    const int64_t tB = c.tB;
    const int64_t wB = (d->flags & FES_WIN_SHARED) ? 1 : d->batch, wS = (d->flags & FES_WIN_SHARED) ? 0 : 1;
    for (int i = 0; i < stage_table_len(S); ++i) eps[i] = (T)stage_table_entry(i, S, h, tauc, d->scheme);
    for (int64_t node = 0; node < ns; ++node)
        for (int64_t b = 0; b < tB; ++b) {
            const double tn = d->t_start + d->dt * (double)node;
            double a = 0.0, bsum = 0.0, prev = 0.0;
            for (int i = 0; i < N; ++i) {
                const double ti = win[(node * N + i) * wB + b * wS];
                const double wa = std::exp(-(tn - ti) / tauc);
                const double wb = (i == 0) ? 0.0 : std::exp(-(ti - prev) / tauc) * wa;
                prev = ti;
                a += wa;
                bsum += wb;
                if (row > 2) {
                    tab[(node * row + 2 + i) * tB + b] = (T)wa;
                    tab[(node * row + 2 + N + i) * tB + b] = (T)wb;
                }
            }
            tab[(node * row + 0) * tB + b] = (T)a;
            tab[(node * row + 1) * tB + b] = (T)bsum;
        }
}

template <typename T, int KIND, bool FAT, int GROUPS>
int ivp_cpu(const IvpArgs<T> &a) {
    using D = Dirs<KIND, FAT, GROUPS>;
    constexpr int NX = D::NX, ND = D::ND;
    // OpenMP only in the TIMED host build that bench.py compiles for the `cpu_baseline.hoisted` row (-fopenmp); the test
    // builds are serial
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < a.B; ++b) {
        Traj<T> tr;
        load_traj(tr, [&](int j) { return a.params[j * a.pB + b * a.pS]; }, a.tauc, a.r0_km);
        T x[NX];
        for (int j = 0; j < NX; ++j) x[j] = a.x0[j * a.B + b];
        Tangent<T, KIND, FAT, GROUPS> tg;
        tg.init(x);
        const bool tr_on = a.trace_mode != 0 && b < a.trace_B;
        auto put_state = [&](int64_t pt) {
            if (a.trace) for (int j = 0; j < NX; ++j) a.trace[(pt * NX + j) * a.trace_B + b] = x[j];
        };
        auto put_sens = [&](int64_t node) {
            if (ND > 0 && a.sens_trace)
                static_for<0, ND>([&](auto ic) {
                    constexpr int d = decltype(ic)::value;
                    for (int j = 0; j < NX; ++j)
                        a.sens_trace[((node * NX + j) * ND + d) * a.trace_B + b] = tg.template get<d>(tr, j);
                });
        };
        if (tr_on) { put_state(0); put_sens(0); }
        LamCache<T> lc;  // as fes_ivp_kernel: tanh of the previous interval's window slots (Hmed2018)
        LamCache<T> *lcp = nullptr;
        if (KIND == HMED2018) { lc.reset(a.N); lcp = &lc; }
        for (int node = 0; node < a.ns; ++node) {
            const T *trow = a.tab + (int64_t)node * a.row * a.tB + b * a.tS;
            const T *crow = a.ctrl + (int64_t)node * a.nu * a.cB + b * a.cS;
            Interval<T, ND> iv;
            interval_setup<T, KIND, FAT, GROUPS>(
                tr, [&](int c) { return trow[c * a.tB]; }, [&](int c) { return crow[c * a.cB]; }, a.N, a.tile_off, iv, lcp);
            for (int i = 0; i < a.S; ++i) {
                rk_step<T, KIND, FAT, GROUPS>(tr, iv, a.scheme, a.h, a.eps[2 * i], a.eps[2 * i + 1], a.eps[2 * i + 2],
                                              a.eps + 2 * a.S + 1 + 4 * i, x, tg);
                if (tr_on && a.trace_mode == 1) put_state((int64_t)node * a.S + i + 1);
            }
            if (tr_on) {
                if (a.trace_mode == 2) put_state(node + 1);
                put_sens(node + 1);
            }
        }
        for (int j = 0; j < NX; ++j) a.x_final[j * a.B + b] = x[j];
        if (ND > 0 && a.sens_final)
            static_for<0, ND>([&](auto ic) {
                constexpr int d = decltype(ic)::value;
                for (int j = 0; j < NX; ++j) a.sens_final[(j * ND + d) * a.B + b] = tg.template get<d>(tr, j);
            });
    }
    return 0;
}

template <typename T, int KIND, bool FAT, int GROUPS>
int shoot_cpu(const ShootArgs<T> &a) {
    using D = Dirs<KIND, FAT, GROUPS>;
    constexpr int NX = D::NX, ND = D::ND;
    for (int64_t k = 0; k < a.ns; ++k)
        for (int64_t b = 0; b < a.B; ++b) {
            Traj<T> tr;
            load_traj(tr, [&](int j) { return a.params[j * a.pB + b * a.pS]; }, a.tauc, a.r0_km);
            T x[NX];
            for (int j = 0; j < NX; ++j) x[j] = a.X[(k * NX + j) * a.B + b];
            Tangent<T, KIND, FAT, GROUPS> tg;
            tg.init(x);
            const T *trow = a.tab + k * a.row * a.tB + b * a.tS;
            const T *crow = a.ctrl + k * a.nu * a.cB + b * a.cS;
            Interval<T, ND> iv;
            interval_setup<T, KIND, FAT, GROUPS>(
                tr, [&](int c) { return trow[c * a.tB]; }, [&](int c) { return crow[c * a.cB]; }, a.N, a.tile_off, iv);
            for (int i = 0; i < a.S; ++i)
                rk_step<T, KIND, FAT, GROUPS>(tr, iv, a.scheme, a.h, a.eps[2 * i], a.eps[2 * i + 1], a.eps[2 * i + 2],
                                              a.eps + 2 * a.S + 1 + 4 * i, x, tg);
            for (int j = 0; j < NX; ++j) {
                if (a.Phi) a.Phi[(k * NX + j) * a.B + b] = x[j];
                if (a.G) a.G[(k * NX + j) * a.B + b] = a.X[((k + 1) * NX + j) * a.B + b] - x[j];
            }
            if (ND > 0 && a.J)
                static_for<0, ND>([&](auto ic) {
                    constexpr int d = decltype(ic)::value;
                    for (int j = 0; j < NX; ++j)
                        a.J[((k * NX + j) * a.j_cols + a.j_col0 + d) * a.B + b] = tg.template get<d>(tr, j);
                });
        }
    return 0;
}

template <typename T, int KIND, bool FAT>
int ivp_kf(int groups, const IvpArgs<T> &a) {
    switch (groups) {
        case 0: return ivp_cpu<T, KIND, FAT, 0>(a);
        case G_X0: return ivp_cpu<T, KIND, FAT, G_X0>(a);
        case G_P: return ivp_cpu<T, KIND, FAT, G_P>(a);
        case G_P | G_PF:
            if constexpr (FAT) return ivp_cpu<T, KIND, FAT, G_P | G_PF>(a);
            else return E_DIRS;
    }
    return E_DIRS;
}
template <typename T, int KIND, bool FAT>
int shoot_kf(int groups, const ShootArgs<T> &a) {
    switch (groups) {
        case 0: return shoot_cpu<T, KIND, FAT, 0>(a);
        case G_X0: return shoot_cpu<T, KIND, FAT, G_X0>(a);
        case G_X0 | G_U:
            if constexpr (KIND != DING2003) return shoot_cpu<T, KIND, FAT, G_X0 | G_U>(a);
            else return shoot_cpu<T, KIND, FAT, G_X0>(a);
        case G_U:
            if constexpr (KIND != DING2003) return shoot_cpu<T, KIND, FAT, G_U>(a);
            else return E_DIRS;
        case G_X0 | G_U | G_C:
            if constexpr (KIND != DING2003) return shoot_cpu<T, KIND, FAT, G_X0 | G_U | G_C>(a);
            else return shoot_cpu<T, KIND, FAT, G_X0 | G_C>(a);
        case G_X0 | G_C:
            if constexpr (KIND == DING2003) return shoot_cpu<T, KIND, FAT, G_X0 | G_C>(a);
            else return E_DIRS;
        case G_U | G_C:  // Hmed2018: the intensity tiles after the first of a call that carries the force factor
            if constexpr (KIND == HMED2018) return shoot_cpu<T, KIND, FAT, G_U | G_C>(a);
            else return E_DIRS;
        case G_P: return shoot_cpu<T, KIND, FAT, G_P>(a);
        case G_P | G_PF:
            if constexpr (FAT) return shoot_cpu<T, KIND, FAT, G_P | G_PF>(a);
            else return E_DIRS;
        case G_W: return shoot_cpu<T, KIND, FAT, G_W>(a);
    }
    return E_DIRS;
}
#define KF_SWITCH(fn, ...)                                    \
    switch (kind * 2 + fat) {                                 \
        case 0: return fn<T, 0, false>(__VA_ARGS__);          \
        case 1: return fn<T, 0, true>(__VA_ARGS__);           \
        case 2: return fn<T, 1, false>(__VA_ARGS__);          \
        case 3: return fn<T, 1, true>(__VA_ARGS__);           \
        case 4: return fn<T, 2, false>(__VA_ARGS__);          \
        case 5: return fn<T, 2, true>(__VA_ARGS__);           \
    }                                                         \
    return E_MODEL;
template <typename T> int disp_ivp(int kind, int fat, int groups, const IvpArgs<T> &a) { KF_SWITCH(ivp_kf, groups, a) }
template <typename T> int disp_shoot(int kind, int fat, int groups, const ShootArgs<T> &a) { KF_SWITCH(shoot_kf, groups, a) }

template <typename T>
int ivp_impl(const FesDesc *d, const FesIvpBuffers *b) {
    int rc = check_desc(d, true);
    if (rc) return rc;
    if (!b || !b->params || !b->x0 || !b->x_final) return E_BUF;
    if (num_controls(d->model_kind, d->n_trunc) > 0 && !b->ctrl) return E_BUF;
    if (d->trace_every != 0 && d->trace_batch > 0 && !b->trace && !b->sens_trace) return E_TRACE;
    const Carve c = carve(d, sizeof(T));
    if (!b->workspace || b->workspace_bytes < (int64_t)c.total) return E_WORKSPACE;
    if (!(d->flags & FES_TABLES_READY)) {
        if (!b->stim_win) return E_BUF;
        tables_cpu<T>(d, b->stim_win, c, b->workspace);
    }
    IvpArgs<T> a;
    fill_ivp_args<T>(d, b, c, a);
    return disp_ivp<T>(d->model_kind, d->fatigue, d->dirs, a);
}
template <typename T>
int shoot_impl(const FesDesc *d, const FesShootBuffers *b) {
    int rc = check_desc(d, false);
    if (rc) return rc;
    if (!b || !b->params || !b->X) return E_BUF;
    if (num_controls(d->model_kind, d->n_trunc) + ((d->dirs & FES_DIR_C) ? 1 : 0) > 0 && !b->ctrl) return E_BUF;
    int groups = d->dirs;
    if (d->model_kind == FES_DING2003) groups &= ~FES_DIR_U;
    const int nd = num_dirs(d->model_kind, d->fatigue, groups);
    if (nd > 0 && b->J && (b->j_col0 < 0 || b->j_col0 + nd > b->j_cols)) return E_SIZE;
    const Carve c = carve(d, sizeof(T));
    if (!b->workspace || b->workspace_bytes < (int64_t)c.total) return E_WORKSPACE;
    if (!(d->flags & FES_TABLES_READY)) {
        if (!b->stim_win) return E_BUF;
        tables_cpu<T>(d, b->stim_win, c, b->workspace);
    }
    ShootArgs<T> a;
    fill_shoot_args<T>(d, b, c, a);
    return disp_shoot<T>(d->model_kind, d->fatigue, groups, a);
}
template <typename T>
int ident_impl(const FesDesc *d, const FesIdentBuffers *b, const FesIdentOptions *o) {
    int rc = check_ident(d, b, o);
    if (rc) return rc;
    const Carve c = carve(d, sizeof(T));
    if (!b->workspace || b->workspace_bytes < (int64_t)c.total) return E_WORKSPACE;
    if (!(d->flags & FES_TABLES_READY)) {
        if (!b->stim_win) return E_BUF;
        tables_cpu<T>(d, b->stim_win, c, b->workspace);
    }
    IdentArgs<T> a;
    fill_ident_args<T>(d, b, o, c, a);
    T acc[8 + 36 + FES_NP];  // per-candidate scratch (ident_scratch<KIND>(), shared memory on the device)
    for (int64_t i = 0; i < a.B; ++i) {
        switch (d->model_kind) {
            case 0: ident_candidate<T, 0>(a, ident_rows_strided(a, i), i, a.eps, acc, 1); break;
            case 1: ident_candidate<T, 1>(a, ident_rows_strided(a, i), i, a.eps, acc, 1); break;
            default:
                // tables and controls without batch axis: the dense unit-stride row source the CUDA kernel stages in
                // shared memory (fes_ident_kernel<.., STAGED>), read here straight from the workspace
                if (a.tS == 0 && a.cS == 0) {
                    const IdentRows<T, true> rows = {a.tab, a.ctrl, a.row, 1, a.nu, 1};
                    ident_candidate<T, 2, true>(a, rows, i, a.eps, acc, 1);
                } else {
                    ident_candidate<T, 2>(a, ident_rows_strided(a, i), i, a.eps, acc, 1);
                }
                break;
        }
    }
    return 0;
}
template <typename T>
int ident_host(const FesDesc *d, const FesIdentBuffers *b, const FesIdentOptions *o) {
    if (!d || !b) return E_DESC;
    FesDesc dd = *d;
    dd.flags &= ~FES_TABLES_READY;
    const int64_t wb = fes_workspace_bytes(&dd, sizeof(T) == 4);
    if (wb < 0) return E_SIZE;
    std::vector<char> ws((size_t)wb + 16);
    FesIdentBuffers bb = *b;
    bb.workspace = ws.data();
    bb.workspace_bytes = wb;
    return ident_impl<T>(&dd, &bb, o);
}
}  // namespace

// second-order tangents: the same hess_unit template, plain loops over (interval, pair, scenario)
template <int KIND, bool FAT>
static void hess_cpu_kf(const HessArgs &a) {
    std::vector<double> eps_v(2 * a.S + 1);
    for (int j = 0; j <= 2 * a.S; ++j) eps_v[j] = hess_eps_entry(a, j);
    const double *eps = eps_v.data();
    for (int k = 0; k < a.ns; ++k)
        for (int p = 0; p < a.npairs; ++p)
            for (int64_t b = 0; b < a.B; ++b) {
                if (a.factor) hess_store<KIND, FAT, true>(a, eps, k, p, b);
                else hess_store<KIND, FAT, false>(a, eps, k, p, b);
            }
}

extern "C" {
int fes_version(void) { return FES_ABI_VERSION; }
const char *fes_strerror(int code) { const char *m = strerror_abi(code); return m ? m : "unknown error"; }
int fes_num_states(int fatigue) { return fatigue ? 5 : 2; }
int fes_num_controls(int k, int n) { return num_controls(k, n); }
int fes_num_dirs(int k, int f, int dirs) {
    if (k < 0 || k > 2) return -1;
    if (k == FES_DING2003 && dirs == (FES_DIR_X0 | FES_DIR_U)) dirs = FES_DIR_X0;
    return num_dirs(k, f ? 1 : 0, dirs);
}
int64_t fes_workspace_bytes(const FesDesc *d, int is_f32) {
    if (!d || d->n_steps < 1 || d->n_shooting < 1 || d->n_trunc < 1 || d->batch < 1) return -1;
    return (int64_t)carve(d, is_f32 ? 4 : 8).total;
}
int fes_ivp_f64(const FesDesc *d, const FesIvpBuffers *b, void *) { return ivp_impl<double>(d, b); }
int fes_ivp_f32(const FesDesc *d, const FesIvpBuffers *b, void *) { return ivp_impl<float>(d, b); }
int fes_shoot_f64(const FesDesc *d, const FesShootBuffers *b, void *) { return shoot_impl<double>(d, b); }
int fes_shoot_f32(const FesDesc *d, const FesShootBuffers *b, void *) { return shoot_impl<float>(d, b); }

// the static-scratch entry point: here the "device" array is a static host array of the same size
int fes_shoot_static_f64(const FesDesc *d, const FesShootBuffers *b, void *) {
    alignas(256) static unsigned char scratch[FES_STATIC_SCRATCH_BYTES];
    if (!d) return E_DESC;
    if (!b) return E_BUF;
    const int64_t wb = fes_workspace_bytes(d, 0);
    if (wb < 0) return E_SIZE;
    if (wb > (int64_t)FES_STATIC_SCRATCH_BYTES) return E_STATIC_SCRATCH;
    FesShootBuffers bb = *b;
    bb.workspace = scratch;
    bb.workspace_bytes = FES_STATIC_SCRATCH_BYTES;
    return shoot_impl<double>(d, &bb);
}

int fes_shoot_hess_num_dirs(int k, int f, int n, int dirs) {
    if (k < 0 || k > 2 || n < 1) return -1;
    return (f ? 5 : 2) + num_controls(k, n) + ((dirs & FES_DIR_C) ? 1 : 0);
}
int fes_shoot_hess_f64(const FesDesc *d, const FesHessBuffers *b, void *) {
    const int rc = check_hess(d, b);
    if (rc) return rc;
    HessArgs a;
    fill_hess_args(d, b, a);
    switch (d->model_kind * 2 + d->fatigue) {
        case 0: hess_cpu_kf<DING2003, false>(a); break;
        case 1: hess_cpu_kf<DING2003, true>(a); break;
        case 2: hess_cpu_kf<DING2007, false>(a); break;
        case 3: hess_cpu_kf<DING2007, true>(a); break;
        case 4: hess_cpu_kf<HMED2018, false>(a); break;
        default: hess_cpu_kf<HMED2018, true>(a); break;
    }
    return 0;
}
int fes_shoot_hess_host_f64(const FesDesc *d, const FesHessBuffers *b, void *s) { return fes_shoot_hess_f64(d, b, s); }

// host-buffer entry points: same loops, workspace allocated here
#define HOST_VARIANT(name, impl, T, Buf)                                                      \
    int name(const FesDesc *d, const Buf *b, void *) {                                        \
        if (!d || !b) return E_DESC;                                                          \
        FesDesc dd = *d;                                                                      \
        dd.flags &= ~FES_TABLES_READY;                                                        \
        const int64_t wb = fes_workspace_bytes(&dd, sizeof(T) == 4);                          \
        if (wb < 0) return E_SIZE;                                                            \
        std::vector<char> ws((size_t)wb + 16);                                                \
        Buf bb = *b;                                                                          \
        bb.workspace = ws.data();                                                             \
        bb.workspace_bytes = wb;                                                              \
        return impl<T>(&dd, &bb);                                                             \
    }
HOST_VARIANT(fes_ivp_host_f64, ivp_impl, double, FesIvpBuffers)
HOST_VARIANT(fes_ivp_host_f32, ivp_impl, float, FesIvpBuffers)
HOST_VARIANT(fes_shoot_host_f64, shoot_impl, double, FesShootBuffers)
HOST_VARIANT(fes_shoot_host_f32, shoot_impl, float, FesShootBuffers)
int fes_identify_f64(const FesDesc *d, const FesIdentBuffers *b, const FesIdentOptions *o, void *) { return ident_impl<double>(d, b, o); }
int fes_identify_f32(const FesDesc *d, const FesIdentBuffers *b, const FesIdentOptions *o, void *) { return ident_impl<float>(d, b, o); }
int fes_identify_host_f64(const FesDesc *d, const FesIdentBuffers *b, const FesIdentOptions *o, void *) { return ident_host<double>(d, b, o); }
int fes_identify_host_f32(const FesDesc *d, const FesIdentBuffers *b, const FesIdentOptions *o, void *) { return ident_host<float>(d, b, o); }
int fes_probe_fma_peak(int, int, void *, double *) { return E_DESC; }
int64_t fes_launch_count(void) { return 0; }
int fes_last_host_timing(double *ms5) { if (!ms5) return E_BUF; for (int i = 0; i < 5; ++i) ms5[i] = 0.0; return 0; }
int fes_host_pool_trim(void) { return 0; }
int fes_is_host_emulation(void) { return 1; }
}
