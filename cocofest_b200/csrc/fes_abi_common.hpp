// fes_abi_common.hpp -- host-only, CUDA-free part of the C ABI: argument checks, workspace carving and the
// translation of (FesDesc, buffers) into kernel argument blocks.  Shared by fes_api.cu and by the
// host build of the kernel arithmetic that the CPU-only CI uses (tests/host_emul, test infrastructure).
#pragma once

#include <algorithm>
#include <cstring>

#include "fes_b200.h"
#include "fes_kernels.cuh"
#include "fes_ident.cuh"
#include "fes_hess.cuh"

namespace fes {
namespace abi {

enum {
    E_OK = 0, E_DESC = -1, E_MODEL = -2, E_SIZE = -3, E_DIRS = -4, E_BUF = -5, E_WORKSPACE = -6, E_TRACE = -7,
    E_STATIC_SCRATCH = -8,
};

inline int num_controls(int kind, int N) { return kind == FES_DING2007 ? 1 : (kind == FES_HMED2018 ? N : 0); }

// number of columns of a direction set; < 0 when the set is not one of the instantiated ones
inline int num_dirs(int kind, int fat, int dirs) {
    const int nx = fat ? 5 : 2;
    const int nu_t = kind == FES_DING2007 ? 1 : (kind == FES_HMED2018 ? FES_TILE : 0);
    const int np = kind == FES_DING2003 ? 4 : (kind == FES_DING2007 ? 6 : 8);
    switch (dirs) {
        case 0: return 0;
        case FES_DIR_X0: return nx;
        case FES_DIR_X0 | FES_DIR_U: return nx + nu_t;
        case FES_DIR_U: return nu_t > 0 ? nu_t : -1;
        case FES_DIR_X0 | FES_DIR_U | FES_DIR_C: return nx + nu_t + 1;
        case FES_DIR_X0 | FES_DIR_C: return kind == FES_DING2003 ? nx + 1 : -1;
        case FES_DIR_U | FES_DIR_C: return kind == FES_HMED2018 ? nu_t + 1 : -1;  // further intensity tiles of a factor call
        case FES_DIR_P: return np;
        case FES_DIR_P | FES_DIR_PF: return fat ? np + 4 : -1;
        case FES_DIR_W: return FES_TILE;
    }
    return -1;
}

inline bool slot_tables(int kind, int dirs) { return kind == FES_HMED2018 || (dirs & FES_DIR_W); }

inline int check_desc(const FesDesc *d, bool ivp) {
    if (d == nullptr || d->abi_version != FES_ABI_VERSION) return E_DESC;
    if (d->model_kind < 0 || d->model_kind > 2 || (d->fatigue != 0 && d->fatigue != 1)) return E_MODEL;
    if (d->scheme != 1 && d->scheme != 2 && d->scheme != 4) return E_MODEL;
    // per-slot window tables (Hmed2018 intensities, pulse-time columns) are staged per thread: 64 slots at most; the other
    // models only see the two sums of the window, whatever its length (the truncation sweep goes to 300 pulses)
    const int max_trunc = slot_tables(d->model_kind, d->dirs) ? 64 : 1024;
    if (d->n_trunc < 1 || d->n_trunc > max_trunc || d->n_shooting < 1 || d->n_steps < 1 || d->n_steps > 256) return E_SIZE;
    if (d->batch < 1 || !(d->dt > 0.0) || !(d->tauc > 0.0)) return E_SIZE;
    if (num_dirs(d->model_kind, d->fatigue, d->dirs) < 0) return E_DIRS;
    if (ivp && (d->dirs & (FES_DIR_U | FES_DIR_W | FES_DIR_C))) return E_DIRS;
    if (d->tile_offset < 0 || d->tile_offset >= d->n_trunc + FES_TILE) return E_SIZE;
    if (ivp) {
        if (d->trace_every != 0 && d->trace_every != 1 && d->trace_every != d->n_steps) return E_TRACE;
        if (d->trace_batch < 0 || d->trace_batch > d->batch) return E_TRACE;
    }
    return E_OK;
}

struct Carve {
    size_t eps_off, tab_off, total;
    int row;
    int64_t tB;
};
inline Carve carve(const FesDesc *d, size_t tsize) {
    Carve c;
    c.row = slot_tables(d->model_kind, d->dirs) ? 2 + 2 * d->n_trunc : 2;
    c.tB = (d->flags & FES_WIN_SHARED) ? 1 : d->batch;
    c.eps_off = 0;
    c.tab_off = ((size_t)stage_table_len(d->n_steps) * tsize + 255) & ~(size_t)255;
    c.total = c.tab_off + (size_t)d->n_shooting * c.row * (size_t)c.tB * tsize;
    return c;
}


template <typename T>
void fill_ivp_args(const FesDesc *d, const FesIvpBuffers *b, const Carve &c, IvpArgs<T> &a) {
    const int nu = num_controls(d->model_kind, d->n_trunc);
    std::memset(&a, 0, sizeof(a));
    a.params = (const T *)b->params;
    a.pB = (d->flags & FES_PARAMS_SHARED) ? 1 : d->batch;
    a.pS = (d->flags & FES_PARAMS_SHARED) ? 0 : 1;
    a.x0 = (const T *)b->x0;
    a.tab = reinterpret_cast<const T *>((char *)b->workspace + c.tab_off);
    a.tB = c.tB;
    a.tS = (d->flags & FES_WIN_SHARED) ? 0 : 1;
    a.row = c.row;
    a.eps = reinterpret_cast<const T *>((char *)b->workspace + c.eps_off);
    a.ctrl = (const T *)b->ctrl;
    a.cB = (d->flags & FES_CTRL_SHARED) ? 1 : d->batch;
    a.cS = (d->flags & FES_CTRL_SHARED) ? 0 : 1;
    a.nu = nu;
    a.x_final = (T *)b->x_final;
    a.trace = (T *)b->trace;
    a.sens_final = (T *)b->sens_final;
    a.sens_trace = (T *)b->sens_trace;
    a.B = d->batch;
    a.trace_B = d->trace_batch;
    a.trace_mode = (d->trace_every == 0 || d->trace_batch == 0) ? 0 : (d->trace_every == 1 ? 1 : 2);
    a.N = d->n_trunc;
    a.ns = d->n_shooting;
    a.S = d->n_steps;
    a.scheme = d->scheme;
    a.tile_off = d->tile_offset;
    a.h = (T)(d->dt / d->n_steps);
    a.tauc = (T)d->tauc;
    a.r0_km = (T)d->r0_km;
    // nodes staged per shared-memory chunk: keep the staged tables + controls within 16 KB
    const size_t per_node = ((a.tS == 0 ? a.row : 0) + ((nu > 0 && a.cS == 0) ? nu : 0)) * sizeof(T);
    a.chunk = per_node ? (int)std::max<size_t>(1, std::min<size_t>((size_t)d->n_shooting, 16384 / per_node)) : d->n_shooting;
}

template <typename T>
void fill_shoot_args(const FesDesc *d, const FesShootBuffers *b, const Carve &c, ShootArgs<T> &a) {
    std::memset(&a, 0, sizeof(a));
    a.params = (const T *)b->params;
    a.pB = (d->flags & FES_PARAMS_SHARED) ? 1 : d->batch;
    a.pS = (d->flags & FES_PARAMS_SHARED) ? 0 : 1;
    a.X = (const T *)b->X;
    a.tab = reinterpret_cast<const T *>((char *)b->workspace + c.tab_off);
    a.tB = c.tB;
    a.tS = (d->flags & FES_WIN_SHARED) ? 0 : 1;
    a.row = c.row;
    a.eps = reinterpret_cast<const T *>((char *)b->workspace + c.eps_off);
    a.ctrl = (const T *)b->ctrl;
    a.cB = (d->flags & FES_CTRL_SHARED) ? 1 : d->batch;
    a.cS = (d->flags & FES_CTRL_SHARED) ? 0 : 1;
    a.nu = num_controls(d->model_kind, d->n_trunc) + ((d->dirs & FES_DIR_C) ? 1 : 0);  // rows of ctrl per node
    a.G = (T *)b->G;
    a.Phi = (T *)b->Phi;
    a.J = (T *)b->J;
    a.j_cols = b->j_cols;
    a.j_col0 = b->j_col0;
    a.B = d->batch;
    a.N = d->n_trunc;
    a.ns = d->n_shooting;
    a.S = d->n_steps;
    a.scheme = d->scheme;
    a.tile_off = d->tile_offset;
    a.h = (T)(d->dt / d->n_steps);
    a.tauc = (T)d->tauc;
    a.r0_km = (T)d->r0_km;
}

inline int ident_num_params(int kind) { return kind == FES_DING2003 ? 4 : (kind == FES_DING2007 ? 6 : 8); }

inline int check_ident(const FesDesc *d, const FesIdentBuffers *b, const FesIdentOptions *o) {
    int rc = check_desc(d, true);
    if (rc) return rc;
    if (d->fatigue != 0) return E_MODEL;   // ding2003_force_parameter_identification.py:196-199
    if (d->dirs != FES_DIR_P) return E_DIRS;
    if (b == nullptr || o == nullptr || b->params == nullptr || b->target == nullptr || b->params_out == nullptr ||
        b->cost == nullptr || b->grad == nullptr || b->hess == nullptr)
        return E_BUF;
    if (num_controls(d->model_kind, d->n_trunc) > 0 && b->ctrl == nullptr) return E_BUF;
    if (o->max_iter < 0 || o->max_iter > 10000 || !(o->lambda0 > 0.0) || !(o->ftol >= 0.0) || !(o->xtol >= 0.0)) return E_SIZE;
    if (o->p_mask < 0 || o->p_mask >= (1 << ident_num_params(d->model_kind))) return E_DIRS;
    return E_OK;
}

template <typename T>
void fill_ident_args(const FesDesc *d, const FesIdentBuffers *b, const FesIdentOptions *o, const Carve &c, IdentArgs<T> &a) {
    std::memset(&a, 0, sizeof(a));
    a.params = (const T *)b->params;
    a.pB = (d->flags & FES_PARAMS_SHARED) ? 1 : d->batch;
    a.pS = (d->flags & FES_PARAMS_SHARED) ? 0 : 1;
    a.x0 = (const T *)b->x0;
    a.tab = reinterpret_cast<const T *>((char *)b->workspace + c.tab_off);
    a.tB = c.tB;
    a.tS = (d->flags & FES_WIN_SHARED) ? 0 : 1;
    a.row = c.row;
    a.eps = reinterpret_cast<const T *>((char *)b->workspace + c.eps_off);
    a.ctrl = (const T *)b->ctrl;
    a.cB = (d->flags & FES_CTRL_SHARED) ? 1 : d->batch;
    a.cS = (d->flags & FES_CTRL_SHARED) ? 0 : 1;
    a.nu = num_controls(d->model_kind, d->n_trunc);
    a.target = (const T *)b->target;
    a.gB = (d->flags & FES_TARGET_SHARED) ? 1 : d->batch;
    a.gS = (d->flags & FES_TARGET_SHARED) ? 0 : 1;
    a.weight = (const T *)b->weight;
    a.lower = (const T *)b->lower;
    a.upper = (const T *)b->upper;
    a.params_out = (T *)b->params_out;
    a.cost = (T *)b->cost;
    a.grad = (T *)b->grad;
    a.hess = (T *)b->hess;
    a.iters = b->iters;
    a.B = d->batch;
    a.N = d->n_trunc;
    a.ns = d->n_shooting;
    a.S = d->n_steps;
    a.scheme = d->scheme;
    a.max_iter = o->max_iter;
    a.p_mask = o->p_mask;
    a.h = (T)(d->dt / d->n_steps);
    a.dt = (T)d->dt;
    a.tauc = (T)d->tauc;
    a.r0_km = (T)d->r0_km;
    a.lambda0 = (T)o->lambda0;
    a.ftol = (T)o->ftol;
    a.xtol = (T)o->xtol;
}

// ---- second-order tangents (fes_shoot_hess_f64) ------------------------------------------------------------------------
// directions z = (x, the model's controls [, force factor]); desc->dirs must be X0 | U or X0 | U | C (U is ignored for ding2003)
inline int hess_num_dirs(const FesDesc *d) {
    return (d->fatigue ? 5 : 2) + num_controls(d->model_kind, d->n_trunc) + ((d->dirs & FES_DIR_C) ? 1 : 0);
}
inline int check_hess(const FesDesc *d, const FesHessBuffers *b) {
    if (d == nullptr || d->abi_version != FES_ABI_VERSION) return E_DESC;
    if (d->model_kind < 0 || d->model_kind > 2 || (d->fatigue != 0 && d->fatigue != 1)) return E_MODEL;
    if (d->scheme != 1 && d->scheme != 2 && d->scheme != 4) return E_MODEL;
    if (d->n_trunc < 1 || d->n_trunc > (d->model_kind == FES_HMED2018 ? 64 : 1024) || d->n_shooting < 1 || d->n_steps < 1 ||
        d->n_steps > 256)
        return E_SIZE;
    if (d->batch < 1 || !(d->dt > 0.0) || !(d->tauc > 0.0)) return E_SIZE;
    if ((d->dirs & ~(FES_DIR_C | FES_DIR_U)) != FES_DIR_X0) return E_DIRS;
    if (d->model_kind != FES_DING2003 && !(d->dirs & FES_DIR_U)) return E_DIRS;  // all of z or nothing
    if (b == nullptr || b->params == nullptr || b->X == nullptr || b->stim_win == nullptr) return E_BUF;
    if (num_controls(d->model_kind, d->n_trunc) + ((d->dirs & FES_DIR_C) ? 1 : 0) > 0 && b->ctrl == nullptr) return E_BUF;
    if (b->D2 == nullptr && b->H == nullptr) return E_BUF;
    if (b->H != nullptr && b->multipliers == nullptr) return E_BUF;
    return E_OK;
}
inline void fill_hess_args(const FesDesc *d, const FesHessBuffers *b, HessArgs &a) {
    std::memset(&a, 0, sizeof(a));
    a.params = (const double *)b->params;
    a.pB = (d->flags & FES_PARAMS_SHARED) ? 1 : d->batch;
    a.pS = (d->flags & FES_PARAMS_SHARED) ? 0 : 1;
    a.X = (const double *)b->X;
    a.win = b->stim_win;
    a.wB = (d->flags & FES_WIN_SHARED) ? 1 : d->batch;
    a.wS = (d->flags & FES_WIN_SHARED) ? 0 : 1;
    a.ctrl = (const double *)b->ctrl;
    a.cB = (d->flags & FES_CTRL_SHARED) ? 1 : d->batch;
    a.cS = (d->flags & FES_CTRL_SHARED) ? 0 : 1;
    a.factor = (d->dirs & FES_DIR_C) ? 1 : 0;
    a.nur = num_controls(d->model_kind, d->n_trunc) + a.factor;
    a.lam = (const double *)b->multipliers;
    a.D2 = (double *)b->D2;
    a.H = (double *)b->H;
    a.B = d->batch;
    a.N = d->n_trunc;
    a.ns = d->n_shooting;
    a.S = d->n_steps;
    a.scheme = d->scheme;
    a.nz = hess_num_dirs(d);
    a.npairs = a.nz * (a.nz + 1) / 2;
    a.h = d->dt / d->n_steps;
    a.dt = d->dt;
    a.t_start = d->t_start;
    a.tauc = d->tauc;
    a.r0_km = d->r0_km;
}

inline const char *strerror_abi(int code) {
    switch (code) {
        case E_OK: return "ok";
        case E_DESC: return "null descriptor or abi_version mismatch";
        case E_MODEL: return "invalid model_kind / fatigue / scheme";
        case E_SIZE: return "invalid size (n_trunc 1..64 [1..1024 for ding2003 / ding2007 without pulse-time columns], n_steps 1..256, n_shooting, batch, dt, tauc, tile_offset, J columns)";
        case E_DIRS: return "unsupported sensitivity direction set for this model / entry point";
        case E_BUF: return "a required buffer is NULL";
        case E_WORKSPACE: return "workspace is NULL or smaller than fes_workspace_bytes";
        case E_TRACE: return "invalid trace request (trace_every must be 0, 1 or n_steps; trace_batch <= batch; trace buffer)";
        case E_STATIC_SCRATCH: return "the tables of this call need more than FES_STATIC_SCRATCH_BYTES of scratch: use fes_shoot_f64 with a caller workspace";
    }
    return nullptr;
}

}  // namespace abi
}  // namespace fes
