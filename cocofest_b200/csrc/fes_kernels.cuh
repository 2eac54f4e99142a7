// fes_kernels.cuh -- sm_100a kernels around fes_core.cuh and their launchers.
//
//   fes_tables_kernel  window tables + stage decay table (2N-1 exp per (node[, trajectory]), once per call)
//   fes_ivp_kernel     one thread per trajectory; single shooting over all nodes
//                      (replaces bioptim interfaces/solve_ivp_interface.py:56-69)
//   fes_shoot_kernel   one thread per (scenario, interval); defect + Jacobian columns
//                      (replaces bioptim limits/penalty.py:1056-1105 + casadi AD)
//
// Data layout: structure of arrays, batch index innermost, so every global access of a warp is one
// coalesced 256-byte (fp64) transaction.  Tables and controls that are shared by all trajectories are
// staged through shared memory in node chunks (ivp) or read through the read-only path with a
// warp-uniform address (shoot).  Trajectory state, parameters and all tangent columns live in registers
// for the whole integration; HBM sees only the initial load, the requested traces and the final store.
#pragma once

#if defined(__CUDACC__)
#include <cuda_runtime.h>
#include <cuda_pipeline.h>
#endif

#include <algorithm>

#include "fes_core.cuh"

namespace fes {

// launch shape: tuned on a B200 (scripts/build_variants.sh + the variants stage of scripts/gpu_run.sh, profiles/r01_tune_ivp_d.txt);
// overridable for that sweep only
#ifndef FES_BLOCK
#define FES_BLOCK 64
#endif
constexpr int BLOCK = FES_BLOCK;
// RK4 steps fused into one basic block (rk4_steps in fes_core.cuh): how many, per kernel family, by measurement on the B200
// (profiles/r02_tune_step_unroll_{r,s}.txt; u2 / u5 / u10 = blocks of 2 / 5 / 10 steps, S = 10):
//   trajectory kernel  fp64 headline +2.0 % (u2), +1.7 % (u5), +0.8 % (u10); config 3 +2.2 %; dense P|PF columns +4.0 % (u2);
//                      fp32 headline +7.7 % (u2), +9.7 % (u5); fp32 config 3 +10.7 % (u2), +7.7 % (u5)   -> pairs (fp32 x0: five)
//   interval kernel    fp64 config 5 (X0|U): 0.484 (u1), 0.486 (u2), 0.469 (u5), 0.571 ms (u10); config 5pf (dense P|PF): 0.790 (u1),
//                      0.758 (u2), 0.804 ms (u5); fp32 config 5: 0.341 (u1), 0.342 (u2), 0.352 ms (u5)  -> 5 / pairs / single (fp32)
//   identification     10.56 (u1), 10.75 (u2), 11.48 (u5), 12.03 ms (u10): it already spills its Lambda cache under the cap -> single
#ifndef FES_IVP_UNROLL
#define FES_IVP_UNROLL 2
#endif
#ifndef FES_IDENT_UNROLL
#define FES_IDENT_UNROLL 1
#endif
template <typename T, bool FAT, int GROUPS>
constexpr int ivp_unroll() {
#ifdef FES_IVP_UNROLL_ALL
    return FES_IVP_UNROLL_ALL;
#else
    // the fp32 fatigue kernel with the x0 columns is the one trajectory kernel that gains from blocks of five
    // (209.0 against 212.9 ms with pairs, 229.3 ms with single steps; config 3 in fp32 prefers pairs)
    return (sizeof(T) == 4 && FAT && GROUPS == G_X0) ? 5 : FES_IVP_UNROLL;
#endif
}
template <typename T, bool FAT, int GROUPS>
constexpr int shoot_unroll() {
#ifdef FES_SHOOT_UNROLL
    return FES_SHOOT_UNROLL;
#else
    if (sizeof(T) == 4) return 1;
    return (FAT && (GROUPS & (G_P | G_PF))) ? 2 : 5;
#endif
}
// Resident blocks per SM the compiler must make room for.  The fp64 fatigue kernels whose columns are all
// "reduced" (fes_core.cuh) take ~188 registers when unconstrained.  Asking for 5 blocks of 64 threads (a cap of
// 204 registers) makes ptxas settle on 162 registers = 6 resident blocks = 12 warps/SM with a schedule that
// measured 4.95e10 trajectory-steps/s on a B200, against 4.45e10 for the 170-register cap of (128, 3) or (64, 6)
// and 4.38e10 unconstrained (profiles/r01_tune_ivp_d.txt).
template <typename T, int KIND, bool FAT, int GROUPS>
constexpr int min_blocks() {
#ifdef FES_MIN_BLOCKS
    return FES_MIN_BLOCKS;
#else
#ifdef FES_F32_MIN_BLOCKS  // sweep knob for the float kernels (scripts/build_variants.sh)
    if (sizeof(T) == 4) return FES_F32_MIN_BLOCKS;
#endif
    // same effect on the two-state kernels (Hmed P columns: 182 registers unconstrained, +16 % under the 204 cap,
    // profiles/r01_tune_ident_g.txt); the dense fatigue parameter columns (P, P|PF: 204-254 registers) stay free
    constexpr int G0 = GROUPS & ~G_C;  // the force-factor column rides with the X0 | U sets and follows their policy
    constexpr bool red = FAT && (G0 == G_X0 || G0 == G_U || G0 == G_W || (G0 == (G_X0 | G_U) && KIND == DING2007));
    // Ding2007 parameter columns (config 3): 7 resident blocks = 120 registers, no spill, +4 % over the 138-register build;
    // every other kernel moved by < 1.5 % or lost (profiles/r01_tune_min_blocks_m.txt)
    if (sizeof(T) == 8 && !FAT && KIND == DING2007 && GROUPS == G_P) return 7;
    // with the 2 x 2 step propagator (PROP2) the reduced kernels fit 6 resident blocks outright: 164 registers, no spill,
    // headline +0.9 %, config 5 +1.5 % over the 5-block request (profiles/r01_tune_shape_ah.txt)
    if (sizeof(T) == 8 && red) return 6;
    return (sizeof(T) == 8 && !FAT) ? 5 : 2;
#endif
}
#define FES_LAUNCH_BOUNDS __launch_bounds__(BLOCK, (min_blocks<T, KIND, FAT, GROUPS>()))

template <typename T>
struct IvpArgs {
    const T *params; int64_t pB, pS;   // element j of trajectory b: params[j*pB + b*pS]
    const T *x0;                       // [nx][B]
    const T *tab; int64_t tB, tS; int row;  // tables [ns][row]([B]): tab[(n*row + c)*tB + b*tS]
    const T *eps;                      // [2S+1] stage decay factors
    const T *ctrl; int64_t cB, cS; int nu;  // [ns][nu]([B])
    T *x_final, *trace, *sens_final, *sens_trace;
    int64_t B, trace_B;
    int trace_mode;                    // 0 none, 1 every RK step, 2 nodes
    int N, ns, S, scheme, tile_off, chunk;
    T h, tauc, r0_km;
};

template <typename T>
struct ShootArgs {
    const T *params; int64_t pB, pS;
    const T *X;                        // [ns+1][nx][B]
    const T *tab; int64_t tB, tS; int row;
    const T *eps;
    const T *ctrl; int64_t cB, cS; int nu;
    T *G, *Phi, *J;
    int j_cols, j_col0;
    int64_t B;
    int N, ns, S, scheme, tile_off;
    T h, tauc, r0_km;
    int stage_tc;                      // set by the launcher: table row + controls are staged in shared memory too
};

#if defined(__CUDACC__)

// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void fes_tables_kernel(const double *__restrict__ win, int64_t wB, int64_t wS, int N, int ns, int row,
                                  double dt, double t_start, double tauc, T *__restrict__ tab, int64_t tB,
                                  int S, double h, int scheme, T *__restrict__ eps) {
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx < stage_table_len(S)) eps[idx] = (T)stage_table_entry((int)idx, S, h, tauc, scheme);
    if (idx >= (int64_t)ns * tB) return;
    const int64_t node = idx / tB, b = idx % tB;
    const double tn = t_start + dt * (double)node;  // optimization_vector.py:373: t_k = dt * k
    double a = 0.0, bsum = 0.0, prev = 0.0;
    for (int i = 0; i < N; ++i) {
        const double ti = win[(node * N + i) * wB + b * wS];
        const double wa = exp(-(tn - ti) / tauc);                          // exp_time_fun, ding2003.py:213
        const double wb = (i == 0) ? 0.0 : exp(-(ti - prev) / tauc) * wa;  // (R_i - 1)/(r0 - 1) * wa, ding2003.py:228,249
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

// ---------------------------------------------------------------------------------------------
template <typename T, int KIND, bool FAT, int GROUPS>
__global__ void FES_LAUNCH_BOUNDS fes_ivp_kernel(const IvpArgs<T> a) {
    using D = Dirs<KIND, FAT, GROUPS>;
    constexpr int NX = D::NX;
    constexpr int ND = D::ND;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T *s_eps = reinterpret_cast<T *>(smem_raw);
    T *s_tab = s_eps + ((stage_table_len(a.S) + 3) & ~3);
    T *s_ctl = s_tab + (a.tS == 0 ? a.chunk * a.row : 0);
    const int tid = threadIdx.x;
    for (int i = tid; i < stage_table_len(a.S); i += BLOCK) s_eps[i] = a.eps[i];

    const int64_t b = (int64_t)blockIdx.x * BLOCK + tid;
    const bool live = b < a.B;
    const int64_t bb = live ? b : a.B - 1;  // idle lanes shadow the last trajectory, they store nothing

    Traj<T> tr;
    load_traj(tr, [&](int j) { return a.params[j * a.pB + bb * a.pS]; }, a.tauc, a.r0_km);
    T x[NX];
#pragma unroll
    for (int j = 0; j < NX; ++j) x[j] = a.x0[j * a.B + bb];
    Tangent<T, KIND, FAT, GROUPS> tg;
    tg.init(x);
    LamCache<T> *lcp = nullptr;
    [[maybe_unused]] LamCache<T> lc;
#ifndef FES_NO_LAMCACHE  // sweep knob: with the branch-free tanh the cache of the previous interval's tanh may not pay any more
    if constexpr (KIND == HMED2018) {
        lc.reset(a.N);
        lcp = &lc;
    }
#endif

    const bool tr_on = live && a.trace_mode != 0 && b < a.trace_B;
    auto put_state = [&](int64_t pt) {
        if (a.trace != nullptr) {
#pragma unroll
            // plain stores: streaming stores (__stcs) made the write-bound config 1 SLOWER, 0.947 against 0.908 ms
            // (profiles/r02_tune_trace_stcs_af.txt)
            for (int j = 0; j < NX; ++j) a.trace[(pt * NX + j) * a.trace_B + b] = x[j];
        }
    };
    auto put_sens = [&](int64_t node) {
        if constexpr (ND > 0) {
            if (a.sens_trace != nullptr) {
                static_for<0, ND>([&](auto ic) {
                    constexpr int d = decltype(ic)::value;
#pragma unroll
                    for (int j = 0; j < NX; ++j) a.sens_trace[((node * NX + j) * ND + d) * a.trace_B + b] = tg.template get<d>(tr, j);
                });
            }
        }
    };
    if (tr_on) {
        put_state(0);
        put_sens(0);
    }

    for (int n0 = 0; n0 < a.ns; n0 += a.chunk) {
        const int nn = min(a.chunk, a.ns - n0);
        __syncthreads();
        if (a.tS == 0)
            for (int i = tid; i < nn * a.row; i += BLOCK) s_tab[i] = a.tab[(int64_t)n0 * a.row + i];
        if (a.nu > 0 && a.cS == 0)
            for (int i = tid; i < nn * a.nu; i += BLOCK) s_ctl[i] = a.ctrl[(int64_t)n0 * a.nu + i];
        __syncthreads();
        for (int m = 0; m < nn; ++m) {
            const int node = n0 + m;
            const T *trow = s_tab + m * a.row;
            int64_t tstr = 1;
            if (a.tS != 0) {
                trow = a.tab + (int64_t)node * a.row * a.tB + bb;
                tstr = a.tB;
            }
            const T *crow = s_ctl + m * a.nu;
            int64_t cstr = 1;
            if (a.cS != 0) {
                crow = a.ctrl + (int64_t)node * a.nu * a.cB + bb;
                cstr = a.cB;
            }
            Interval<T, ND> iv;
            interval_setup<T, KIND, FAT, GROUPS>(
                tr, [&](int c) { return trow[c * tstr]; }, [&](int c) { return crow[c * cstr]; }, a.N, a.tile_off, iv, lcp);
            constexpr int UNROLL = ivp_unroll<T, FAT, GROUPS>();
            int i = 0;
            if constexpr (UNROLL > 1)
                if (a.scheme == 4 && !(tr_on && a.trace_mode == 1))
                    for (; i + UNROLL <= a.S; i += UNROLL) rk4_steps<T, KIND, FAT, GROUPS, UNROLL>(tr, iv, a.h, s_eps, a.S, i, x, tg);
            for (; i < a.S; ++i) {
                rk_step<T, KIND, FAT, GROUPS>(tr, iv, a.scheme, a.h, s_eps[2 * i], s_eps[2 * i + 1], s_eps[2 * i + 2],
                                              s_eps + 2 * a.S + 1 + 4 * i, x, tg);
                if (tr_on && a.trace_mode == 1) put_state((int64_t)node * a.S + i + 1);
            }
            if (tr_on) {
                if (a.trace_mode == 2) put_state(node + 1);
                put_sens(node + 1);
            }
        }
    }
    if (live) {
#pragma unroll
        for (int j = 0; j < NX; ++j) a.x_final[j * a.B + b] = x[j];
        if constexpr (ND > 0) {
            if (a.sens_final != nullptr) {
                static_for<0, ND>([&](auto ic) {
                    constexpr int d = decltype(ic)::value;
#pragma unroll
                    for (int j = 0; j < NX; ++j) a.sens_final[(j * ND + d) * a.B + b] = tg.template get<d>(tr, j);
                });
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Every thread integrates ONE interval of S steps, so the latency of its input loads is paid once per ~3000 issued
// instructions instead of once per trajectory: left to the compiler (under the register cap the loads sink next to their
// first use) four dependent round trips to HBM were exposed per thread -- long_scoreboard 1.33 cycles per issue, fp64 pipe
// 70 % active (profiles/r01_ncu_shoot_cfg5_ai_summary.tsv).  Here the thread's inputs (node state, next node state, and the
// table row + controls when they fit) are copied global -> shared with cp.async (LDGSTS, no registers) as the first thing
// the thread does, all in flight at once and overlapped with the stage-table load and the parameter set-up; the
// integration then reads them from its own shared-memory column [value][thread] (conflict-free).  Measured on config 5:
// 0.539 -> 0.522 ms.  A PERSISTENT grid with the next item's inputs prefetched into a double buffer was tried on top of
// this and lost (0.627 ms, profiles/r02_tune_persistent_shoot_d_f64.jsonl): identical work items keep the resident warps in
// phase, so their set-up / store phases no longer overlap other warps' RK phases; the hardware's own staggered block launch
// does that better.
template <typename T, int KIND, bool FAT, int GROUPS>
__global__ void FES_LAUNCH_BOUNDS fes_shoot_kernel(const ShootArgs<T> a) {
    using D = Dirs<KIND, FAT, GROUPS>;
    constexpr int NX = D::NX;
    constexpr int ND = D::ND;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T *s_eps = reinterpret_cast<T *>(smem_raw);
    T *s_in = s_eps + ((stage_table_len(a.S) + 3) & ~3) + threadIdx.x;   // this thread's column: element c at s_in[c * BLOCK]
    const int tid = threadIdx.x;
    const int64_t idx = (int64_t)blockIdx.x * BLOCK + tid;
    const bool live = idx < a.B * a.ns;
    const int64_t k = live ? idx / a.B : 0, b = live ? idx % a.B : 0;  // batch innermost: a warp shares k when 32 | B
    const T *trow = a.tab + k * a.row * a.tB + b * a.tS;
    const T *crow = a.ctrl + k * a.nu * a.cB + b * a.cS;
    const bool want_g = a.G != nullptr;
    T *s_tab = s_in + (want_g ? 2 * NX : NX) * BLOCK, *s_ctl = s_tab + a.row * BLOCK;
    if (live) {
#pragma unroll
        for (int j = 0; j < NX; ++j) __pipeline_memcpy_async(s_in + j * BLOCK, a.X + (k * NX + j) * a.B + b, sizeof(T));
        if (want_g) {
#pragma unroll
            for (int j = 0; j < NX; ++j)
                __pipeline_memcpy_async(s_in + (NX + j) * BLOCK, a.X + ((k + 1) * NX + j) * a.B + b, sizeof(T));
        }
        if (a.stage_tc) {
            for (int c = 0; c < a.row; ++c) __pipeline_memcpy_async(s_tab + c * BLOCK, trow + c * a.tB, sizeof(T));
            for (int c = 0; c < a.nu; ++c) __pipeline_memcpy_async(s_ctl + c * BLOCK, crow + c * a.cB, sizeof(T));
        }
    }
    __pipeline_commit();
    for (int i = tid; i < stage_table_len(a.S); i += BLOCK) s_eps[i] = a.eps[i];
    Traj<T> tr;
    load_traj(tr, [&](int j) { return __ldg(a.params + j * a.pB + b * a.pS); }, a.tauc, a.r0_km);
    __pipeline_wait_prior(0);
    __syncthreads();
    if (!live) return;

    T x[NX];
#pragma unroll
    for (int j = 0; j < NX; ++j) x[j] = s_in[j * BLOCK];
    Tangent<T, KIND, FAT, GROUPS> tg;
    tg.init(x);

    Interval<T, ND> iv;
    if (a.stage_tc)
        interval_setup<T, KIND, FAT, GROUPS>(
            tr, [&](int c) { return s_tab[c * BLOCK]; }, [&](int c) { return s_ctl[c * BLOCK]; }, a.N, a.tile_off, iv);
    else
        interval_setup<T, KIND, FAT, GROUPS>(
            tr, [&](int c) { return __ldg(trow + c * a.tB); }, [&](int c) { return __ldg(crow + c * a.cB); }, a.N,
            a.tile_off, iv);
    int i = 0;
    constexpr int UNROLL = shoot_unroll<T, FAT, GROUPS>();
    if constexpr (UNROLL > 1)
        if (a.scheme == 4)
            for (; i + UNROLL <= a.S; i += UNROLL) rk4_steps<T, KIND, FAT, GROUPS, UNROLL>(tr, iv, a.h, s_eps, a.S, i, x, tg);
    for (; i < a.S; ++i)
        rk_step<T, KIND, FAT, GROUPS>(tr, iv, a.scheme, a.h, s_eps[2 * i], s_eps[2 * i + 1], s_eps[2 * i + 2],
                                      s_eps + 2 * a.S + 1 + 4 * i, x, tg);

#pragma unroll
    for (int j = 0; j < NX; ++j) {
        if (a.Phi != nullptr) a.Phi[(k * NX + j) * a.B + b] = x[j];
        if (want_g) a.G[(k * NX + j) * a.B + b] = s_in[(NX + j) * BLOCK] - x[j];
    }
    if constexpr (ND > 0) {
        if (a.J != nullptr) {
            static_for<0, ND>([&](auto ic) {
                constexpr int d = decltype(ic)::value;
#pragma unroll
                for (int j = 0; j < NX; ++j) a.J[((k * NX + j) * a.j_cols + a.j_col0 + d) * a.B + b] = tg.template get<d>(tr, j);
            });
        }
    }
}

// ---------------------------------------------------------------------------------------------
template <typename T>
inline size_t ivp_smem_bytes(const IvpArgs<T> &a) {
    size_t n = (stage_table_len(a.S) + 3) & ~3;
    if (a.tS == 0) n += (size_t)a.chunk * a.row;
    if (a.nu > 0 && a.cS == 0) n += (size_t)a.chunk * a.nu;
    return n * sizeof(T);
}

template <typename T, int KIND, bool FAT, int GROUPS>
int launch_ivp_t(const IvpArgs<T> &a, cudaStream_t s) {
    const unsigned grid = (unsigned)((a.B + BLOCK - 1) / BLOCK);
    fes_ivp_kernel<T, KIND, FAT, GROUPS><<<grid, BLOCK, ivp_smem_bytes(a), s>>>(a);
    return (int)cudaGetLastError();
}

// shared memory a shoot block may spend on staged table rows + controls (the node states are always staged): with 6
// resident blocks this stays well inside the SM's 227 KB
constexpr size_t SHOOT_STAGE_BUDGET = 24 * 1024;

template <typename T, int KIND, bool FAT, int GROUPS>
int launch_shoot_t(const ShootArgs<T> &a0, cudaStream_t s) {
    ShootArgs<T> a = a0;
    const int64_t total = a.B * a.ns;
    const unsigned grid = (unsigned)((total + BLOCK - 1) / BLOCK);
    constexpr int NX = Dirs<KIND, FAT, GROUPS>::NX;
    size_t vals = (size_t)(a.G != nullptr ? 2 * NX : NX);
    a.stage_tc = (vals + a.row + a.nu) * BLOCK * sizeof(T) <= SHOOT_STAGE_BUDGET ? 1 : 0;
    if (a.stage_tc) vals += (size_t)a.row + a.nu;
    const size_t smem = (((stage_table_len(a.S) + 3) & ~3) + vals * BLOCK) * sizeof(T);
    fes_shoot_kernel<T, KIND, FAT, GROUPS><<<grid, BLOCK, smem, s>>>(a);
    return (int)cudaGetLastError();
}

constexpr int ERR_DIRS = -4;

// one translation unit per (KIND, FAT) instantiates these two for double and float
template <typename T, int KIND, bool FAT>
int launch_ivp_kf(int groups, const IvpArgs<T> &a, cudaStream_t s) {
    switch (groups) {
        case 0: return launch_ivp_t<T, KIND, FAT, 0>(a, s);
        case G_X0: return launch_ivp_t<T, KIND, FAT, G_X0>(a, s);
        case G_P: return launch_ivp_t<T, KIND, FAT, G_P>(a, s);
        case G_P | G_PF:
            if constexpr (FAT) return launch_ivp_t<T, KIND, FAT, G_P | G_PF>(a, s);
            else return ERR_DIRS;
        default: return ERR_DIRS;
    }
}

template <typename T, int KIND, bool FAT>
int launch_shoot_kf(int groups, const ShootArgs<T> &a, cudaStream_t s) {
    switch (groups) {
        case 0: return launch_shoot_t<T, KIND, FAT, 0>(a, s);
        case G_X0: return launch_shoot_t<T, KIND, FAT, G_X0>(a, s);
        case G_X0 | G_U:
            if constexpr (KIND != DING2003) return launch_shoot_t<T, KIND, FAT, G_X0 | G_U>(a, s);
            else return launch_shoot_t<T, KIND, FAT, G_X0>(a, s);
        case G_U:
            if constexpr (KIND != DING2003) return launch_shoot_t<T, KIND, FAT, G_U>(a, s);
            else return ERR_DIRS;
        case G_X0 | G_U | G_C:  // with the force factor of the interval as an input and a column (FES_DIR_C)
            if constexpr (KIND != DING2003) return launch_shoot_t<T, KIND, FAT, G_X0 | G_U | G_C>(a, s);
            else return launch_shoot_t<T, KIND, FAT, G_X0 | G_C>(a, s);
        case G_X0 | G_C:
            if constexpr (KIND == DING2003) return launch_shoot_t<T, KIND, FAT, G_X0 | G_C>(a, s);
            else return ERR_DIRS;
        case G_U | G_C:  // Hmed2018: the intensity tiles after the first of a call that carries the force factor
            if constexpr (KIND == HMED2018) return launch_shoot_t<T, KIND, FAT, G_U | G_C>(a, s);
            else return ERR_DIRS;
        case G_P: return launch_shoot_t<T, KIND, FAT, G_P>(a, s);
        case G_P | G_PF:
            if constexpr (FAT) return launch_shoot_t<T, KIND, FAT, G_P | G_PF>(a, s);
            else return ERR_DIRS;
        case G_W: return launch_shoot_t<T, KIND, FAT, G_W>(a, s);
        default: return ERR_DIRS;
    }
}

#endif  // __CUDACC__

}  // namespace fes
