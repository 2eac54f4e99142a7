// fes_ident.cuh -- force-tracking cost assembly and batched Levenberg-Marquardt identification on device.
//
// What the reference does (one candidate at a time, on the host):
//   - residuals r_j = F(t_j) - F_data(t_j) at all ns+1 nodes, cost = sum_j w_j r_j^2 with w_j = dt
//     (Lagrange TRACK_STATE on "F", Node.ALL, quadratic: optimization/fes_identification_ocp.py:248-266;
//     the data are resampled at the nodes by identification_method.py:253-278, force_at_node_in_ocp)
//   - decision parameters a_rest, km_rest, tau1_rest, tau2 (+ pd0, pdt | ar, bs, Is, cr) with box bounds
//     (fes_identification_ocp.py:269-299; defaults ding2003_force_parameter_identification.py:121-150,
//     ding2007_...:105-148, hmed2018_...:103-160), fatigue models rejected (ding2003_...:196-199)
//   - one Ipopt solve per candidate through casadi AD.
//
// Here: ONE thread owns one candidate for the WHOLE identification.  It integrates the two-state model with
// the chained parameter tangents (fes_core.cuh, direction group P), folds every node's residual into the
// cost, its gradient and the Gauss-Newton matrix J^T W J held in registers (the [ns+1] x np Jacobian is never
// written to HBM), then iterates a bounded Levenberg-Marquardt loop (Marquardt-scaled Cholesky solve in
// registers, projection on the box, accept / reject with lambda update) without leaving the kernel.
// max_iter = 0 turns the same kernel into the plain cost / gradient / Gauss-Newton evaluation.
#pragma once

#include "fes_kernels.cuh"

namespace fes {

template <int KIND>
struct IdentDims {
    static constexpr int NPK = KIND == DING2003 ? 4 : (KIND == DING2007 ? 6 : 8);  // identifiable parameters
    static constexpr int NH = NPK * (NPK + 1) / 2;                                 // upper triangle, row-major
    // canonical parameter row of P-group column d (same order as Dirs<KIND, false, G_P>::code)
    static constexpr int canon(int d) { return d < 4 ? d : (KIND == DING2007 ? P_PD0 + (d - 4) : P_AR + (d - 4)); }
    static constexpr int hidx(int d, int e) { return d * NPK - d * (d - 1) / 2 + (e - d); }  // d <= e
};

template <typename T>
struct IdentArgs {
    const T *params; int64_t pB, pS;   // starting points: params[j*pB + b*pS]
    const T *x0;                       // [2][B] | nullptr (rest: Cn = F = 0)
    const T *tab; int64_t tB, tS; int row;
    const T *eps;
    const T *ctrl; int64_t cB, cS; int nu;
    const T *target; int64_t gB, gS;   // force data at the nodes: target[node*gB + b*gS]
    const T *weight;                   // [ns+1] | nullptr (every node weighs dt)
    const T *lower, *upper;            // [NP] box of the canonical parameters | nullptr
    T *params_out, *cost, *grad, *hess;
    int32_t *iters;
    int64_t B;
    int N, ns, S, scheme, max_iter, p_mask;
    T h, dt, tauc, r0_km, lambda0, ftol, xtol;
};

// cost, gradient and Gauss-Newton matrix of one candidate (parameters pc[NP]) -------------------------
// The np + np(np+1)/2 running sums live in `acc` (element i at acc[i * stride]): on the device that is a
// conflict-free column of shared memory per thread (stride = block size), which keeps 88 registers (Hmed, fp64)
// out of the RK loop; they are touched once per node, i.e. once per S RK steps.  pc[j * pstride] = parameter j.
// Where a candidate reads its window-table rows and controls from.  UNIT: a dense [ns][row] / [ns][nu] copy shared by all
// candidates (the kernel stages it in shared memory when the tables and controls carry no batch axis): unit column stride
// and 32-bit indices, which takes the 64-bit address arithmetic of the general strided layout out of the per-slot loop.
template <typename T, bool UNIT>
struct IdentRows {
    const T *tab, *ctl;
    int64_t t_node, t_col, c_node, c_col;  // UNIT: t_node = row, c_node = nu, column strides 1
};

template <typename T, int KIND, bool UNIT = false>
FES_HD void ident_eval(const IdentArgs<T> &a, const IdentRows<T, UNIT> &rows, int64_t b, const T *pc, int pstride, const T *eps,
                       T &cost, T *acc, int stride) {
    using I = IdentDims<KIND>;
    constexpr int NPK = I::NPK;
    Traj<T> tr;
    load_traj(tr, [&](int j) { return pc[j * pstride]; }, a.tauc, a.r0_km);
    T x[2] = {T(0), T(0)};
    if (a.x0 != nullptr) {
        x[0] = a.x0[b];
        x[1] = a.x0[a.B + b];
    }
    Tangent<T, KIND, false, G_P> tg;
    tg.init(x);
    LamCache<T> *lcp = nullptr;
    [[maybe_unused]] LamCache<T> lc;
    // measured (profiles/r01_tune_ident_lamcache_x.txt): +2.5 % for the plain evaluation, -2.4 % inside the LM loop,
    // where the cache would be refilled at every evaluation (bs and Is move) next to ~1.3 KB of register spills
#ifndef FES_NO_LAMCACHE
    if (KIND == HMED2018 && a.max_iter == 0) {
        lc.reset(a.N);
        lcp = &lc;
    }
#endif
    cost = T(0);
    T *g = acc, *H = acc + NPK * stride;
    static_for<0, NPK + I::NH>([&](auto ic) { acc[decltype(ic)::value * stride] = T(0); });
    auto fold = [&](int node) {
        const T w = a.weight != nullptr ? a.weight[node] : a.dt;
        const T r = x[1] - a.target[(int64_t)node * a.gB + b * a.gS];
        const T wr = w * r;
        cost = xfma(wr, r, cost);
        static_for<0, NPK>([&](auto id) {
            constexpr int d = decltype(id)::value;
            const T vd = tg.V[d][1];
            g[d * stride] = xfma(wr + wr, vd, g[d * stride]);
            const T wv = (w + w) * vd;
            static_for<d, NPK>([&](auto ie) {
                constexpr int e = decltype(ie)::value;
                H[I::hidx(d, e) * stride] = xfma(wv, tg.V[e][1], H[I::hidx(d, e) * stride]);
            });
        });
    };
    fold(0);
    for (int node = 0; node < a.ns; ++node) {
        Interval<T, NPK> iv;
        if constexpr (UNIT) {
            const T *trow = rows.tab + node * (int)rows.t_node, *crow = rows.ctl + node * (int)rows.c_node;
            interval_setup<T, KIND, false, G_P>(tr, [&](int c) { return trow[c]; }, [&](int c) { return crow[c]; }, a.N, 0, iv, lcp);
        } else {
            const T *trow = rows.tab + (int64_t)node * rows.t_node, *crow = rows.ctl + (int64_t)node * rows.c_node;
            interval_setup<T, KIND, false, G_P>(
                tr, [&](int c) { return trow[c * rows.t_col]; }, [&](int c) { return crow[c * rows.c_col]; }, a.N, 0, iv, lcp);
        }
        int i = 0;
        if constexpr (FES_IDENT_UNROLL > 1)
            if (a.scheme == 4)
                for (; i + FES_IDENT_UNROLL <= a.S; i += FES_IDENT_UNROLL)
                    rk4_steps<T, KIND, false, G_P, FES_IDENT_UNROLL>(tr, iv, a.h, eps, a.S, i, x, tg);
        for (; i < a.S; ++i)
            rk_step<T, KIND, false, G_P>(tr, iv, a.scheme, a.h, eps[2 * i], eps[2 * i + 1], eps[2 * i + 2],
                                         eps + 2 * a.S + 1 + 4 * i, x, tg);
        fold(node + 1);
    }
}

// Marquardt-scaled step: solves (H + lambda diag(H)) delta = -g on the free parameters (p_mask: identified and
// not held on a bound) by an in-register Cholesky factorisation.  gH(i) reads element i of the current
// [gradient | Gauss-Newton upper triangle] record.  Returns false when the matrix is not numerically positive
// definite.
template <typename T, int KIND, class GH>
FES_HD bool ident_solve(GH gH, T lambda, int p_mask, T *delta) {
    using I = IdentDims<KIND>;
    constexpr int NPK = I::NPK;
    T L[I::NH];
    T rhs[NPK];
    int on_mask = 0;
    static_for<0, NPK>([&](auto id) {
        constexpr int d = decltype(id)::value;
        L[I::hidx(d, d)] = gH(NPK + I::hidx(d, d));
        if (((p_mask >> d) & 1) && L[I::hidx(d, d)] > T(0)) on_mask |= 1 << d;
    });
    static_for<0, NPK>([&](auto id) {
        constexpr int d = decltype(id)::value;
        const bool on = (on_mask >> d) & 1;
        static_for<d + 1, NPK>([&](auto ie) {
            constexpr int e = decltype(ie)::value;
            L[I::hidx(d, e)] = (on && ((on_mask >> e) & 1)) ? gH(NPK + I::hidx(d, e)) : T(0);
        });
        L[I::hidx(d, d)] = on ? L[I::hidx(d, d)] * (T(1) + lambda) : T(1);
        rhs[d] = on ? -gH(d) : T(0);
    });
    bool ok = true;
    // upper Cholesky H = U^T U, row by row
    static_for<0, NPK>([&](auto id) {
        constexpr int d = decltype(id)::value;
        T piv = L[I::hidx(d, d)];
        if (!(piv > T(0))) ok = false;
        const T inv = T(1) / xsqrt(piv > T(0) ? piv : T(1));
        static_for<d, NPK>([&](auto ie) {
            constexpr int e = decltype(ie)::value;
            L[I::hidx(d, e)] *= inv;
        });
        static_for<d + 1, NPK>([&](auto ir) {
            constexpr int r = decltype(ir)::value;
            static_for<r, NPK>([&](auto ie) {
                constexpr int e = decltype(ie)::value;
                L[I::hidx(r, e)] = xfma(-L[I::hidx(d, r)], L[I::hidx(d, e)], L[I::hidx(r, e)]);
            });
        });
    });
    // forward: U^T y = rhs
    static_for<0, NPK>([&](auto id) {
        constexpr int d = decltype(id)::value;
        T s = rhs[d];
        static_for<0, d>([&](auto ik) {
            constexpr int k = decltype(ik)::value;
            s = xfma(-L[I::hidx(k, d)], rhs[k], s);
        });
        rhs[d] = s / L[I::hidx(d, d)];
    });
    // backward: U delta = y
    static_for<0, NPK>([&](auto id) {
        constexpr int d = NPK - 1 - decltype(id)::value;
        T s = rhs[d];
        static_for<d + 1, NPK>([&](auto ik) {
            constexpr int k = decltype(ik)::value;
            s = xfma(-L[I::hidx(d, k)], delta[k], s);
        });
        delta[d] = s / L[I::hidx(d, d)];
    });
    return ok;
}

// the whole identification of candidate b ------------------------------------------------------------------
// Per-candidate scratch `acc` (element i at acc[i * stride]; a shared-memory column per thread on the device):
//   [0, np)                gradient sums of the trial point      [np, np + nh)   Gauss-Newton sums of the trial point
//   [np + nh, + 14)        trial parameters
// The CURRENT point lives in the output arrays (params_out, cost, grad, hess), so that only the integration state
// occupies registers inside the RK loop.
template <int KIND>
constexpr int ident_scratch() { return IdentDims<KIND>::NPK + IdentDims<KIND>::NH + NP; }

// the general (strided, per-candidate) row source of candidate b
template <typename T>
FES_HD IdentRows<T, false> ident_rows_strided(const IdentArgs<T> &a, int64_t b) {
    return {a.tab + b * a.tS, a.ctrl + b * a.cS, (int64_t)a.row * a.tB, a.tB, (int64_t)a.nu * a.cB, a.cB};
}

template <typename T, int KIND, bool UNIT = false>
FES_HD void ident_candidate(const IdentArgs<T> &a, const IdentRows<T, UNIT> &rows, int64_t b, const T *eps, T *acc, int stride) {
    using I = IdentDims<KIND>;
    constexpr int NPK = I::NPK;
    constexpr int NH = I::NH;
    T *pt = acc + (NPK + NH) * stride;
    auto cur_p = [&](int j) -> T & { return a.params_out[j * a.B + b]; };
#pragma unroll
    for (int j = 0; j < NP; ++j) pt[j * stride] = a.params[j * a.pB + b * a.pS];
    if (a.max_iter > 0) {  // the iteration starts inside the box
        static_for<0, NPK>([&](auto id) {
            constexpr int d = decltype(id)::value;
            constexpr int j = I::canon(d);
            if (!((a.p_mask >> d) & 1)) return;
            T v = pt[j * stride];
            if (a.lower != nullptr && v < a.lower[j]) v = a.lower[j];
            if (a.upper != nullptr && v > a.upper[j]) v = a.upper[j];
            pt[j * stride] = v;
        });
    }
    T cost;
    ident_eval<T, KIND, UNIT>(a, rows, b, pt, stride, eps, cost, acc, stride);
    auto publish = [&]() {  // the trial point becomes the current one
        a.cost[b] = cost;
        static_for<0, NPK + NH>([&](auto ic) {
            constexpr int i = decltype(ic)::value;
            (i < NPK ? a.grad[i * a.B + b] : a.hess[(i - NPK) * a.B + b]) = acc[i * stride];
        });
#pragma unroll
        for (int j = 0; j < NP; ++j) cur_p(j) = pt[j * stride];
    };
    publish();
    T cur = cost;
    T lambda = a.lambda0;
    int evals = 1;
    auto gH = [&](int i) { return i < NPK ? a.grad[i * a.B + b] : a.hess[(i - NPK) * a.B + b]; };
    for (int it = 0; it < a.max_iter; ++it) {
        if (!(cur > T(0)) || !(lambda < T(1e15))) break;  // exact fit, NaN start, or no acceptable step left
        // active set: a parameter sitting on a bound whose gradient pushes it further out is held for this step
        int free_mask = a.p_mask;
        static_for<0, NPK>([&](auto id) {
            constexpr int d = decltype(id)::value;
            constexpr int j = I::canon(d);
            const T gd = gH(d), pj = cur_p(j);
            if (a.lower != nullptr && pj <= a.lower[j] && gd > T(0)) free_mask &= ~(1 << d);
            if (a.upper != nullptr && pj >= a.upper[j] && gd < T(0)) free_mask &= ~(1 << d);
        });
        T delta[NPK];
        if (!ident_solve<T, KIND>(gH, lambda, free_mask, delta)) {
            lambda *= T(10);
            continue;
        }
        T step = T(0);
        static_for<0, NPK>([&](auto id) {
            constexpr int d = decltype(id)::value;
            constexpr int j = I::canon(d);
            // a parameter that is not identified is not projected on the box either (same rule as the initial projection):
            // its trial value stays the starting value even when that lies outside [lower, upper]
            if (!((a.p_mask >> d) & 1)) return;
            const T pj = cur_p(j);
            T v = pj + delta[d];
            if (a.lower != nullptr) v = v < a.lower[j] ? a.lower[j] : v;
            if (a.upper != nullptr) v = v > a.upper[j] ? a.upper[j] : v;
            const T rel = xabs(v - pj) / (xabs(pj) + T(1e-30));
            step = rel > step ? rel : step;
            pt[j * stride] = v;  // rows that are not P-columns keep the current (= starting) value
        });
        if (!(step > T(0))) break;  // pinned on the box in every identified direction
        ident_eval<T, KIND, UNIT>(a, rows, b, pt, stride, eps, cost, acc, stride);
        ++evals;
        if (cost < cur) {
            const T gain = cur - cost;
            publish();
            const bool done = gain <= a.ftol * cur || step <= a.xtol;
            cur = cost;
            lambda = lambda * T(1.0 / 3.0);
            if (lambda < T(1e-12)) lambda = T(1e-12);
            if (done) break;
        } else {
            lambda *= T(4);
        }
    }
    if (a.iters != nullptr) a.iters[b] = evals;
}

#if defined(__CUDACC__)

// Resident blocks the compiler must make room for.  6 blocks of 64 threads cap the kernel at 168 registers: the Hmed
// instance then spills ~1 KB per thread to (L1-resident) local memory, and still gains 4.5 % over the 253-register /
// 4-block build because 12 warps instead of 8 cover the dependent-issue latency (profiles/r01_tune_ident_t.txt).
#ifndef FES_IDENT_MIN_BLOCKS  // sweep knob (scripts/build_variants.sh)
#define FES_IDENT_MIN_BLOCKS 6
#endif
// STAGED: window tables and controls carry no batch axis (one stimulation train and one control sequence for all
// candidates -- the identification case): the block copies them into shared memory once, [ns][row] and [ns][nu], and
// every candidate reads its interval set-up from there.  Measured before (profiles/r01_ncu_ident_cfg4_ak_summary.tsv): 373
// warp instructions executed per trajectory-step against 211 in the RK4 path -- the per-slot loop of the Hmed2018 set-up
// (two table weights + one intensity per slot through 64-bit strided global addresses) was 43 % of the kernel.
constexpr size_t IDENT_STAGE_BUDGET = 8 * 1024;   // bytes of staged rows per block, on top of the 30 KB of running sums

template <typename T, int KIND, bool STAGED>
__global__ void __launch_bounds__(BLOCK, FES_IDENT_MIN_BLOCKS) fes_ident_kernel(const IdentArgs<T> a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T *s_eps = reinterpret_cast<T *>(smem_raw);
    T *s_acc = s_eps + ((stage_table_len(a.S) + 3) & ~3);  // [ident_scratch][BLOCK]
    T *s_tab = s_acc + ident_scratch<KIND>() * BLOCK;       // STAGED: [ns][row], then [ns][nu]
    T *s_ctl = s_tab + a.ns * a.row;
    for (int i = threadIdx.x; i < stage_table_len(a.S); i += BLOCK) s_eps[i] = a.eps[i];
    if constexpr (STAGED) {
        for (int i = threadIdx.x; i < a.ns * a.row; i += BLOCK) s_tab[i] = a.tab[i];
        for (int i = threadIdx.x; i < a.ns * a.nu; i += BLOCK) s_ctl[i] = a.ctrl[i];
    }
    __syncthreads();
    const int64_t b = (int64_t)blockIdx.x * BLOCK + threadIdx.x;
    if (b >= a.B) return;
    if constexpr (STAGED) {
        const IdentRows<T, true> rows = {s_tab, s_ctl, a.row, 1, a.nu, 1};
        ident_candidate<T, KIND, true>(a, rows, b, s_eps, s_acc + threadIdx.x, BLOCK);
    } else {
        ident_candidate<T, KIND, false>(a, ident_rows_strided(a, b), b, s_eps, s_acc + threadIdx.x, BLOCK);
    }
}

template <typename T, int KIND>
int launch_ident_k(const IdentArgs<T> &a, cudaStream_t s) {
    const unsigned grid = (unsigned)((a.B + BLOCK - 1) / BLOCK);
    size_t smem = (((stage_table_len(a.S) + 3) & ~3) + (size_t)ident_scratch<KIND>() * BLOCK) * sizeof(T);
    const size_t rows = (size_t)a.ns * (a.row + a.nu) * sizeof(T);
    const bool staged = a.tS == 0 && (a.nu == 0 || a.cS == 0) && rows <= IDENT_STAGE_BUDGET;
    if (staged) fes_ident_kernel<T, KIND, true><<<grid, BLOCK, smem + rows, s>>>(a);
    else fes_ident_kernel<T, KIND, false><<<grid, BLOCK, smem, s>>>(a);
    return (int)cudaGetLastError();
}

#endif  // __CUDACC__

}  // namespace fes
