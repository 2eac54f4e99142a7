// fes_core.cuh -- per-trajectory arithmetic of the FES muscle models and their tangent (forward
// sensitivity) propagation, written once as __host__ __device__ templates.
//
// What the reference computes (all paths under the cocofest tree):
//   S(t)  = sum_i R_i exp(-(t - t_i)/tauc) lambda_i                  models/ding2003.py:230-252
//   Cn'   = S/tauc - Cn/tauc                                         models/ding2003.py:254-266
//   F'    = A_eff q - F/(tau1 + tau2 q),  q = Cn/(Km + Cn)           models/ding2003.py:274-311
//   A_eff = A (1 - exp(-(pw - pd0)/pdt))                             models/ding2007.py:172-188
//   lambda_i = ar (tanh(bs (I_i - Is)) + cr)                         models/hmed2018.py:169-180
//   A', Tau1', Km' fatigue rows                                      models/ding2003_with_fatigue.py:197-240
//   RK1/RK2/RK4 with controls and window constant over an interval   external/bioptim/bioptim/dynamics/integrator.py:300-365
//
// How it is computed here (B200-first, not a translation): the window and the controls are constant
// over an interval, so S(t) factorises as S(t) = E_k * eps(t - t_k) with eps(s) = exp(-s/tauc).  E_k is
// ONE number per (trajectory, interval), assembled from per-interval window tables (fes_tables kernel);
// eps at the 2S+1 stage offsets is a table shared by every trajectory and interval.  The inner RK loop
// therefore contains no exp at all; the two divisions of F' share one reciprocal (MUFU.RCP64H + 5 DFMA,
// no slow path).  Sensitivities are the exact tangent of the discrete RK map (what casadi's forward AD of
// the unrolled integrator yields): the nominal stages record the few non-trivial Jacobian entries of F'
// and every direction column is advanced with them -- for RK4 through a step propagator shared by all columns
// (what the scheme does to the linear tangent rows, written out once per step: scalar for the two-state models,
// 2 x 2 for the reduced fatigue columns; rk_step_scheme), otherwise stage by stage.  Structural zeros (Cn
// decouples from every state, fatigue rows are linear) are removed at compile time through the direction
// code of each column.
#pragma once

#include <stdint.h>
#include <math.h>
#include <type_traits>

#if defined(__CUDACC__)
#define FES_HD __host__ __device__ __forceinline__
#else
#define FES_HD inline
#endif

namespace fes {

enum { DING2003 = 0, DING2007 = 1, HMED2018 = 2 };
enum { G_X0 = 1, G_U = 2, G_P = 4, G_W = 8, G_PF = 16, G_C = 32 };
// G_C: the force factor c = fl fv + fp of f_dot_fun (ding2003.py:309-311), F' = c (A q - F/(tau1 + tau2 q)), enters as a
// per-interval input (the row after the model's own controls) and gets one tangent column d / d c.  Kernels without G_C
// integrate c = 1 and do not execute a single instruction for it.
constexpr int TILE = 5;
constexpr int NP = 14;
enum {
    P_A_REST, P_KM_REST, P_TAU1_REST, P_TAU2, P_ALPHA_A, P_ALPHA_TAU1, P_ALPHA_KM, P_TAU_FAT,
    P_PD0, P_PDT, P_AR, P_BS, P_IS, P_CR
};

// direction codes: what a tangent column differentiates with respect to
enum {
    D_CN0, D_F0, D_A0, D_TAU10, D_KM0,          // initial state of the interval
    D_PW, D_INT,                                // controls: pulse width, pulse intensity of a window slot
    D_WT,                                       // pulse time of a window slot
    D_A_REST, D_KM_REST, D_TAU1_REST, D_TAU2,   // identifiable parameters (fes_identification_ocp.py:269-299)
    D_PD0, D_PDT, D_AR, D_BS, D_IS, D_CR,
    D_ALPHA_A, D_ALPHA_TAU1, D_ALPHA_KM, D_TAU_FAT,
    D_FC                                        // the force factor c of the interval (G_C)
};

template <int KIND, bool FAT, int GROUPS>
struct Dirs {
    static constexpr int NX = FAT ? 5 : 2;
    static constexpr int n_x0 = (GROUPS & G_X0) ? NX : 0;
    static constexpr int n_u = (GROUPS & G_U) ? (KIND == DING2007 ? 1 : (KIND == HMED2018 ? TILE : 0)) : 0;
    static constexpr int n_c = (GROUPS & G_C) ? 1 : 0;
    static constexpr int n_p = (GROUPS & G_P) ? (KIND == DING2003 ? 4 : (KIND == DING2007 ? 6 : 8)) : 0;
    static constexpr int n_pf = ((GROUPS & G_PF) && FAT) ? 4 : 0;
    static constexpr int n_w = (GROUPS & G_W) ? TILE : 0;
    static constexpr int ND = n_x0 + n_u + n_c + n_p + n_pf + n_w;
    static constexpr bool FC = (GROUPS & G_C) != 0;
    // window tables carry per-slot weights (not only their sums)
    static constexpr bool SLOT = (KIND == HMED2018) || (GROUPS & G_W);

    static constexpr int code(int d) {
        if (d < n_x0) return D_CN0 + d;
        d -= n_x0;
        if (d < n_u) return KIND == DING2007 ? D_PW : D_INT;
        d -= n_u;
        if (d < n_c) return D_FC;
        d -= n_c;
        if (d < n_p) {
            if (d < 4) return D_A_REST + d;
            return KIND == DING2007 ? D_PD0 + (d - 4) : D_AR + (d - 4);
        }
        d -= n_p;
        if (d < n_pf) return D_ALPHA_A + d;
        return D_WT;
    }
    // window-slot index inside the tile for D_INT / D_WT columns
    static constexpr int slot(int d) {
        if (d >= n_x0 && d < n_x0 + n_u) return d - n_x0;
        return d - (n_x0 + n_u + n_c + n_p + n_pf);
    }
};

constexpr bool dir_has_cn(int c) {
    return c == D_CN0 || c == D_INT || c == D_WT || c == D_KM_REST || c == D_AR || c == D_BS || c == D_IS || c == D_CR;
}
constexpr bool dir_has_dg(int c) { return c == D_PW || c == D_PD0 || c == D_PDT; }
// Columns whose Cn component is FORCED through the pulse sum (vCn' = (sE eps(t) - vCn)/tauc, sE constant over the
// interval).  That row is linear and identical for all of them, so over one RK step
//     vCn(stage s) = c dl_s + sE psi_s,     vCn(end) = c dl_NS + sE psi_NS        (c = vCn at the step start)
// with dl = the scheme's stage values of v' = -v/tauc from 1 (loop-invariant) and psi = those of
// psi' = (eps - psi)/tauc from 0 (shared by every column of the step).  Explicit RK maps are linear in (c, sE) on a
// linear row, so this IS the tangent of the discrete map; the column stores c only and its F-row takes two FMAs.
constexpr bool dir_forced_cn(int c) { return dir_has_cn(c) && c != D_CN0; }
// The shared modes cost ~20 instructions per step (psi 12, the two products per stage 8) and save ~10 per forced column:
// used from three such columns on (Hmed parameter / intensity tiles, pulse-time tiles); below that each column
// integrates its own Cn component (Ding2003 / Ding2007 km_rest column: 202 instead of 213 instructions per step).
template <class D>
constexpr bool dirs_forced_cn() {
    int n = 0;
    for (int d = 0; d < D::ND; ++d)
        if (dir_forced_cn(D::code(d))) ++n;
    return n >= 3;
}
// Fatigue models: the rows A', Tau1', Km' are linear with the SAME decay 1/tau_fat and are all driven by F
// alone, so the tangent (vA, vTau1, vKm) of EVERY direction stays in an affine family around one scalar W:
//     vA = alpha_a W + mA,  vTau1 = alpha_tau1 W + mT1,  vKm = alpha_km W + mKm,
//     W' = vF - W/tau_fat  (W = 0 at the seed),   m' = -m/tau_fat + f_d(t)  (m = the seeded component at the seed),
// where f_d is what the direction adds to the fatigue rows THEMSELVES -- it does not depend on the tangent:
//     initial states, controls, pulse times, tau2, pd0, pdt, ar..cr   f = 0:  m = the seed times the decaying mode mu
//     a_rest | tau1_rest | km_rest      f = 1/tau_fat on its own row:  m = 1 - mu            (the same mode again)
//     alpha_a | alpha_tau1 | alpha_km   f = F(t) on its own row:  m = M, M' = F - M/tau_fat  (ONE scalar for the three)
//     tau_fat                           f = (x - x_rest)/tau_fat^2 on the three rows: three scalars of its own
// Runge-Kutta maps commute with constant linear embeddings, so pushing (vCn, vF, W) and the few shared decaying
// scalars through the stages gives the tangent of the discrete map exactly (not an approximation of it);
// such a "reduced" column carries 2 or 3 numbers instead of 5 (tau_fat: 5).  The Cn0 column's own Cn component is
// another shared scalar (vCn' = -vCn/tauc), which leaves that column with (vF, W) as well.
constexpr bool dir_reduced(bool fat, int c) { return fat && c >= 0; }
constexpr bool dir_rest(int c) { return c == D_A_REST || c == D_TAU1_REST || c == D_KM_REST; }
constexpr bool dir_alpha(int c) { return c == D_ALPHA_A || c == D_ALPHA_TAU1 || c == D_ALPHA_KM; }
constexpr bool dir_own_m(int c) { return c == D_TAU_FAT; }  // mA, mT1, mKm kept in V[d][0], V[d][3], V[d][4]
template <class D>
constexpr bool dirs_any_forced_cn() {
    for (int d = 0; d < D::ND; ++d)
        if (dir_forced_cn(D::code(d))) return true;
    return false;
}
template <class D>
constexpr bool dirs_have(int c) {
    for (int d = 0; d < D::ND; ++d)
        if (D::code(d) == c) return true;
    return false;
}
// is component j of V[d][...] advanced stage by stage for a column of code c?  (sh: the forced Cn components of this
// direction set ride on the shared modes and are advanced once per step instead)
constexpr bool dir_stores(bool fat, int c, int j, bool sh) {
    if (j == 0) {
        if (fat && dir_own_m(c)) return true;
        if (fat && c == D_CN0) return false;  // the shared scalar nu
        return dir_has_cn(c) && !(sh && dir_forced_cn(c));
    }
    if (j == 1) return true;
    if (!fat) return false;
    return j == 2 || dir_own_m(c);
}

template <int I, int N, class F>
FES_HD void static_for(F &&f) {
    if constexpr (I < N) {
        f(std::integral_constant<int, I>{});
        static_for<I + 1, N>(f);
    }
}

// ---- stage table: shared by every trajectory and interval of a call -------------------------------
// [0, 2S]            eps_j  = exp(-(j h/2)/tauc), the decay of the pulse sum at the stage offsets
// [2S+1 + 4 i + q]   psi    = the scheme's own integration of psi' = (eps - psi)/tauc over step i of the interval from
//                    psi = 0: q = 0..2 its values at stages 1..3 (as many as the scheme has), q = 3 its value after
//                    the step.  Every Cn-like linear row, x' = (g eps - x)/tauc with g constant over the interval,
//                    is then  x(stage s) = x_i dl_s + g psi_s,  x_{i+1} = x_i dl_end + g psi_end  EXACTLY as the
//                    scheme would integrate it (explicit RK maps are linear on linear rows): the nominal Cn (g = E)
//                    and the forced Cn components of the tangent columns (g = sE) need no stages of their own.
constexpr int stage_table_len(int S) { return 6 * S + 1; }
FES_HD double stage_table_entry(int idx, int S, double h, double tauc, int scheme) {
    if (idx <= 2 * S) return exp(-(0.5 * h * (double)idx) / tauc);
    const int i = (idx - (2 * S + 1)) / 4, q = (idx - (2 * S + 1)) % 4;
    const double z = h / tauc;
    const double e0 = exp(-(h * (double)i) / tauc), e1 = exp(-(h * ((double)i + 0.5)) / tauc), e2 = exp(-(h * (double)(i + 1)) / tauc);
    if (scheme == 1) return q == 3 ? z * e0 : 0.0;
    const double p1 = 0.5 * z * e0;  // stage 1 value (RK2 midpoint and RK4)
    if (scheme == 2) return q == 0 ? p1 : (q == 3 ? z * (e1 - p1) : 0.0);
    const double k2 = e1 - p1, p2 = 0.5 * z * k2, k3 = e1 - p2, p3 = z * k3, k4 = e2 - p3;
    if (q == 0) return p1;
    if (q == 1) return p2;
    if (q == 2) return p3;
    return z * (e0 + 2.0 * k2 + 2.0 * k3 + k4) / 6.0;
}

// ---- scalar helpers ---------------------------------------------------------------------------
FES_HD double xfma(double a, double b, double c) { return fma(a, b, c); }
FES_HD float xfma(float a, float b, float c) { return fmaf(a, b, c); }
FES_HD double xexp(double a) { return exp(a); }
FES_HD float xexp(float a) { return expf(a); }
// tanh without the branch of the library version (|x| < 0.55: polynomial, else exp): lanes of a warp hold different
// candidates, so both sides were executed (~95 instructions per window slot).  th = (1 - t)/(1 + t), t = exp(-2|x|), has an
// ABSOLUTE error of ~1e-16 (near 0 the subtraction 1 - t is exact, t itself carries the rounding), which is what its uses
// need: th is only ever added to cr ~ 0.8 (lambda_i_calculation, hmed2018.py:179) or squared into 1 - th^2.
FES_HD double xrcp(double x);
FES_HD double xtanh(double a) {
    const double t = exp(-2.0 * fabs(a));
    return copysign((1.0 - t) * xrcp(1.0 + t), a);
}
FES_HD float xtanh(float a) { return tanhf(a); }
FES_HD double xsqrt(double a) { return sqrt(a); }
FES_HD float xsqrt(float a) { return sqrtf(a); }
FES_HD double xabs(double a) { return fabs(a); }
FES_HD float xabs(float a) { return fabsf(a); }

// reciprocal without the denormal / overflow slow path: the arguments here are products of
// physiological quantities, (Km + Cn) * (tau1 (Km + Cn) + tau2 Cn), far from the range limits.
// MUFU.RCP64H seeds ~20 bits; r (1 + e + e^2) leaves a relative error ~e^3 < 2^-58, below one ulp.
FES_HD double xrcp(double x) {
#if defined(__CUDA_ARCH__)
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    double e = fma(-x, r, 1.0);
    e = fma(e, e, e);
    r = fma(r, e, r);
#ifdef FES_RCP_LONG  // a second Newton step: measured no change of any parity figure (profiles/r01_rcp_short_margin.txt)
    e = fma(-x, r, 1.0);
    r = fma(r, e, r);
#endif
    return r;
#else
    return 1.0 / x;
#endif
}
FES_HD float xrcp(float x) {
#if defined(__CUDA_ARCH__)
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    float e = fmaf(-x, r, 1.0f);
    r = fmaf(r, e, r);
    return r;
#else
    return 1.0f / x;
#endif
}

// ---- per-trajectory constants ---------------------------------------------------------------
template <typename T>
struct Traj {
    T a_rest, km_rest, tau1_rest, tau2;
    T alpha_a, alpha_tau1, alpha_km, kappa, kappa2;  // kappa = 1/tau_fat
    T a_rest_k, tau1_rest_k, km_rest_k;              // rest value * kappa
    T r0m1;                                          // r0 - 1 = km_rest + r0_km - 1   (ding2003.py:145-147, :269-271)
    T pd0, pdt, ar, bs, Is, cr;
    T inv_tauc;
};

// p(j) returns canonical parameter j of this trajectory
template <typename T, class PF>
FES_HD void load_traj(Traj<T> &tr, PF p, T tauc, T r0_km) {
    tr.a_rest = p(P_A_REST);
    tr.km_rest = p(P_KM_REST);
    tr.tau1_rest = p(P_TAU1_REST);
    tr.tau2 = p(P_TAU2);
    tr.alpha_a = p(P_ALPHA_A);
    tr.alpha_tau1 = p(P_ALPHA_TAU1);
    tr.alpha_km = p(P_ALPHA_KM);
    // reciprocal without the IEEE-division slow path (once per (scenario, interval) in shoot); a non-finite or huge
    // tau_fat (fatigue switched off by the caller) takes the exact division
    const T tf = p(P_TAU_FAT);
    tr.kappa = (xabs(tf) <= T(3.0e38)) ? xrcp(tf) : T(1) / tf;
    tr.kappa2 = tr.kappa * tr.kappa;
    tr.a_rest_k = tr.a_rest * tr.kappa;
    tr.tau1_rest_k = tr.tau1_rest * tr.kappa;
    tr.km_rest_k = tr.km_rest * tr.kappa;
    tr.r0m1 = tr.km_rest + r0_km - T(1);
    tr.pd0 = p(P_PD0);
    tr.pdt = p(P_PDT);
    tr.ar = p(P_AR);
    tr.bs = p(P_BS);
    tr.Is = p(P_IS);
    tr.cr = p(P_CR);
    tr.inv_tauc = xrcp(tauc);
}

// ---- per-interval quantities ------------------------------------------------------------------
template <typename T, int ND>
struct Interval {
    T E;                       // Cn forcing amplitude: S(t) = E * eps(t - t_k)
    T g;                       // pulse-width scaling of A (ding2007), 1 otherwise
    T fc;                      // force factor fl fv + fp of the interval (G_C), 1 otherwise
    T sE[ND > 0 ? ND : 1];     // d E / d direction
    T dg[ND > 0 ? ND : 1];     // d g / d direction
};

// Window table row of one node (fes_tables kernel):  [a, b]  or  [a, b, wa_0..wa_{N-1}, wb_0..wb_{N-1}]
//   wa_i = exp(-(t_k - t_i)/tauc),  wb_i = exp(-(t_i - t_{i-1})/tauc) wa_i (0 for i = 0),  a = sum wa, b = sum wb
//   so that  E = sum_i lambda_i (wa_i + (r0 - 1) wb_i)   [= a + (r0 - 1) b when lambda = 1].
// tab(c) / ctl(c) return element c of this trajectory's row for the current node.
// Hmed2018: tanh(bs (I - Is)) of the window slots of the previous interval, keyed by the intensity itself.  The
// window slides by at most a few pulses per node, so slot i of a node usually holds the intensity that slot i or
// i + 1 held at the previous node; its tanh (the one expensive operation of the set-up, N of them per interval) is
// then reused -- the very same number, whatever pulse it belongs to.  Lives in per-thread local memory; reset
// whenever bs or Is change (every cost evaluation of fes_identify).
template <typename T>
struct LamCache {
    T I[66], th[66];
    FES_HD void reset(int N) {
        for (int i = 0; i <= N; ++i) I[i] = T(NAN);  // NaN never compares equal: nothing cached
    }
};

template <typename T, int KIND, bool FAT, int GROUPS, class TF, class CF>
FES_HD void interval_setup(const Traj<T> &tr, TF tab, CF ctl, int N, int tile_off,
                           Interval<T, Dirs<KIND, FAT, GROUPS>::ND> &iv, LamCache<T> *lc = nullptr) {
    using D = Dirs<KIND, FAT, GROUPS>;
    constexpr int ND = D::ND;
    static_for<0, ND>([&](auto ic) {
        constexpr int d = decltype(ic)::value;
        iv.sE[d] = T(0);
        iv.dg[d] = T(0);
    });
    iv.g = T(1);
    iv.fc = T(1);
    if constexpr (D::FC) iv.fc = ctl(KIND == DING2007 ? 1 : (KIND == HMED2018 ? N : 0));  // the row after the model's controls
    if constexpr (KIND == DING2007) {
        // a_calculation, ding2007.py:188; the control is the width of the LAST pulse (ding2007.py:280)
        const T pw = ctl(0);
        const T ipdt = xrcp(tr.pdt);
        const T z = (pw - tr.pd0) * ipdt;
        const T ew = xexp(-z);
        iv.g = T(1) - ew;
        static_for<0, ND>([&](auto ic) {
            constexpr int d = decltype(ic)::value;
            constexpr int c = D::code(d);
            if constexpr (c == D_PW) iv.dg[d] = ew * ipdt;
            if constexpr (c == D_PD0) iv.dg[d] = -(ew * ipdt);
            if constexpr (c == D_PDT) iv.dg[d] = -(ew * z * ipdt);
        });
    }
    if constexpr (!D::SLOT) {
        const T a = tab(0), b = tab(1);
        iv.E = xfma(tr.r0m1, b, a);
        static_for<0, ND>([&](auto ic) {
            constexpr int d = decltype(ic)::value;
            if constexpr (D::code(d) == D_KM_REST) iv.sE[d] = b;  // d r0 / d km_rest = 1
        });
    } else {
        T E = T(0);
        for (int i = 0; i < N; ++i) {
            const T wa = tab(2 + i), wb = tab(2 + N + i);
            // a pad slot (-10 000 000, ding2003.py:394-397) has wa = wb = 0 exactly and adds exact zeros to every sum
            // below: skipped, with its tanh (early nodes of a train are mostly pads)
            if (wa == T(0) && wb == T(0)) continue;
            const T W = xfma(tr.r0m1, wb, wa);
            T lam = T(1), th = T(0), du = T(0);
            if constexpr (KIND == HMED2018) {
                const T Ii = ctl(i);
                du = Ii - tr.Is;
                bool hit = false;
                if (lc != nullptr) {
                    if (lc->I[i] == Ii) {
                        th = lc->th[i];
                        hit = true;
                    } else if (lc->I[i + 1] == Ii) {
                        th = lc->th[i + 1];
                        hit = true;
                    }
                }
                if (!hit) th = xtanh(tr.bs * du);  // lambda_i_calculation, hmed2018.py:179
                if (lc != nullptr) {
                    lc->I[i] = Ii;
                    lc->th[i] = th;
                }
                lam = tr.ar * (th + tr.cr);
            }
            E = xfma(lam, W, E);
            const T sech2 = T(1) - th * th;
            static_for<0, ND>([&](auto ic) {
                constexpr int d = decltype(ic)::value;
                constexpr int c = D::code(d);
                if constexpr (c == D_KM_REST) iv.sE[d] = xfma(lam, wb, iv.sE[d]);
                if constexpr (c == D_AR) iv.sE[d] = xfma(th + tr.cr, W, iv.sE[d]);
                if constexpr (c == D_BS) iv.sE[d] = xfma(tr.ar * sech2 * du, W, iv.sE[d]);
                if constexpr (c == D_IS) iv.sE[d] = xfma(-(tr.ar * tr.bs * sech2), W, iv.sE[d]);
                if constexpr (c == D_CR) iv.sE[d] = xfma(tr.ar, W, iv.sE[d]);
                if constexpr (c == D_INT) {
                    if (i == tile_off + D::slot(d)) iv.sE[d] = tr.ar * tr.bs * sech2 * W;
                }
                if constexpr (c == D_WT) {
                    // d E / d t_j = (lambda_j wa_j + (r0 - 1) lambda_{j+1} wb_{j+1}) / tauc
                    if (i == tile_off + D::slot(d)) iv.sE[d] = xfma(lam * wa, tr.inv_tauc, iv.sE[d]);
                    if (i == tile_off + D::slot(d) + 1) iv.sE[d] = xfma(tr.r0m1 * lam * wb, tr.inv_tauc, iv.sE[d]);
                }
            });
        }
        iv.E = E;
    }
}

// ---- one RK stage of the nominal trajectory ------------------------------------------------
template <typename T>
struct Stage {  // the non-trivial entries of d f / d (x, params) at a stage point
    T cF, cCn, cKm, cT1, cA, cT2, cG, dA, dT1, dKm;
    T cW;  // fatigue: d F' / d W = cA alpha_a + cT1 alpha_tau1 + cKm alpha_km  (reduced columns)
    T fU;  // G_C: F' / c, the forcing of the d / d c column
};

// FC: the F row and every entry of its Jacobian carry the interval's force factor fc (f_dot_fun, ding2003.py:309-311)
template <typename T, int KIND, bool FAT, bool NEED, bool FC = false>
FES_HD void nominal_rhs(const Traj<T> &tr, T Ee, T g, const T *x, T *k, Stage<T> &st, T fc = T(1)) {
    const T cn = x[0], f = x[1];
    const T A = FAT ? x[2] : tr.a_rest;
    const T t1 = FAT ? x[3] : tr.tau1_rest;
    const T km = FAT ? x[4] : tr.km_rest;
    k[0] = xfma(-tr.inv_tauc, cn, Ee);       // Ee already carries the 1/tauc factor (rk_step_scheme)
    const T s = km + cn;
    const T M = xfma(tr.tau2, cn, t1 * s);  // (tau1 + tau2 q) * s
    const T r = xrcp(s * M);
    const T is = r * M;                      // 1 / (Km + Cn)
    const T invD1 = s * (r * s);             // 1 / (tau1 + tau2 q)
    const T q = cn * is;
    const T Aeff1 = (KIND == DING2007) ? A * g : A;
    if constexpr (FC && NEED) st.fU = xfma(Aeff1, q, -(f * invD1));
    // with a force factor the two coefficients of the F row carry it, and with them every Jacobian entry below
    const T invD = FC ? invD1 * fc : invD1;  // c / (tau1 + tau2 q)
    const T Aeff = FC ? Aeff1 * fc : Aeff1;
    k[1] = xfma(Aeff, q, -(f * invD));
    if constexpr (FAT) {
        k[2] = xfma(tr.alpha_a, f, xfma(-tr.kappa, A, tr.a_rest_k));
        k[3] = xfma(tr.alpha_tau1, f, xfma(-tr.kappa, t1, tr.tau1_rest_k));
        k[4] = xfma(tr.alpha_km, f, xfma(-tr.kappa, km, tr.km_rest_k));
    }
    if constexpr (NEED) {
        st.cF = -invD;
        const T ct1 = f * invD * invD1;      // d F' / d tau1
        st.cT1 = ct1;
        const T Gc = xfma(ct1, tr.tau2, Aeff);  // d F' / d q
        const T u = Gc * is;                    // d q / d Cn = (1 - q)/(Km + Cn),  d q / d Km = -q/(Km + Cn)
        st.cKm = -(u * q);
        st.cCn = u + st.cKm;
        st.cA = (KIND == DING2007) ? q * g : q;
        st.cT2 = ct1 * q;
        st.cG = A * q;
        if constexpr (FC) {
            st.cA *= fc;
            st.cG *= fc;
        }
        if constexpr (FAT) {
            st.cW = xfma(st.cA, tr.alpha_a, xfma(st.cT1, tr.alpha_tau1, st.cKm * tr.alpha_km));
            st.dA = (A - tr.a_rest) * tr.kappa2;   // d/d tau_fat of -(A - A_rest)/tau_fat
            st.dT1 = (t1 - tr.tau1_rest) * tr.kappa2;
            st.dKm = (km - tr.km_rest) * tr.kappa2;
        }
    }
}

// ---- one RK stage of one tangent column of a TWO-STATE model (Cn, F) ---------------------------
// pc_dl = cCn dl_s, pc_psi = cCn psi_s (forced-Cn columns, see dir_forced_cn; S0: first stage, psi_0 = 0);
// v[0] of such a column is its Cn component at the step start and k[0] is not produced.
// SH: the forced Cn components of this direction set use the shared modes.
// (Every column of a fatigue model goes through column_rhs_red below.)
template <typename T, int KIND, int CODE, bool S0, bool SH>
FES_HD void column_rhs(const Traj<T> &tr, const Stage<T> &st, T eps, T sE, T dg, T pc_dl, T pc_psi, const T *v, T *k) {
    constexpr bool OWN = CODE == D_CN0 || (dir_forced_cn(CODE) && !SH);  // integrates its own Cn component
    if constexpr (CODE == D_CN0) k[0] = -(v[0] * tr.inv_tauc);
    else if constexpr (OWN) k[0] = xfma(sE, eps, -v[0]) * tr.inv_tauc;
    else k[0] = T(0);
    T kf = st.cF * v[1];
    if constexpr (OWN) kf = xfma(st.cCn, v[0], kf);
    if constexpr (dir_forced_cn(CODE) && SH) {
        kf = xfma(pc_dl, v[0], kf);
        if constexpr (!S0) kf = xfma(pc_psi, sE, kf);
    }
    if constexpr (CODE == D_A_REST) kf += st.cA;
    if constexpr (CODE == D_TAU1_REST) kf += st.cT1;
    if constexpr (CODE == D_KM_REST) kf += st.cKm;
    if constexpr (CODE == D_TAU2) kf += st.cT2;
    if constexpr (CODE == D_FC) kf += st.fU;
    if constexpr (dir_has_dg(CODE)) kf = xfma(st.cG, dg, kf);
    k[1] = kf;
}

// ---- one RK stage of one REDUCED tangent column (fatigue models, see dir_reduced) ------------
// v = (vCn | mA, vF, W, mT1, mKm); mu / nu = the shared decaying scalars (fatigue mode, Cn mode) and Ms = the shared
// alpha-column scalar M and rest-column scalar R, all at this stage point.
template <typename T, int KIND, int CODE, bool S0, bool SH>
FES_HD void column_rhs_red(const Traj<T> &tr, const Stage<T> &st, T eps, T sE, T dg, T pc_dl, T pc_psi, T mu, T nu, T Ms,
                           T Rs, const T *v, T *k) {
    T kf = xfma(st.cW, v[2], st.cF * v[1]);
    if constexpr (CODE == D_CN0) {
        kf = xfma(st.cCn, nu, kf);
    } else if constexpr (dir_forced_cn(CODE) && SH) {
        kf = xfma(pc_dl, v[0], kf);
        if constexpr (!S0) kf = xfma(pc_psi, sE, kf);
    } else if constexpr (dir_forced_cn(CODE)) {
        k[0] = xfma(sE, eps, -v[0]) * tr.inv_tauc;
        kf = xfma(st.cCn, v[0], kf);
    }
    if constexpr (CODE == D_A0) kf = xfma(st.cA, mu, kf);
    if constexpr (CODE == D_TAU10) kf = xfma(st.cT1, mu, kf);
    if constexpr (CODE == D_KM0) kf = xfma(st.cKm, mu, kf);
    if constexpr (dir_rest(CODE)) kf = xfma(CODE == D_A_REST ? st.cA : (CODE == D_TAU1_REST ? st.cT1 : st.cKm), Rs, kf);
    if constexpr (dir_alpha(CODE)) kf = xfma(CODE == D_ALPHA_A ? st.cA : (CODE == D_ALPHA_TAU1 ? st.cT1 : st.cKm), Ms, kf);
    if constexpr (dir_own_m(CODE)) {
        kf = xfma(st.cA, v[0], xfma(st.cT1, v[3], xfma(st.cKm, v[4], kf)));
        k[0] = xfma(-tr.kappa, v[0], st.dA);
        k[3] = xfma(-tr.kappa, v[3], st.dT1);
        k[4] = xfma(-tr.kappa, v[4], st.dKm);
    }
    if constexpr (CODE == D_TAU2) kf += st.cT2;
    if constexpr (CODE == D_FC) kf += st.fU;
    if constexpr (dir_has_dg(CODE)) kf = xfma(st.cG, dg, kf);
    k[1] = kf;
    k[2] = xfma(-tr.kappa, v[2], v[1]);
}

// ---- tangent state of one trajectory ----------------------------------------------------------
// V[d] holds the stored components of column d (dir_stores); mu, nu are the two shared decaying
// scalars of the reduced columns.  They are kept in double in the f32 build too: a float cannot represent
// 1 - h/tau_fat (h/tau_fat ~ 2e-5) to better than 0.3 % of the decay rate, and the fp64 pipe is idle there.
template <typename T, int KIND, bool FAT, int GROUPS>
struct Tangent {
    using D = Dirs<KIND, FAT, GROUPS>;
    T V[D::ND > 0 ? D::ND : 1][D::NX];
    double mu, nu;
    T M;  // shared by the alpha_a / alpha_tau1 / alpha_km columns: M' = F - M/tau_fat, 0 at the seed
    T R;  // shared by the a_rest / tau1_rest / km_rest columns:     R' = (1 - R)/tau_fat, 0 at the seed (= 1 - mu, kept on
          // its own so that it is exact near the seed, where 1 - mu cancels)

    // f32 build, fatigue models: A, Tau1, Km are ACCUMULATED in double.  Their per-step increments (alpha_X F h ~ 1e-4 of the
    // state) sit a few bits above a float's resolution of the state, so a float accumulator loses ~3e-4 of every increment:
    // measured 9e-5 relative drift of A after 20 s of the fatigue protocol, growing linearly.  The increments themselves are
    // formed in float; only the three running sums (3 DADD per step on the otherwise idle fp64 pipe) are double.
    // (Pairs of floats with TwoSum instead -- no F64 <-> F32 conversions, 9 fewer XU-pipe instructions but 30 more FP32
    // ones per step -- were built and measured: 1064.8 against 1045.0 ms, profiles/r02_tune_f32_float_pair_acc_ak.json.  The
    // float kernel is bound by its FP32 issue slots; conversions on the XU pipe are close to free.)
    double xd[3];

    // x = the state the integration starts from (seeds the double accumulators of the f32 build)
    FES_HD void init(const T *x) {
        if constexpr (FAT && sizeof(T) == 4) {
            xd[0] = (double)x[2];
            xd[1] = (double)x[3];
            xd[2] = (double)x[4];
        }
        static_for<0, D::ND>([&](auto ic) {
            constexpr int d = decltype(ic)::value;
            constexpr int C = D::code(d);
#pragma unroll
            for (int j = 0; j < D::NX; ++j) {
                if constexpr (dir_reduced(FAT, C)) V[d][j] = (C == D_F0 && j == 1) ? T(1) : T(0);
                else V[d][j] = (d < D::n_x0 && j == d) ? T(1) : T(0);
            }
        });
        mu = 1.0;
        nu = 1.0;
        M = T(0);
        R = T(0);
    }
    // component j of the actual tangent column d (what the caller sees)
    template <int d>
    FES_HD T get(const Traj<T> &tr, int j) const {
        constexpr int C = D::code(d);
        if constexpr (!dir_reduced(FAT, C)) {
            return V[d][j];
        } else {
            if (j == 0) return C == D_CN0 ? (T)nu : ((dir_has_cn(C) && !dir_own_m(C)) ? V[d][0] : T(0));
            if (j == 1) return V[d][1];
            // the component the direction seeds or forces on the fatigue row j (see dir_reduced)
            T m = T(0);
            if constexpr (C == D_A0 || C == D_TAU10 || C == D_KM0) m = (j == 2 + (C - D_A0)) ? (T)mu : T(0);
            if constexpr (dir_rest(C)) m = (j == (C == D_A_REST ? 2 : (C == D_TAU1_REST ? 3 : 4))) ? R : T(0);
            if constexpr (dir_alpha(C)) m = (j == 2 + (C - D_ALPHA_A)) ? M : T(0);
            if constexpr (dir_own_m(C)) m = j == 2 ? V[d][0] : V[d][j];
            const T al = j == 2 ? tr.alpha_a : (j == 3 ? tr.alpha_tau1 : tr.alpha_km);
            return xfma(al, V[d][2], m);
        }
    }
};

// ---- one RK step of the state and of every tangent column ------------------------------------
// Butcher data of the three explicit schemes bioptim offers (integrator.py:324-365), in units of h:
//   stage s is evaluated at x + alpha_s h k_{s-1};  the step adds beta_s h k_s.
template <int SCHEME> struct Tableau;
template <> struct Tableau<1> {
    static constexpr int NS = 1;
    static constexpr double alpha(int) { return 0.0; }
    static constexpr double beta(int) { return 1.0; }
    static constexpr int eps_at(int) { return 0; }
};
template <> struct Tableau<2> {  // midpoint
    static constexpr int NS = 2;
    static constexpr double alpha(int s) { return s == 1 ? 0.5 : 0.0; }
    static constexpr double beta(int s) { return s == 1 ? 1.0 : 0.0; }
    static constexpr int eps_at(int s) { return s; }
};
template <> struct Tableau<4> {
    static constexpr int NS = 4;
    static constexpr double alpha(int s) { return s == 0 ? 0.0 : (s == 3 ? 1.0 : 0.5); }
    static constexpr double beta(int s) { return (s == 0 || s == 3) ? 1.0 / 6.0 : 1.0 / 3.0; }
    static constexpr int eps_at(int s) { return s == 0 ? 0 : (s == 3 ? 2 : 1); }
};

// Straight-line code, stage-major: the nominal stage s produces the Jacobian entries that every column's
// stage s consumes, so only ONE Stage record is live at a time and the columns (independent of each other)
// sit in one basic block where the compiler interleaves their FMA chains.
// e[0], e[1], e[2] = eps at the step start, midpoint and end (relative to the interval start).
template <typename T, int KIND, bool FAT, int GROUPS, int SCHEME>
FES_HD void rk_step_scheme(const Traj<T> &tr, const Interval<T, Dirs<KIND, FAT, GROUPS>::ND> &iv, T h, const T *e,
                           const T *ps, T *x, Tangent<T, KIND, FAT, GROUPS> &tg) {
    using D = Dirs<KIND, FAT, GROUPS>;
    using TB = Tableau<SCHEME>;
    constexpr int NX = D::NX;
    constexpr int ND = D::ND;
    constexpr int NS = TB::NS;
    constexpr bool NEED = ND > 0;
    constexpr bool ANY_REST = dirs_have<D>(D_A_REST) || dirs_have<D>(D_TAU1_REST) || dirs_have<D>(D_KM_REST);
    constexpr bool ANY_ALPHA = dirs_have<D>(D_ALPHA_A) || dirs_have<D>(D_ALPHA_TAU1) || dirs_have<D>(D_ALPHA_KM);
    constexpr bool MODES = FAT && (D::n_x0 > 0);  // the shared decaying scalars mu, nu are advanced
    T xs[NX], xn[NX];
#pragma unroll
    for (int j = 0; j < NX; ++j) xs[j] = xn[j] = x[j];
    T vs[ND > 0 ? ND : 1][NX], vn[ND > 0 ? ND : 1][NX];
    static_for<0, ND>([&](auto ic) {
        constexpr int d = decltype(ic)::value;
#pragma unroll
        for (int j = 0; j < NX; ++j) vs[d][j] = vn[d][j] = tg.V[d][j];
    });
    // the two shared decaying scalars of the reduced columns obey m' = -(z / h) m: same scheme, in double
    const double zf = (double)tr.kappa * (double)h, zc = (double)tr.inv_tauc * (double)h;
    double mus = tg.mu, mun = tg.mu, nus = tg.nu, nun = tg.nu;
    // Cn' = (E eps - Cn)/tauc is a linear row with tabulated forcing: Cn(stage s) = Cn dl_s + E psi_s (stage table)
    // Nominal fatigue rows, reduced within the step exactly like the tangent ones: with the offsets d = x - rest at
    // the step start,  A(stage) = A_rest + dA m_s + alpha_a Wn = A - dA (1 - m_s) + alpha_a Wn,  Wn' = F - Wn/tau_fat (Wn = 0 at the step start),
    // m_s = the scheme's stage values of m' = -m/tau_fat from m = 1 (loop-invariant, hoisted by the compiler).
    constexpr int NXN = FAT ? 2 : NX;  // nominal components advanced by the generic update below
    T fm[NS + 1];                      // fm[s] = m at stage s, fm[NS] = m after the step
    T dA = T(0), dT1 = T(0), dKm = T(0), wn_s = T(0), wn_n = T(0);
    T wnv[NS + 1];  // Wn at the stage points (wnv[0] = 0) and after the step: the forcing of the alpha / tau_fat columns
    wnv[0] = T(0);
    T em[NS + 1];  // em[s] = 1 - fm[s], formed WITHOUT the cancellation (a float holds z = h/tau_fat ~ 2e-5 exactly enough,
                   // but not 1 - z): the modes that start at 0 (R, M, the tau_fat scalars) are advanced with em
    if constexpr (FAT) {
        const T z = tr.kappa * h;
        fm[0] = T(1);
        em[0] = T(0);
        T acc = T(1), eacc = T(0);
        static_for<0, NS>([&](auto is) {
            constexpr int s = decltype(is)::value;
            if constexpr (TB::beta(s) != 0.0) {
                acc = xfma(-z * T(TB::beta(s)), fm[s], acc);
                eacc = xfma(z * T(TB::beta(s)), fm[s], eacc);
            }
            if constexpr (s + 1 < NS) {
                fm[s + 1] = xfma(-z * T(TB::alpha(s + 1)), fm[s], T(1));
                em[s + 1] = (z * T(TB::alpha(s + 1))) * fm[s];
            }
        });
        fm[NS] = acc;
        em[NS] = eacc;
        if constexpr (sizeof(T) == 4) {
            dA = (T)(tg.xd[0] - (double)tr.a_rest);
            dT1 = (T)(tg.xd[1] - (double)tr.tau1_rest);
            dKm = (T)(tg.xd[2] - (double)tr.km_rest);
        } else {
            dA = x[2] - tr.a_rest;
            dT1 = x[3] - tr.tau1_rest;
            dKm = x[4] - tr.km_rest;
        }
    }

    // dl[s] = the scheme's stage values of v' = -v/tauc from 1 (loop-invariant); ps[s-1] = psi at stage s and ps[3] =
    // psi after the step come from the stage table (stage_table_entry)
    constexpr bool CNF = dirs_forced_cn<D>();
    T dl[NS + 1];
    {
        const T zt = tr.inv_tauc * h;
        dl[0] = T(1);
        T acc = T(1);
        static_for<0, NS>([&](auto is) {
            constexpr int s = decltype(is)::value;
            if constexpr (TB::beta(s) != 0.0) acc = xfma(-zt * T(TB::beta(s)), dl[s], acc);
            if constexpr (s + 1 < NS) dl[s + 1] = xfma(-zt * T(TB::alpha(s + 1)), dl[s], T(1));
        });
        dl[NS] = acc;
    }

    auto finish_fatigue = [&]() {  // the fatigue states after the step: x - em d + alpha_X Wn
        if constexpr (FAT && sizeof(T) == 4) {
            tg.xd[0] += (double)xfma(tr.alpha_a, wn_n, -(em[NS] * dA));
            tg.xd[1] += (double)xfma(tr.alpha_tau1, wn_n, -(em[NS] * dT1));
            tg.xd[2] += (double)xfma(tr.alpha_km, wn_n, -(em[NS] * dKm));
            x[2] = (T)tg.xd[0];
            x[3] = (T)tg.xd[1];
            x[4] = (T)tg.xd[2];
        } else if constexpr (FAT) {
            x[2] = xfma(tr.alpha_a, wn_n, xfma(dA, fm[NS], tr.a_rest));
            x[3] = xfma(tr.alpha_tau1, wn_n, xfma(dT1, fm[NS], tr.tau1_rest));
            x[4] = xfma(tr.alpha_km, wn_n, xfma(dKm, fm[NS], tr.km_rest));
        }
    };
    // the two halves of a stage: the nominal RHS (fills the Jacobian record st) and the tangent columns (consume it)
    auto nominal_stage = [&](auto is, Stage<T> &st) {
        constexpr int s = decltype(is)::value;
        constexpr bool LAST = s == NS - 1;
        const T hb = h * T(TB::beta(s));
        const T ha = h * T(LAST ? 0.0 : TB::alpha(s + 1));
        {
            T k[NX];
            nominal_rhs<T, KIND, FAT, NEED, D::FC>(tr, T(0), iv.g, xs, k, st, iv.fc);  // k[0] is not used: Cn rides on the stage table
            if constexpr (FAT) {  // before xs[1] moves on: Wn' at this stage (k[2..4] of nominal_rhs are dead here)
                const T kw = s == 0 ? xs[1] : xfma(-tr.kappa, wn_s, xs[1]);
                if constexpr (TB::beta(s) != 0.0) wn_n = s == 0 ? hb * kw : xfma(hb, kw, wn_n);
                if constexpr (!LAST) {
                    wn_s = ha * kw;
                    wnv[s + 1] = wn_s;
                    // f32 build: rest + d fm_s = x - em_s d, the decay applied as a small correction of the state itself (em = 1 - fm
                    // is exact in a float, fm is not: its rounding would bias the recovery rate by up to 0.1 %)
                    if constexpr (sizeof(T) == 4) {
                        xs[2] = xfma(tr.alpha_a, wn_s, xfma(-em[s + 1], dA, x[2]));
                        xs[3] = xfma(tr.alpha_tau1, wn_s, xfma(-em[s + 1], dT1, x[3]));
                        xs[4] = xfma(tr.alpha_km, wn_s, xfma(-em[s + 1], dKm, x[4]));
                    } else {  // fp64 holds fm to 1e-16: the direct form keeps em out of the register budget
                        xs[2] = xfma(tr.alpha_a, wn_s, xfma(dA, fm[s + 1], tr.a_rest));
                        xs[3] = xfma(tr.alpha_tau1, wn_s, xfma(dT1, fm[s + 1], tr.tau1_rest));
                        xs[4] = xfma(tr.alpha_km, wn_s, xfma(dKm, fm[s + 1], tr.km_rest));
                    }
                }
            }
            if constexpr (!LAST) xs[0] = xfma(x[0], dl[s + 1], iv.E * ps[s]);
#pragma unroll
            for (int j = 1; j < NXN; ++j) {
                if constexpr (TB::beta(s) != 0.0) xn[j] = xfma(hb, k[j], xn[j]);
                if constexpr (!LAST) xs[j] = xfma(ha, k[j], x[j]);
            }
        }
    };
    auto column_stage = [&](auto is, const Stage<T> &st) {
        constexpr int s = decltype(is)::value;
        constexpr bool LAST = s == NS - 1;
        const T hb = h * T(TB::beta(s));
        const T ha = h * T(LAST ? 0.0 : TB::alpha(s + 1));
        const T eps = e[TB::eps_at(s)];
        T mu_s = T(0), nu_s = T(0), M_s = T(0), R_s = T(0);
        // M(s) = M fm_s + Wn(s),  R(s) = R fm_s + (1 - fm_s), written with em = 1 - fm
        if constexpr (FAT && ANY_ALPHA) M_s = s == 0 ? tg.M : xfma(-em[s], tg.M, tg.M) + wnv[s];
        if constexpr (FAT && ANY_REST) R_s = s == 0 ? tg.R : xfma(em[s], T(1) - tg.R, tg.R);
        if constexpr (MODES && sizeof(T) == 8) {
            // the decaying scalars are the step-start values times the loop-invariant stage factors fm / dl
            mu_s = s == 0 ? (T)tg.mu : (T)tg.mu * fm[s];
            nu_s = s == 0 ? (T)tg.nu : (T)tg.nu * dl[s];
        } else if constexpr (MODES) {  // f32 build: the stage factors would be floats, keep the double recurrences
            mu_s = (T)mus;
            nu_s = (T)nus;
            if constexpr (TB::beta(s) != 0.0) {
                mun = fma(-zf * TB::beta(s), mus, mun);
                nun = fma(-zc * TB::beta(s), nus, nun);
            }
            if constexpr (!LAST) {
                mus = fma(-zf * TB::alpha(s + 1), mus, tg.mu);
                nus = fma(-zc * TB::alpha(s + 1), nus, tg.nu);
            }
        }
        T pc_dl = T(0), pc_psi = T(0);
        if constexpr (CNF) {
            pc_dl = s == 0 ? st.cCn : st.cCn * dl[s];
            if constexpr (s > 0) pc_psi = st.cCn * ps[s - 1];
        }
        static_for<0, ND>([&](auto ic) {
            constexpr int d = decltype(ic)::value;
            constexpr int C = D::code(d);
            // stored components of this column: Cn only for the directions that reach it; (vF, W) when reduced;
            // a forced Cn component (vs[d][0] = its value at the step start) is advanced once, after the stages
            T k[NX];
            if constexpr (dir_reduced(FAT, C))
                column_rhs_red<T, KIND, C, s == 0, CNF>(tr, st, eps, iv.sE[d], iv.dg[d], pc_dl, pc_psi, mu_s, nu_s, M_s, R_s, vs[d], k);
            else
                column_rhs<T, KIND, C, s == 0, CNF>(tr, st, eps, iv.sE[d], iv.dg[d], pc_dl, pc_psi, vs[d], k);
            static_for<0, NX>([&](auto ij) {
                constexpr int j = decltype(ij)::value;
                if constexpr (dir_stores(FAT, C, j, CNF)) {
                    if constexpr (TB::beta(s) != 0.0) vn[d][j] = xfma(hb, k[j], vn[d][j]);
                    if constexpr (!LAST) vs[d][j] = xfma(ha, k[j], tg.V[d][j]);
                }
            });
        });
    };
#ifndef FES_NO_PROPAGATOR  // measured against the stage-by-stage columns: config 3 +20 %, config 4 +15 %, fes_identify +13 %
    constexpr bool PROP = !FAT && ND > 0 && SCHEME == 4;  // (profiles/r01_tune_propagator_af.txt); RK1 / RK2 keep the stages
#else
    constexpr bool PROP = false;
#endif
    if constexpr (PROP) {
        // Two-state models: the F-row of EVERY column is the same scalar linear ODE v' = cF(t) v + g_d(t), so one RK4 step
        // is  v_end = Phi v + sum_s W_s g_d(s)  with Phi, W_s functions of cF at the four stages only -- shared by all
        // columns (what the scheme does to a linear row, written out once).  The Cn components come from the stage table.
        Stage<T> sq[NS];
        static_for<0, NS>([&](auto is) { nominal_stage(is, sq[decltype(is)::value]); });
        const T h2 = h * T(0.5), h6 = h * T(1.0 / 6.0);
        const T p1 = h2 * sq[1].cF, p2 = h2 * sq[2].cF, p3 = h * sq[3].cF;
        const T A1 = sq[0].cF, A2 = xfma(p1, A1, sq[1].cF), A3 = xfma(p2, A2, sq[2].cF), A4 = xfma(p3, A3, sq[3].cF);
        const T Phi = xfma(h6, xfma(T(2), A2 + A3, A1 + A4), T(1));
        const T w2 = T(2) + p3, w1 = xfma(p2, w2, T(2)), w0 = xfma(p1, w1, T(1));
        const T W[NS] = {h6 * w0, h6 * w1, h6 * w2, h6};
        T WC[NS], WGsum = T(0);
        static_for<0, NS>([&](auto is) {
            constexpr int q = decltype(is)::value;
            WC[q] = W[q] * sq[q].cCn;
            WGsum = xfma(W[q], sq[q].cG, WGsum);
        });
        static_for<0, ND>([&](auto ic) {
            constexpr int d = decltype(ic)::value;
            constexpr int C = D::code(d);
            T acc = Phi * tg.V[d][1];
            if constexpr (dir_has_cn(C)) {
                const T c = tg.V[d][0];
                acc = xfma(WC[0], c, acc);
                static_for<1, NS>([&](auto is) {
                    constexpr int q = decltype(is)::value;
                    T vcn = c * dl[q];
                    if constexpr (dir_forced_cn(C)) vcn = xfma(iv.sE[d], ps[q - 1], vcn);
                    acc = xfma(WC[q], vcn, acc);
                });
                tg.V[d][0] = dir_forced_cn(C) ? xfma(c, dl[NS], iv.sE[d] * ps[3]) : c * dl[NS];
            }
            static_for<0, NS>([&](auto is) {
                constexpr int q = decltype(is)::value;
                if constexpr (C == D_A_REST) acc = xfma(W[q], sq[q].cA, acc);
                if constexpr (C == D_TAU1_REST) acc = xfma(W[q], sq[q].cT1, acc);
                if constexpr (C == D_KM_REST) acc = xfma(W[q], sq[q].cKm, acc);
                if constexpr (C == D_TAU2) acc = xfma(W[q], sq[q].cT2, acc);
                if constexpr (C == D_FC) acc = xfma(W[q], sq[q].fU, acc);
            });
            if constexpr (dir_has_dg(C)) acc = xfma(WGsum, iv.dg[d], acc);
            tg.V[d][1] = acc;
        });
        x[0] = xfma(x[0], dl[NS], iv.E * ps[3]);
        x[1] = xn[1];
        return;
    }
#ifndef FES_NO_PROP2  // measured against the stage-by-stage reduced columns (profiles/r01_tune_prop2_ag.txt): headline +15 %
    constexpr bool PROP2 = FAT && ND > 0 && SCHEME == 4;
#else
    constexpr bool PROP2 = false;
#endif
    if constexpr (PROP2) {
        // Fatigue models, reduced columns: every column is the same 2 x 2 linear system
        //     (vF, W)' = M(t) (vF, W) + g_d(t) e_F,    M = [[cF, cW], [1, -1/tau_fat]],
        // so one RK4 step is  v_end = Phi v + sum_s (Lambda_s e_F) g_d(s),  Lambda_s = beta_s h I + a_{s+1} h Lambda_{s+1} M_{s+1}
        // (the total weight of stage s in the step), Phi = I + sum_s Lambda_s M_s -- functions of cF, cW at the four stages
        // only, written out once for all columns.  p_s = Lambda_s e_F, r_s = Lambda_s e_W; T_s1 = Lambda_s M_s e_F, T_s2 = Lambda_s M_s e_W.
        Stage<T> sq[NS];
        static_for<0, NS>([&](auto is) { nominal_stage(is, sq[decltype(is)::value]); });
        const T h2 = h * T(0.5), h3 = h * T(1.0 / 3.0), h6 = h * T(1.0 / 6.0), kp = tr.kappa;
        const T t41_0 = h6 * sq[3].cF, t41_1 = h6, t42_0 = h6 * sq[3].cW, t42_1 = -(kp * h6);
        const T p3_0 = xfma(h, t41_0, h3), p3_1 = h * t41_1, r3_0 = h * t42_0, r3_1 = xfma(h, t42_1, h3);
        const T t31_0 = xfma(p3_0, sq[2].cF, r3_0), t31_1 = xfma(p3_1, sq[2].cF, r3_1);
        const T t32_0 = xfma(p3_0, sq[2].cW, -(kp * r3_0)), t32_1 = xfma(p3_1, sq[2].cW, -(kp * r3_1));
        const T p2_0 = xfma(h2, t31_0, h3), p2_1 = h2 * t31_1, r2_0 = h2 * t32_0, r2_1 = xfma(h2, t32_1, h3);
        const T t21_0 = xfma(p2_0, sq[1].cF, r2_0), t21_1 = xfma(p2_1, sq[1].cF, r2_1);
        const T t22_0 = xfma(p2_0, sq[1].cW, -(kp * r2_0)), t22_1 = xfma(p2_1, sq[1].cW, -(kp * r2_1));
        const T p1_0 = xfma(h2, t21_0, h6), p1_1 = h2 * t21_1, r1_0 = h2 * t22_0, r1_1 = xfma(h2, t22_1, h6);
        const T t11_0 = xfma(p1_0, sq[0].cF, r1_0), t11_1 = xfma(p1_1, sq[0].cF, r1_1);
        const T t12_0 = xfma(p1_0, sq[0].cW, -(kp * r1_0)), t12_1 = xfma(p1_1, sq[0].cW, -(kp * r1_1));
        const T Phi00 = T(1) + ((t11_0 + t21_0) + (t31_0 + t41_0)), Phi10 = (t11_1 + t21_1) + (t31_1 + t41_1);
        // f32 build: W decays by ~2e-5 per step, a factor a float cannot hold (1 - 2e-5 rounds to 0.1 % of the rate): W is
        // advanced as W + (dPhi11 W + ...) with dPhi11 = Phi11 - 1 formed without the cancellation
        constexpr bool WINC = sizeof(T) == 4;
        const T dPhi11 = (t12_1 + t22_1) + (t32_1 + t42_1);
        const T Phi01 = (t12_0 + t22_0) + (t32_0 + t42_0), Phi11 = WINC ? dPhi11 : T(1) + dPhi11;
        const T pF[NS] = {p1_0, p2_0, p3_0, h6}, pW[NS] = {p1_1, p2_1, p3_1, T(0)};
        // S(X) = sum_s p_s X_s for the forcing profiles the columns of this set use
        auto wsum = [&](auto coef, T &sF, T &sW) {
            const T c0 = coef(std::integral_constant<int, 0>{}), c1 = coef(std::integral_constant<int, 1>{});
            const T c2 = coef(std::integral_constant<int, 2>{}), c3 = coef(std::integral_constant<int, 3>{});
            sF = xfma(pF[3], c3, xfma(pF[2], c2, xfma(pF[1], c1, pF[0] * c0)));
            sW = xfma(pW[2], c2, xfma(pW[1], c1, pW[0] * c0));
        };
        constexpr bool ANY_CN = dirs_have<D>(D_CN0) || dirs_any_forced_cn<D>();
        T sCdF = T(0), sCdW = T(0), sCpF = T(0), sCpW = T(0), sAF = T(0), sAW = T(0), sT1F = T(0), sT1W = T(0);
        T sKmF = T(0), sKmW = T(0), sGF = T(0), sGW = T(0);
        constexpr bool ANY_DG = dirs_have<D>(D_PW) || dirs_have<D>(D_PD0) || dirs_have<D>(D_PDT);
        if constexpr (ANY_CN)
            wsum([&](auto q) { constexpr int s = decltype(q)::value; return s == 0 ? sq[0].cCn : sq[s].cCn * dl[s]; }, sCdF, sCdW);
        if constexpr (dirs_any_forced_cn<D>()) {
            sCpF = xfma(pF[3], sq[3].cCn * ps[2], xfma(pF[2], sq[2].cCn * ps[1], pF[1] * (sq[1].cCn * ps[0])));
            sCpW = xfma(pW[2], sq[2].cCn * ps[1], pW[1] * (sq[1].cCn * ps[0]));
        }
        // Forcing of the fatigue rows (dir_reduced): every such column adds  sum_s p_s c_X(s) m(s)  with c_X the Jacobian entry
        // of its row (cA | cT1 | cKm) and m(s) = m0 fm_s + (what its own forcing adds within the step):
        //     seeds A0 / Tau10 / Km0        m(s) = mu fm_s
        //     a_rest / tau1_rest / km_rest  m(s) = R fm_s + em_s                    (em = 1 - fm)
        //     alpha_a / alpha_tau1 / _km    m(s) = M fm_s + Wn(s)
        // so the sums S_fm(c_X) (weights p_s fm_s), S_em(c_X) (weights p_s em_s) and S_wn(c_X) (weights p_s Wn(s)) are formed
        // once per row and shared by the columns that use them.
        constexpr bool ROW_A = dirs_have<D>(D_A0) || dirs_have<D>(D_A_REST) || dirs_have<D>(D_ALPHA_A);
        constexpr bool ROW_T1 = dirs_have<D>(D_TAU10) || dirs_have<D>(D_TAU1_REST) || dirs_have<D>(D_ALPHA_TAU1);
        constexpr bool ROW_KM = dirs_have<D>(D_KM0) || dirs_have<D>(D_KM_REST) || dirs_have<D>(D_ALPHA_KM);
        constexpr bool OWN_M = dirs_have<D>(D_TAU_FAT);
        const T qF[NS] = {pF[0], pF[1] * fm[1], pF[2] * fm[2], pF[3] * fm[3]}, qW[NS] = {pW[0], pW[1] * fm[1], pW[2] * fm[2], T(0)};
        auto msum = [&](auto coef, T &sF, T &sW) {
            sF = xfma(qF[3], coef(3), xfma(qF[2], coef(2), xfma(qF[1], coef(1), qF[0] * coef(0))));
            sW = xfma(qW[2], coef(2), xfma(qW[1], coef(1), qW[0] * coef(0)));
        };
        if constexpr (ROW_A) msum([&](int s) { return sq[s].cA; }, sAF, sAW);
        if constexpr (ROW_T1) msum([&](int s) { return sq[s].cT1; }, sT1F, sT1W);
        if constexpr (ROW_KM) msum([&](int s) { return sq[s].cKm; }, sKmF, sKmW);
        T eAF = T(0), eAW = T(0), eT1F = T(0), eT1W = T(0), eKmF = T(0), eKmW = T(0);   // S_em
        if constexpr (ANY_REST) {
            const T wF[NS] = {T(0), pF[1] * em[1], pF[2] * em[2], pF[3] * em[3]}, wW[NS] = {T(0), pW[1] * em[1], pW[2] * em[2], T(0)};
            auto esum = [&](auto coef, T &sF, T &sW) {
                sF = xfma(wF[3], coef(3), xfma(wF[2], coef(2), wF[1] * coef(1)));
                sW = xfma(wW[2], coef(2), wW[1] * coef(1));
            };
            if constexpr (dirs_have<D>(D_A_REST)) esum([&](int s) { return sq[s].cA; }, eAF, eAW);
            if constexpr (dirs_have<D>(D_TAU1_REST)) esum([&](int s) { return sq[s].cT1; }, eT1F, eT1W);
            if constexpr (dirs_have<D>(D_KM_REST)) esum([&](int s) { return sq[s].cKm; }, eKmF, eKmW);
        }
        T nAF = T(0), nAW = T(0), nT1F = T(0), nT1W = T(0), nKmF = T(0), nKmW = T(0);   // S_wn
        if constexpr (ANY_ALPHA) {
            const T wF[NS] = {T(0), pF[1] * wnv[1], pF[2] * wnv[2], pF[3] * wnv[3]}, wW[NS] = {T(0), pW[1] * wnv[1], pW[2] * wnv[2], T(0)};
            auto nsum = [&](auto coef, T &sF, T &sW) {
                sF = xfma(wF[3], coef(3), xfma(wF[2], coef(2), wF[1] * coef(1)));
                sW = xfma(wW[2], coef(2), wW[1] * coef(1));
            };
            if constexpr (dirs_have<D>(D_ALPHA_A)) nsum([&](int s) { return sq[s].cA; }, nAF, nAW);
            if constexpr (dirs_have<D>(D_ALPHA_TAU1)) nsum([&](int s) { return sq[s].cT1; }, nT1F, nT1W);
            if constexpr (dirs_have<D>(D_ALPHA_KM)) nsum([&](int s) { return sq[s].cKm; }, nKmF, nKmW);
        }
        T sT2F = T(0), sT2W = T(0), sFcF = T(0), sFcW = T(0);
        if constexpr (dirs_have<D>(D_TAU2)) wsum([&](auto q) { return sq[decltype(q)::value].cT2; }, sT2F, sT2W);
        if constexpr (dirs_have<D>(D_FC)) wsum([&](auto q) { return sq[decltype(q)::value].fU; }, sFcF, sFcW);
        // d / d tau_fat: the three rows are forced by (x - x_rest)/tau_fat^2 = kappa^2 (d_X fm(t) + alpha_X Wn(t)), hence
        //     m_X(s) = m_X fm_s + kappa^2 (d_X sigma_s + alpha_X omega_s),
        // sigma / omega = the scheme's stage values of y' = -y/tau_fat + fm(t) | + Wn(t) from 0 (sigma is loop-invariant).
        T sg[NS + 1] = {}, om[NS + 1] = {};
        if constexpr (OWN_M) {
            const T ks1 = xfma(-kp, h2, fm[1]);                 // sigma_1 = h/2 (k_0 = fm_0 = 1)
            sg[1] = h2;
            sg[2] = h2 * ks1;
            const T ks2 = xfma(-kp, sg[2], fm[2]);
            sg[3] = h * ks2;
            const T ks3 = xfma(-kp, sg[3], fm[3]);
            sg[4] = h6 * ((T(1) + ks3) + T(2) * (ks1 + ks2));
            const T ko1 = wnv[1];                               // omega_1 = 0 (k_0 = Wn(0) = 0)
            om[2] = h2 * ko1;
            const T ko2 = xfma(-kp, om[2], wnv[2]);
            om[3] = h * ko2;
            const T ko3 = xfma(-kp, om[3], wnv[3]);
            om[4] = h6 * (ko3 + T(2) * (ko1 + ko2));
        }
        if constexpr (ANY_DG)
            wsum([&](auto q) { return sq[decltype(q)::value].cG; }, sGF, sGW);
        const T mu0 = (T)tg.mu, nu0 = (T)tg.nu;
        static_for<0, ND>([&](auto ic) {
            constexpr int d = decltype(ic)::value;
            constexpr int C = D::code(d);
            const T vF = tg.V[d][1], vW = tg.V[d][2];
            T aF = xfma(Phi01, vW, Phi00 * vF), aW = xfma(Phi11, vW, Phi10 * vF);
            if constexpr (C == D_CN0) {
                aF = xfma(sCdF, nu0, aF);
                aW = xfma(sCdW, nu0, aW);
            }
            if constexpr (dir_forced_cn(C)) {
                const T c = tg.V[d][0];
                aF = xfma(sCpF, iv.sE[d], xfma(sCdF, c, aF));
                aW = xfma(sCpW, iv.sE[d], xfma(sCdW, c, aW));
                tg.V[d][0] = xfma(c, dl[NS], iv.sE[d] * ps[3]);
            }
            if constexpr (C == D_A0) {
                aF = xfma(sAF, mu0, aF);
                aW = xfma(sAW, mu0, aW);
            }
            if constexpr (C == D_TAU10) {
                aF = xfma(sT1F, mu0, aF);
                aW = xfma(sT1W, mu0, aW);
            }
            if constexpr (C == D_KM0) {
                aF = xfma(sKmF, mu0, aF);
                aW = xfma(sKmW, mu0, aW);
            }
            if constexpr (dir_has_dg(C)) {
                aF = xfma(sGF, iv.dg[d], aF);
                aW = xfma(sGW, iv.dg[d], aW);
            }
            if constexpr (dir_rest(C)) {
                const T R0 = tg.R;
                const T fF = C == D_A_REST ? sAF : (C == D_TAU1_REST ? sT1F : sKmF), fW = C == D_A_REST ? sAW : (C == D_TAU1_REST ? sT1W : sKmW);
                const T gF = C == D_A_REST ? eAF : (C == D_TAU1_REST ? eT1F : eKmF), gW = C == D_A_REST ? eAW : (C == D_TAU1_REST ? eT1W : eKmW);
                aF = xfma(fF, R0, aF + gF);
                aW = xfma(fW, R0, aW + gW);
            }
            if constexpr (dir_alpha(C)) {
                const T M0 = tg.M;
                const T fF = C == D_ALPHA_A ? sAF : (C == D_ALPHA_TAU1 ? sT1F : sKmF), fW = C == D_ALPHA_A ? sAW : (C == D_ALPHA_TAU1 ? sT1W : sKmW);
                const T gF = C == D_ALPHA_A ? nAF : (C == D_ALPHA_TAU1 ? nT1F : nKmF), gW = C == D_ALPHA_A ? nAW : (C == D_ALPHA_TAU1 ? nT1W : nKmW);
                aF = xfma(fF, M0, aF + gF);
                aW = xfma(fW, M0, aW + gW);
            }
            if constexpr (C == D_TAU2) {
                aF += sT2F;
                aW += sT2W;
            }
            if constexpr (C == D_FC) {
                aF += sFcF;
                aW += sFcW;
            }
            if constexpr (dir_own_m(C)) {
                const T mA = tg.V[d][0], mT1 = tg.V[d][3], mKm = tg.V[d][4], k2 = tr.kappa2;
                T gF = T(0), gW = T(0);
                static_for<0, NS>([&](auto is) {
                    constexpr int q = decltype(is)::value;
                    const T Z = xfma(sq[q].cA, mA, xfma(sq[q].cT1, mT1, sq[q].cKm * mKm));
                    T G = q == 0 ? Z : Z * fm[q];
                    if constexpr (q > 0) {
                        const T Y = xfma(sq[q].cA, dA, xfma(sq[q].cT1, dT1, sq[q].cKm * dKm));
                        G = xfma(k2, xfma(Y, sg[q], sq[q].cW * om[q]), G);
                    }
                    gF = xfma(pF[q], G, gF);
                    if constexpr (q < NS - 1) gW = xfma(pW[q], G, gW);
                });
                aF += gF;
                aW += gW;
                tg.V[d][0] = xfma(k2, xfma(dA, sg[NS], tr.alpha_a * om[NS]), xfma(-em[NS], mA, mA));
                tg.V[d][3] = xfma(k2, xfma(dT1, sg[NS], tr.alpha_tau1 * om[NS]), xfma(-em[NS], mT1, mT1));
                tg.V[d][4] = xfma(k2, xfma(dKm, sg[NS], tr.alpha_km * om[NS]), xfma(-em[NS], mKm, mKm));
            }
            tg.V[d][1] = aF;
            tg.V[d][2] = WINC ? vW + aW : aW;
        });
        if constexpr (ANY_ALPHA) tg.M = xfma(-em[NS], tg.M, tg.M) + wn_n;
        if constexpr (ANY_REST) tg.R = xfma(em[NS], T(1) - tg.R, tg.R);
        x[0] = xfma(x[0], dl[NS], iv.E * ps[3]);
        x[1] = xn[1];
        finish_fatigue();
        if constexpr (MODES && sizeof(T) == 8) {
            tg.mu *= (double)fm[NS];
            tg.nu *= (double)dl[NS];
        } else if constexpr (MODES) {  // f32 build: a float cannot hold 1 - h/tau_fat, the step factors are formed in double
            auto step_factor = [](double z) {
                const double f1 = fma(-0.5 * z, 1.0, 1.0), f2 = fma(-0.5 * z, f1, 1.0), f3 = fma(-z, f2, 1.0);
                return fma(-z * (1.0 / 6.0), (1.0 + f3) + 2.0 * (f1 + f2), 1.0);
            };
            tg.mu *= step_factor(zf);
            tg.nu *= step_factor(zc);
        }
        return;
    }
    static_for<0, NS>([&](auto is) {
        Stage<T> st;
        nominal_stage(is, st);
        column_stage(is, st);
    });
    x[0] = xfma(x[0], dl[NS], iv.E * ps[3]);
#pragma unroll
    for (int j = 1; j < NXN; ++j) x[j] = xn[j];
    if constexpr (FAT) {
        finish_fatigue();
    }
    static_for<0, ND>([&](auto ic) {
        constexpr int d = decltype(ic)::value;
        constexpr int C = D::code(d);
        if constexpr (CNF && dir_forced_cn(C)) tg.V[d][0] = xfma(tg.V[d][0], dl[NS], iv.sE[d] * ps[3]);
        static_for<0, NX>([&](auto ij) {
            constexpr int j = decltype(ij)::value;
            if constexpr (dir_stores(FAT, C, j, CNF)) tg.V[d][j] = vn[d][j];
        });
    });
    if constexpr (FAT && ANY_ALPHA) tg.M = xfma(-em[NS], tg.M, tg.M) + wn_n;
    if constexpr (FAT && ANY_REST) tg.R = xfma(em[NS], T(1) - tg.R, tg.R);
    if constexpr (MODES && sizeof(T) == 8) {
        tg.mu *= (double)fm[NS];
        tg.nu *= (double)dl[NS];
    } else if constexpr (MODES) {
        tg.mu = mun;
        tg.nu = nun;
    }
}

// U consecutive RK4 steps (i0 .. i0 + U - 1 of the interval) as ONE basic block: no scheme branch and no loop back edge between
// them, so that the instruction scheduler can start the nominal stage chain of step i + 1 (a serial chain: reciprocal -> q -> F'
// -> next stage state) while the propagator / column phase of step i is still in flight.  tbl = the stage table of the call.
template <typename T, int KIND, bool FAT, int GROUPS, int U>
FES_HD void rk4_steps(const Traj<T> &tr, const Interval<T, Dirs<KIND, FAT, GROUPS>::ND> &iv, T h, const T *tbl, int S, int i0,
                      T *x, Tangent<T, KIND, FAT, GROUPS> &tg) {
    static_for<0, U>([&](auto iu) {
        const int i = i0 + decltype(iu)::value;
        const T e[3] = {tbl[2 * i], tbl[2 * i + 1], tbl[2 * i + 2]};
        rk_step_scheme<T, KIND, FAT, GROUPS, 4>(tr, iv, h, e, tbl + 2 * S + 1 + 4 * i, x, tg);
    });
}

// e0, e1, e2 = eps at the step start, midpoint and end.  ONE uniform branch per step selects the scheme.
template <typename T, int KIND, bool FAT, int GROUPS>
FES_HD void rk_step(const Traj<T> &tr, const Interval<T, Dirs<KIND, FAT, GROUPS>::ND> &iv, int scheme, T h,
                    T e0, T e1, T e2, const T *ps, T *x, Tangent<T, KIND, FAT, GROUPS> &tg) {
    const T e[3] = {e0, e1, e2};
    if (scheme == 4) rk_step_scheme<T, KIND, FAT, GROUPS, 4>(tr, iv, h, e, ps, x, tg);
    else if (scheme == 2) rk_step_scheme<T, KIND, FAT, GROUPS, 2>(tr, iv, h, e, ps, x, tg);
    else rk_step_scheme<T, KIND, FAT, GROUPS, 1>(tr, iv, h, e, ps, x, tg);
}

}  // namespace fes
