// fes_hess.cuh -- second-order tangents of the interval map: d2 Phi_k / d z_k d z_k, z_k = (x_k, u_k [, c_k]).
//
// What it serves: the reference's Ipopt runs on the EXACT Hessian of the Lagrangian by default
// (external/bioptim/bioptim/interfaces/ipopt_options.py:88), which casadi obtains by differentiating the unrolled RK
// graph of every continuity constraint g_k = x_{k+1} - Phi_k(x_k, u_k) (external/bioptim/bioptim/limits/penalty.py:1056-1105)
// twice.  The interval kernel (fes_kernels.cuh) gives the first derivatives; this kernel gives the second ones, exactly:
// the second-order tangent of the DISCRETE map, not a difference quotient.
//
// How (B200-first): interval k depends on (x_k, u_k) only, so the second derivative is block diagonal and every
// (scenario, interval, pair of directions i <= j) is an independent unit of work: ONE THREAD PER UNIT, batch index
// innermost (coalesced loads of X / ctrl, coalesced stores).  A thread integrates the interval once in hyper-dual
// arithmetic -- numbers (v, a, b, c) with a = d/dz_i, b = d/dz_j, c = d2/dz_i dz_j, closed under + - * 1/x exp tanh -- so
// the `c` parts of the end state ARE the wanted entries (forward-over-forward differentiation of the same RK stages as
// external/bioptim/bioptim/dynamics/integrator.py:300-365).  No workspace and no tables: the window sum
// S(t) = E_k eps(t - t_k) is assembled per thread (2 N exp per interval), eps at the stage offsets is a table of 2 S + 1
// doubles that every block builds in shared memory (the same numbers as fes_core.cuh's stage table).
// The unit is small (20 doubles of state per stage vector) and the grid is nz (nz + 1) / 2 times the interval kernel's.
// Cost: each pair recomputes value and first-order parts, so the call costs ~3.7 x the difference
// quotients of the exact Jacobian (12 interval-kernel calls); what it buys is rounding-level accuracy.  fp64 only.
#pragma once

#include "fes_core.cuh"

namespace fes {

struct HD {
    double v, a, b, c;
};
FES_HD HD hd(double v) { return HD{v, 0.0, 0.0, 0.0}; }
FES_HD HD operator+(const HD &x, const HD &y) { return HD{x.v + y.v, x.a + y.a, x.b + y.b, x.c + y.c}; }
FES_HD HD operator-(const HD &x, const HD &y) { return HD{x.v - y.v, x.a - y.a, x.b - y.b, x.c - y.c}; }
FES_HD HD operator+(const HD &x, double s) { return HD{x.v + s, x.a, x.b, x.c}; }
FES_HD HD operator+(double s, const HD &x) { return x + s; }
FES_HD HD operator-(const HD &x, double s) { return HD{x.v - s, x.a, x.b, x.c}; }
FES_HD HD operator-(double s, const HD &x) { return HD{s - x.v, -x.a, -x.b, -x.c}; }
FES_HD HD operator*(const HD &x, double s) { return HD{x.v * s, x.a * s, x.b * s, x.c * s}; }
FES_HD HD operator*(double s, const HD &x) { return x * s; }
FES_HD HD operator*(const HD &x, const HD &y) {
    return HD{x.v * y.v, fma(x.a, y.v, x.v * y.a), fma(x.b, y.v, x.v * y.b),
              fma(x.c, y.v, fma(x.a, y.b, fma(x.b, y.a, x.v * y.c)))};
}
// y + s x  (the stage states and the RK combination)
FES_HD HD axpy(double s, const HD &x, const HD &y) { return HD{fma(s, x.v, y.v), fma(s, x.a, y.a), fma(s, x.b, y.b), fma(s, x.c, y.c)}; }
FES_HD HD axpy(double s, const HD &x, double y) { return HD{fma(s, x.v, y), s * x.a, s * x.b, s * x.c}; }
// 1 / x: f' = -r^2, f'' = 2 r^3
FES_HD HD hrcp(const HD &x) {
    const double r = xrcp(x.v), f1 = -r * r;
    const double a = f1 * x.a;
    return HD{r, a, f1 * x.b, fma(f1, x.c, -2.0 * r * a * x.b)};
}
FES_HD double hrcp(double x) { return xrcp(x); }
FES_HD HD hexp(const HD &x) {
    const double e = exp(x.v);
    return HD{e, e * x.a, e * x.b, e * fma(x.a, x.b, x.c)};
}
FES_HD HD htanh(const HD &x) {
    const double th = xtanh(x.v), d1 = fma(-th, th, 1.0), d2 = -2.0 * th * d1;
    return HD{th, d1 * x.a, d1 * x.b, fma(d1, x.c, d2 * x.a * x.b)};
}

struct HessArgs {
    const double *params;  // [NP][B] | [NP]
    int64_t pB, pS;
    const double *X;       // [ns+1][nx][B]
    const double *win;     // [>=ns][N][B] | [>=ns][N]
    int64_t wB, wS;
    const double *ctrl;    // [ns][nur][B] | [ns][nur]
    int64_t cB, cS;
    int nur;               // rows of ctrl per node: the model's controls (+ 1 with the force factor)
    const double *lam;     // [ns][nx][B] multipliers | NULL
    double *D2;            // [ns][nx][npairs][B] | NULL
    double *H;             // [ns][npairs][B] = sum_r lam_r D2_r | NULL
    int64_t B;
    int N, ns, S, scheme, factor, nz, npairs;
    double h, dt, t_start, tauc, r0_km;
};

// pair index p of the upper triangle, row-major: (0,0), (0,1), ..., (0,nz-1), (1,1), ...
FES_HD void hess_pair_of(int p, int nz, int &i, int &j) {
    i = 0;
    while (p >= nz - i) {
        p -= nz - i;
        ++i;
    }
    j = i + p;
}

// One (interval k, pair (pi, pj), scenario b): writes d2 Phi_k[r] / dz_pi dz_pj for r = 0..NX-1 into out[r].
// z = (Cn, F [, A, Tau1, Km], controls of the model [, force factor]).
template <int KIND, bool FAT, bool FC>
FES_HD void hess_unit(const HessArgs &a, const double *eps, int k, int pi, int pj, int64_t b, double *out) {
    constexpr int NX = FAT ? 5 : 2;
    auto par = [&](int j) { return a.params[j * a.pB + b * a.pS]; };
    auto seed = [&](double v, int idx) { return HD{v, idx == pi ? 1.0 : 0.0, idx == pj ? 1.0 : 0.0, 0.0}; };
    const double inv_tauc = 1.0 / a.tauc;
    const double tau2 = par(P_TAU2);
    const double r0m1 = par(P_KM_REST) + a.r0_km - 1.0;
    const int nu = KIND == DING2007 ? 1 : (KIND == HMED2018 ? a.N : 0);
    auto ctl = [&](int c) { return a.ctrl[((int64_t)k * a.nur + c) * a.cB + b * a.cS]; };

    HD x[NX];
#pragma unroll
    for (int r = 0; r < NX; ++r) x[r] = seed(a.X[((int64_t)k * NX + r) * a.B + b], r);

    // window of the node: E = sum_i lambda_i wa_i R_i  (ding2003.py:230-252; hmed2018.py:169-180 for lambda)
    using ET = std::conditional_t<KIND == HMED2018, HD, double>;
    ET E = ET();
    {
        const double tn = a.t_start + a.dt * (double)k;
        double prev = 0.0;
        for (int i = 0; i < a.N; ++i) {
            const double ti = a.win[((int64_t)k * a.N + i) * a.wB + b * a.wS];
            const double wa = exp(-(tn - ti) * inv_tauc);
            const double R = i == 0 ? 1.0 : fma(r0m1, exp(-(ti - prev) * inv_tauc), 1.0);
            prev = ti;
            if constexpr (KIND == HMED2018) {
                const HD I = seed(ctl(i), NX + i);
                const HD lam = (htanh((I - par(P_IS)) * par(P_BS)) + par(P_CR)) * par(P_AR);
                E = E + lam * (wa * R);
            } else {
                E += wa * R;
            }
        }
    }
    // pulse-width scaling of A (ding2007.py:172-188)
    using GT = std::conditional_t<KIND == DING2007, HD, double>;
    GT g;
    if constexpr (KIND == DING2007) {
        const HD pw = seed(ctl(0), NX);
        g = 1.0 - hexp((par(P_PD0) - pw) * (1.0 / par(P_PDT)));
    } else {
        g = 1.0;
    }
    HD fc = hd(1.0);
    if constexpr (FC) fc = seed(ctl(nu), NX + nu);

    const double a_rest = par(P_A_REST), tau1_rest = par(P_TAU1_REST), km_rest = par(P_KM_REST);
    const double kappa = FAT ? 1.0 / par(P_TAU_FAT) : 0.0;
    const double al_a = par(P_ALPHA_A), al_t = par(P_ALPHA_TAU1), al_k = par(P_ALPHA_KM);

    // linear rows as fused multiply-adds on the four parts: y' = rest kappa - kappa y + alpha F, Cn' = (E eps - Cn) / tauc
    auto lin = [&](double c0, double m, const HD &y, double al, const HD &F) {
        return HD{fma(m, y.v, fma(al, F.v, c0)), fma(m, y.a, al * F.a), fma(m, y.b, al * F.b), fma(m, y.c, al * F.c)};
    };
    const double a_rest_k = a_rest * kappa, tau1_rest_k = tau1_rest * kappa, km_rest_k = km_rest * kappa;
    auto rhs = [&](const HD *y, double eps, HD *dy) {
        const HD &Cn = y[0], &F = y[1];
        dy[0] = axpy(-inv_tauc, Cn, E * (eps * inv_tauc));
        if constexpr (FAT) {
            const HD &A = y[2], &T1 = y[3], &Km = y[4];
            const HD q = Cn * hrcp(Km + Cn);
            HD f;
            if constexpr (KIND == DING2007) f = (A * g) * q - F * hrcp(axpy(tau2, q, T1));
            else f = A * q - F * hrcp(axpy(tau2, q, T1));
            if constexpr (FC) f = f * fc;
            dy[1] = f;
            dy[2] = lin(a_rest_k, -kappa, A, al_a, F);
            dy[3] = lin(tau1_rest_k, -kappa, T1, al_t, F);
            dy[4] = lin(km_rest_k, -kappa, Km, al_k, F);
        } else {
            const HD q = Cn * hrcp(Cn + km_rest);
            HD f = (a_rest * g) * q - F * hrcp(q * tau2 + tau1_rest);
            if constexpr (FC) f = f * fc;
            dy[1] = f;
        }
    };

    // One loop per scheme, so that a step is straight-line code.  RK4: x, the running combination, the stage state and ONE
    // stage derivative are live (not k1..k4): 80 doubles.  eps[j] = exp(-(j h/2)/tauc), j = 0..2S, is fes_core.cuh's stage
    // table (hess_eps_table): the same for every unit of a call, built once per block in shared memory by the kernel.
    const double h = a.h;
    auto eps_at = [&](int idx) { return eps[idx]; };
    if (a.scheme == 4) {
        double e0 = 1.0;
        for (int s = 0; s < a.S; ++s) {
            const double e1 = eps_at(2 * s + 1), e2 = eps_at(2 * s + 2);
            HD kk[NX], y[NX], acc[NX];
            rhs(x, e0, kk);
#pragma unroll
            for (int r = 0; r < NX; ++r) {
                acc[r] = axpy(h / 6.0, kk[r], x[r]);
                y[r] = axpy(0.5 * h, kk[r], x[r]);
            }
            rhs(y, e1, kk);
#pragma unroll
            for (int r = 0; r < NX; ++r) {
                acc[r] = axpy(h / 3.0, kk[r], acc[r]);
                y[r] = axpy(0.5 * h, kk[r], x[r]);
            }
            rhs(y, e1, kk);
#pragma unroll
            for (int r = 0; r < NX; ++r) {
                acc[r] = axpy(h / 3.0, kk[r], acc[r]);
                y[r] = axpy(h, kk[r], x[r]);
            }
            rhs(y, e2, kk);
#pragma unroll
            for (int r = 0; r < NX; ++r) x[r] = axpy(h / 6.0, kk[r], acc[r]);
            e0 = e2;
        }
    } else if (a.scheme == 2) {  // midpoint (integrator.py:324-349)
        for (int s = 0; s < a.S; ++s) {
            const double e0 = eps_at(2 * s), e1 = eps_at(2 * s + 1);
            HD kk[NX], y[NX];
            rhs(x, e0, kk);
#pragma unroll
            for (int r = 0; r < NX; ++r) y[r] = axpy(0.5 * h, kk[r], x[r]);
            rhs(y, e1, kk);
#pragma unroll
            for (int r = 0; r < NX; ++r) x[r] = axpy(h, kk[r], x[r]);
        }
    } else {
        for (int s = 0; s < a.S; ++s) {
            HD kk[NX];
            rhs(x, eps_at(2 * s), kk);
#pragma unroll
            for (int r = 0; r < NX; ++r) x[r] = axpy(h, kk[r], x[r]);
        }
    }
#pragma unroll
    for (int r = 0; r < NX; ++r) out[r] = x[r].c;
}

FES_HD double hess_eps_entry(const HessArgs &a, int idx) { return exp(-(0.5 * a.h * (double)idx) / a.tauc); }

template <int KIND, bool FAT, bool FC>
FES_HD void hess_store(const HessArgs &a, const double *eps, int k, int p, int64_t b) {
    constexpr int NX = FAT ? 5 : 2;
    int pi, pj;
    hess_pair_of(p, a.nz, pi, pj);
    double o[NX];
    hess_unit<KIND, FAT, FC>(a, eps, k, pi, pj, b, o);
    if (a.D2) {
#pragma unroll
        for (int r = 0; r < NX; ++r) a.D2[(((int64_t)k * NX + r) * a.npairs + p) * a.B + b] = o[r];
    }
    if (a.H) {
        double acc = 0.0;
#pragma unroll
        for (int r = 0; r < NX; ++r) acc = fma(a.lam[((int64_t)k * NX + r) * a.B + b], o[r], acc);
        a.H[((int64_t)k * a.npairs + p) * a.B + b] = acc;
    }
}

#if defined(__CUDACC__)
#ifndef FES_HESS_MIN_BLOCKS
#define FES_HESS_MIN_BLOCKS 2   // resident blocks of 128 threads the register budget is sized for (fatigue instances: ~216 registers)
#endif
template <int KIND, bool FAT, bool FC>
__global__ void __launch_bounds__(128, FES_HESS_MIN_BLOCKS) fes_hess_kernel(const HessArgs a) {
    extern __shared__ double s_eps[];   // 2 S + 1 doubles (dynamic)
    for (int j = threadIdx.x; j <= 2 * a.S; j += blockDim.x) s_eps[j] = hess_eps_entry(a, j);
    __syncthreads();
    const int64_t total = (int64_t)a.ns * a.npairs * a.B;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
        const int64_t b = idx % a.B, rest = idx / a.B;
        const int p = (int)(rest % a.npairs), k = (int)(rest / a.npairs);
        hess_store<KIND, FAT, FC>(a, s_eps, k, p, b);
    }
}

int launch_hess(int kind, int fat, const HessArgs &a, cudaStream_t s);
#endif

}  // namespace fes
