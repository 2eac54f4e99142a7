/*
 * CPU oracle for the FES muscle-ODE hot path, C restatement  --  TEST INFRASTRUCTURE ONLY.
 *
 * Same algorithm as oracle/fes_oracle.py (the numpy restatement that is pinned against the
 * reference's golden vectors), in scalar C so that (a) parity can be checked at sizes numpy cannot
 * reach in seconds and (b) bench.py has a CPU baseline ("kind": "port") to time on the GPU box's host
 * cores (OpenMP over trajectories).  Nothing under cocofest_b200/ links or loads this file.
 *
 * It follows the REFERENCE FORMULATION on purpose (what casadi's SX VM executes per RHS call):
 *   - the N-term sum  S(t) = sum_i R_i * exp(-(t - t_i)/tauc) * lambda_i  is re-evaluated at every
 *     stage with R_i recomputed                      (cocofest/models/ding2003.py:230-252, :213, :228)
 *   - Cn' = S/tauc - Cn/tauc                          (ding2003.py:266)
 *   - F'  = A q - F/(tau1 + tau2 q), q = Cn/(Km+Cn), two true divisions     (ding2003.py:309)
 *   - A_eff = A (1 - exp(-(pw - pd0)/pdt))            (ding2007.py:188)
 *   - lambda_i = ar (tanh(bs (I_i - Is)) + cr)        (hmed2018.py:179)
 *   - fatigue rows                                    (ding2003_with_fatigue.py:197-240,
 *                                                      ding2007_with_fatigue.py:198-241 with a_scale as rest value)
 *   - RK1 / RK2 / RK4 stages, t = t0 + h (i-1), controls and window constant over the interval
 *                                                     (external/bioptim/bioptim/dynamics/integrator.py:300-365)
 *   - single shooting over nodes, t0 = dt*node        (interfaces/solve_ivp_interface.py:56-69,
 *                                                      optimization/optimization_vector.py:373)
 *   - continuity defect g_k = x_{k+1} - Phi_k(x_k)    (limits/penalty.py:1092-1099)
 *
 * Sensitivities are obtained by forward-mode dual numbers pushed through exactly the same expression
 * tree (this is what casadi's forward AD of the unrolled integrator computes).  The reference pins no
 * derivative ("parity unpinned" for that half); tests cross-check this file against complex-step
 * differentiation of the numpy oracle.
 *
 * Parameter vector order (include/fes_b200.h): a_rest, km_rest, tau1_rest, tau2, alpha_a, alpha_tau1,
 * alpha_km, tau_fat, pd0, pdt, ar, bs, Is, cr.
 */
#include <math.h>
#include <stddef.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

enum { P_A_REST, P_KM_REST, P_TAU1_REST, P_TAU2, P_ALPHA_A, P_ALPHA_TAU1, P_ALPHA_KM, P_TAU_FAT,
       P_PD0, P_PDT, P_AR, P_BS, P_IS, P_CR, NP_FULL };
enum { DING2003 = 0, DING2007 = 1, HMED2018 = 2 };
#define MAXN 320  /* window length limit of the oracle (the truncation sweep reaches 100) */
#define MAXD 40   /* max tangent directions of the oracle */

/* ------------------------------------------------------------------ plain double path ---------- */
static void rhs_plain(int kind, int fatigue, int N, double t, const double *x, const double *u, const double *p,
                      const double *win, double tauc, double r0_km, double *xdot)
{
    double r0 = p[P_KM_REST] + r0_km;
    double s = 0.0;
    for (int i = 0; i < N; ++i) {
        double ri = (i == 0) ? 1.0 : 1.0 + (r0 - 1.0) * exp(-(win[i] - win[i - 1]) / tauc);
        double e = exp(-(t - win[i]) / tauc);
        double lam = (kind == HMED2018) ? p[P_AR] * (tanh(p[P_BS] * (u[i] - p[P_IS])) + p[P_CR]) : 1.0;
        s += ri * e * lam;
    }
    double cn = x[0], f = x[1];
    xdot[0] = (1.0 / tauc) * s - (cn / tauc);
    double a = fatigue ? x[2] : p[P_A_REST];
    double tau1 = fatigue ? x[3] : p[P_TAU1_REST];
    double km = fatigue ? x[4] : p[P_KM_REST];
    if (kind == DING2007) a = a * (1.0 - exp(-(u[0] - p[P_PD0]) / p[P_PDT]));
    xdot[1] = a * (cn / (km + cn)) - (f / (tau1 + p[P_TAU2] * (cn / (km + cn))));
    if (fatigue) {
        xdot[2] = -(x[2] - p[P_A_REST]) / p[P_TAU_FAT] + p[P_ALPHA_A] * f;
        xdot[3] = -(x[3] - p[P_TAU1_REST]) / p[P_TAU_FAT] + p[P_ALPHA_TAU1] * f;
        xdot[4] = -(x[4] - p[P_KM_REST]) / p[P_TAU_FAT] + p[P_ALPHA_KM] * f;
    }
}

static void interval_plain(int kind, int fatigue, int scheme, int N, int S, double t0, double dt, double *x,
                           const double *u, const double *p, const double *win, double tauc, double r0_km,
                           double *xall /* may be NULL; [S][nx] sub-step states 1..S */)
{
    int nx = fatigue ? 5 : 2;
    double h = dt / S, k1[5], k2[5], k3[5], k4[5], xs[5];
    for (int i = 1; i <= S; ++i) {
        double t = t0 + h * (i - 1);
        rhs_plain(kind, fatigue, N, t, x, u, p, win, tauc, r0_km, k1);
        if (scheme == 1) {
            for (int j = 0; j < nx; ++j) x[j] = x[j] + h * k1[j];
        } else if (scheme == 2) {
            for (int j = 0; j < nx; ++j) xs[j] = x[j] + h / 2 * k1[j];
            rhs_plain(kind, fatigue, N, t + h / 2, xs, u, p, win, tauc, r0_km, k2);
            for (int j = 0; j < nx; ++j) x[j] = x[j] + h * k2[j];
        } else {
            for (int j = 0; j < nx; ++j) xs[j] = x[j] + h / 2 * k1[j];
            rhs_plain(kind, fatigue, N, t + h / 2, xs, u, p, win, tauc, r0_km, k2);
            for (int j = 0; j < nx; ++j) xs[j] = x[j] + h / 2 * k2[j];
            rhs_plain(kind, fatigue, N, t + h / 2, xs, u, p, win, tauc, r0_km, k3);
            for (int j = 0; j < nx; ++j) xs[j] = x[j] + h * k3[j];
            rhs_plain(kind, fatigue, N, t + h, xs, u, p, win, tauc, r0_km, k4);
            for (int j = 0; j < nx; ++j) x[j] = x[j] + h / 6 * (k1[j] + 2 * k2[j] + 2 * k3[j] + k4[j]);
        }
        if (xall) memcpy(xall + (size_t)(i - 1) * nx, x, sizeof(double) * nx);
    }
}

/* ------------------------------------------------------------------ dual-number path ----------- */
typedef struct { double v; double d[MAXD]; } dual;
static int g_dummy_;
#define ND_LOOP for (int k_ = 0; k_ < nd; ++k_)
static inline dual d_const(double v, int nd) { dual r; r.v = v; ND_LOOP r.d[k_] = 0.0; return r; }
static inline dual d_add(dual a, dual b, int nd) { dual r; r.v = a.v + b.v; ND_LOOP r.d[k_] = a.d[k_] + b.d[k_]; return r; }
static inline dual d_sub(dual a, dual b, int nd) { dual r; r.v = a.v - b.v; ND_LOOP r.d[k_] = a.d[k_] - b.d[k_]; return r; }
static inline dual d_mul(dual a, dual b, int nd) { dual r; r.v = a.v * b.v; ND_LOOP r.d[k_] = a.d[k_] * b.v + a.v * b.d[k_]; return r; }
static inline dual d_div(dual a, dual b, int nd) { dual r; r.v = a.v / b.v; ND_LOOP r.d[k_] = (a.d[k_] - r.v * b.d[k_]) / b.v; return r; }
static inline dual d_neg(dual a, int nd) { dual r; r.v = -a.v; ND_LOOP r.d[k_] = -a.d[k_]; return r; }
static inline dual d_exp(dual a, int nd) { dual r; r.v = exp(a.v); ND_LOOP r.d[k_] = r.v * a.d[k_]; return r; }
static inline dual d_tanh(dual a, int nd) { dual r; r.v = tanh(a.v); double s = 1.0 - r.v * r.v; ND_LOOP r.d[k_] = s * a.d[k_]; return r; }
static inline dual d_scale(dual a, double c, int nd) { dual r; r.v = a.v * c; ND_LOOP r.d[k_] = a.d[k_] * c; return r; }
static inline dual d_addc(dual a, double c, int nd) { dual r = a; (void)nd; r.v += c; return r; }

/* fc: the force factor fl*fv + fp of f_dot_fun (ding2003.py:309-311), NULL = the default arguments (1 * 1 + 0) */
static void rhs_dual(int kind, int fatigue, int N, dual t, const dual *x, const dual *u, const dual *p,
                     const dual *win, double tauc, double r0_km, dual *xdot, int nd, const dual *fc)
{
    dual r0 = d_addc(p[P_KM_REST], r0_km, nd);
    dual s = d_const(0.0, nd);
    for (int i = 0; i < N; ++i) {
        dual ri;
        if (i == 0) ri = d_const(1.0, nd);
        else ri = d_addc(d_mul(d_addc(r0, -1.0, nd), d_exp(d_scale(d_neg(d_sub(win[i], win[i - 1], nd), nd), 1.0 / tauc, nd), nd), nd), 1.0, nd);
        dual e = d_exp(d_scale(d_neg(d_sub(t, win[i], nd), nd), 1.0 / tauc, nd), nd);
        dual term = d_mul(ri, e, nd);
        if (kind == HMED2018) {
            dual lam = d_mul(p[P_AR], d_add(d_tanh(d_mul(p[P_BS], d_sub(u[i], p[P_IS], nd), nd), nd), p[P_CR], nd), nd);
            term = d_mul(term, lam, nd);
        }
        s = d_add(s, term, nd);
    }
    dual cn = x[0], f = x[1];
    xdot[0] = d_sub(d_scale(s, 1.0 / tauc, nd), d_scale(cn, 1.0 / tauc, nd), nd);
    dual a = fatigue ? x[2] : p[P_A_REST];
    dual tau1 = fatigue ? x[3] : p[P_TAU1_REST];
    dual km = fatigue ? x[4] : p[P_KM_REST];
    if (kind == DING2007) {
        dual ew = d_exp(d_div(d_neg(d_sub(u[0], p[P_PD0], nd), nd), p[P_PDT], nd), nd);
        a = d_mul(a, d_addc(d_neg(ew, nd), 1.0, nd), nd);
    }
    dual q = d_div(cn, d_add(km, cn, nd), nd);
    xdot[1] = d_sub(d_mul(a, q, nd), d_div(f, d_add(tau1, d_mul(p[P_TAU2], q, nd), nd), nd), nd);
    if (fc) xdot[1] = d_mul(xdot[1], *fc, nd);
    if (fatigue) {
        xdot[2] = d_add(d_div(d_neg(d_sub(x[2], p[P_A_REST], nd), nd), p[P_TAU_FAT], nd), d_mul(p[P_ALPHA_A], f, nd), nd);
        xdot[3] = d_add(d_div(d_neg(d_sub(x[3], p[P_TAU1_REST], nd), nd), p[P_TAU_FAT], nd), d_mul(p[P_ALPHA_TAU1], f, nd), nd);
        xdot[4] = d_add(d_div(d_neg(d_sub(x[4], p[P_KM_REST], nd), nd), p[P_TAU_FAT], nd), d_mul(p[P_ALPHA_KM], f, nd), nd);
    }
}

static void interval_dual(int kind, int fatigue, int scheme, int N, int S, double t0, double dt, dual *x,
                          const dual *u, const dual *p, const dual *win, double tauc, double r0_km, int nd, const dual *fc)
{
    int nx = fatigue ? 5 : 2;
    double h = dt / S;
    dual k1[5], k2[5], k3[5], k4[5], xs[5];
    for (int i = 1; i <= S; ++i) {
        double t = t0 + h * (i - 1);
        rhs_dual(kind, fatigue, N, d_const(t, nd), x, u, p, win, tauc, r0_km, k1, nd, fc);
        if (scheme == 1) {
            for (int j = 0; j < nx; ++j) x[j] = d_add(x[j], d_scale(k1[j], h, nd), nd);
        } else if (scheme == 2) {
            for (int j = 0; j < nx; ++j) xs[j] = d_add(x[j], d_scale(k1[j], h / 2, nd), nd);
            rhs_dual(kind, fatigue, N, d_const(t + h / 2, nd), xs, u, p, win, tauc, r0_km, k2, nd, fc);
            for (int j = 0; j < nx; ++j) x[j] = d_add(x[j], d_scale(k2[j], h, nd), nd);
        } else {
            for (int j = 0; j < nx; ++j) xs[j] = d_add(x[j], d_scale(k1[j], h / 2, nd), nd);
            rhs_dual(kind, fatigue, N, d_const(t + h / 2, nd), xs, u, p, win, tauc, r0_km, k2, nd, fc);
            for (int j = 0; j < nx; ++j) xs[j] = d_add(x[j], d_scale(k2[j], h / 2, nd), nd);
            rhs_dual(kind, fatigue, N, d_const(t + h / 2, nd), xs, u, p, win, tauc, r0_km, k3, nd, fc);
            for (int j = 0; j < nx; ++j) xs[j] = d_add(x[j], d_scale(k3[j], h, nd), nd);
            rhs_dual(kind, fatigue, N, d_const(t + h, nd), xs, u, p, win, tauc, r0_km, k4, nd, fc);
            for (int j = 0; j < nx; ++j) {
                dual sum = d_add(d_add(k1[j], d_scale(k2[j], 2.0, nd), nd), d_add(d_scale(k3[j], 2.0, nd), k4[j], nd), nd);
                x[j] = d_add(x[j], d_scale(sum, h / 6, nd), nd);
            }
        }
    }
}

/* ------------------------------------------------------------------ exported entry points ------ */
/*
 * Layouts (CPU-friendly, row-major, batch OUTERMOST -- unlike the device library's SoA):
 *   params [B or 1][NP_FULL]      x0 [B][nx]       win [B or 1][ns+1][N]      u [B or 1][ns][nu]
 *   x_final [B][nx]
 *   trace   [trace_B][npts][nx]   npts = ns*S+1 (trace_every = 1) or ns+1 (trace_every = S); may be NULL
 * Tangent directions (IVP: chained through all nodes):
 *   seed_x0 != 0 -> nx directions d/dx0 first, then n_pdir directions d/dparams[pdir[j]]
 *   sens_final [B][nx][nd];  sens_trace [trace_B][ns+1][nx][nd] at nodes (may be NULL)
 */
int fes_oracle_ivp(int kind, int fatigue, int scheme, int N, int ns, int S, double dt, double tauc, double r0_km,
                   long long B, const double *params, int params_shared, const double *x0,
                   const double *win, int win_shared, const double *u, int u_shared, int nu,
                   int seed_x0, int n_pdir, const int *pdir,
                   double *x_final, double *trace, long long trace_B, int trace_every,
                   double *sens_final, double *sens_trace, int n_threads)
{
    int nx = fatigue ? 5 : 2;
    int nd = (seed_x0 ? nx : 0) + n_pdir;
    if (N > MAXN || nd > MAXD || (kind == HMED2018 && nu != N) || (kind == DING2007 && nu != 1)) return -1;
    if (trace_every != 1 && trace_every != S) return -1;
    if (trace && trace_every == 1 && S != 1 && (S > 64 || nd > 0)) return -1; /* sub-step traces: forward only */
    int npts = (trace_every == 1) ? ns * S + 1 : ns + 1;
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
#endif
    (void)g_dummy_;
#pragma omp parallel for schedule(static)
    for (long long b = 0; b < B; ++b) {
        const double *p = params + (params_shared ? 0 : (size_t)b * NP_FULL);
        const double *w = win + (win_shared ? 0 : (size_t)b * (ns + 1) * N);
        const double *uu = u ? u + (u_shared ? 0 : (size_t)b * ns * nu) : NULL;
        if (nd == 0) {
            double x[5];
            for (int j = 0; j < nx; ++j) x[j] = x0[(size_t)b * nx + j];
            double *tr = (trace && b < trace_B) ? trace + (size_t)b * npts * nx : NULL;
            if (tr) memcpy(tr, x, sizeof(double) * nx);
            double xall[64 * 5];
            for (int n = 0; n < ns; ++n) {
                int want_all = tr && trace_every == 1 && S <= 64;
                double t0 = dt * n;
                interval_plain(kind, fatigue, scheme, N, S, t0, (t0 + dt) - t0, x, uu ? uu + (size_t)n * nu : NULL, p,
                               w + (size_t)n * N, tauc, r0_km, want_all ? xall : NULL);
                if (tr) {
                    if (trace_every == 1) memcpy(tr + ((size_t)n * S + 1) * nx, xall, sizeof(double) * nx * S);
                    else memcpy(tr + (size_t)(n + 1) * nx, x, sizeof(double) * nx);
                }
            }
            for (int j = 0; j < nx; ++j) x_final[(size_t)b * nx + j] = x[j];
        } else {
            dual x[5], pd[NP_FULL], wd[MAXN], ud[MAXN];
            for (int j = 0; j < NP_FULL; ++j) pd[j] = d_const(p[j], nd);
            int off = 0;
            for (int j = 0; j < nx; ++j) { x[j] = d_const(x0[(size_t)b * nx + j], nd); if (seed_x0) x[j].d[j] = 1.0; }
            if (seed_x0) off = nx;
            for (int j = 0; j < n_pdir; ++j) pd[pdir[j]].d[off + j] = 1.0;
            double *tr = (trace && b < trace_B) ? trace + (size_t)b * npts * nx : NULL;
            double *st = (sens_trace && b < trace_B) ? sens_trace + (size_t)b * (ns + 1) * nx * nd : NULL;
            if (tr && trace_every != 1) for (int j = 0; j < nx; ++j) tr[j] = x[j].v;
            if (st) for (int j = 0; j < nx; ++j) for (int k = 0; k < nd; ++k) st[(size_t)j * nd + k] = x[j].d[k];
            for (int n = 0; n < ns; ++n) {
                for (int i = 0; i < N; ++i) wd[i] = d_const(w[(size_t)n * N + i], nd);
                for (int i = 0; i < nu; ++i) ud[i] = d_const(uu[(size_t)n * nu + i], nd);
                double t0 = dt * n;
                interval_dual(kind, fatigue, scheme, N, S, t0, (t0 + dt) - t0, x, ud, pd, wd, tauc, r0_km, nd, NULL);
                if (tr && trace_every != 1) for (int j = 0; j < nx; ++j) tr[(size_t)(n + 1) * nx + j] = x[j].v;
                if (st) for (int j = 0; j < nx; ++j) for (int k = 0; k < nd; ++k)
                    st[((size_t)(n + 1) * nx + j) * nd + k] = x[j].d[k];
            }
            for (int j = 0; j < nx; ++j) {
                x_final[(size_t)b * nx + j] = x[j].v;
                if (sens_final) for (int k = 0; k < nd; ++k) sens_final[((size_t)b * nx + j) * nd + k] = x[j].d[k];
            }
        }
    }
    return 0;
}

/*
 * Multiple shooting: every (scenario, interval) pair independently.
 *   X [B][ns+1][nx]   U [B or 1][ns][nu]   win [B or 1][ns(+1)][N] (row stride N, node k used for interval k)
 *   G  [B][ns][nx]            g_k = x_{k+1} - Phi_k
 *   Jx [B][ns][nx][nx]        dPhi_k/dx_k          (NULL to skip)
 *   Ju [B][ns][nx][nu]        dPhi_k/du_k          (NULL to skip)
 *   Jp [B][ns][nx][n_pdir]    dPhi_k/dparams[pdir] (NULL to skip)
 *   Jw [B][ns][nx][N]         dPhi_k/dwin_k (pulse-time sensitivities; NULL to skip)
 */
static int shoot_core(int kind, int fatigue, int scheme, int N, int ns, int S, double dt, double t_start, double tauc,
                      double r0_km, long long B, const double *params, int params_shared, const double *X,
                      const double *win, int win_shared, const double *U, int u_shared, int nu,
                      int n_pdir, const int *pdir, double *G, double *Jx, double *Ju, double *Jp, double *Jw, int n_threads,
                      const double *Fc, int fc_shared, double *Jc)
{
    int nx = fatigue ? 5 : 2;
    int ndx = Jx ? nx : 0, ndu = Ju ? nu : 0, ndp = Jp ? n_pdir : 0, ndw = Jw ? N : 0, ndc = (Fc && Jc) ? 1 : 0;
    int nd = ndx + ndu + ndp + ndw + ndc;
    if (N > MAXN || nd > MAXD || (kind == HMED2018 && nu != N) || (kind == DING2007 && nu != 1)) return -1;
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
#endif
#pragma omp parallel for schedule(static)
    for (long long bk = 0; bk < B * ns; ++bk) {
        long long b = bk / ns;
        int k = (int)(bk % ns);
        const double *p = params + (params_shared ? 0 : (size_t)b * NP_FULL);
        const double *w = win + (win_shared ? 0 : (size_t)b * (ns + 1) * N) + (size_t)k * N;
        const double *uu = U ? U + (u_shared ? 0 : (size_t)b * ns * nu) + (size_t)k * nu : NULL;
        const double *xk = X + ((size_t)b * (ns + 1) + k) * nx;
        dual x[5], pd[NP_FULL], wd[MAXN], ud[MAXN];
        for (int j = 0; j < NP_FULL; ++j) pd[j] = d_const(p[j], nd);
        for (int j = 0; j < nx; ++j) { x[j] = d_const(xk[j], nd); if (ndx) x[j].d[j] = 1.0; }
        for (int i = 0; i < nu; ++i) { ud[i] = d_const(uu[i], nd); if (ndu) ud[i].d[ndx + i] = 1.0; }
        for (int j = 0; j < ndp; ++j) pd[pdir[j]].d[ndx + ndu + j] = 1.0;
        for (int i = 0; i < N; ++i) { wd[i] = d_const(w[i], nd); if (ndw) wd[i].d[ndx + ndu + ndp + i] = 1.0; }
        dual fcd;
        if (Fc) {
            fcd = d_const(Fc[(fc_shared ? 0 : (size_t)b * ns) + k], nd);
            if (ndc) fcd.d[ndx + ndu + ndp + ndw] = 1.0;
        }
        interval_dual(kind, fatigue, scheme, N, S, t_start + dt * k, dt, x, ud, pd, wd, tauc, r0_km, nd, Fc ? &fcd : NULL);
        for (int j = 0; j < nx; ++j) {
            size_t o = ((size_t)b * ns + k) * nx + j;
            G[o] = xk[nx + j] - x[j].v;
            for (int c = 0; c < ndx; ++c) Jx[o * nx + c] = x[j].d[c];
            for (int c = 0; c < ndu; ++c) Ju[o * nu + c] = x[j].d[ndx + c];
            for (int c = 0; c < ndp; ++c) Jp[o * n_pdir + c] = x[j].d[ndx + ndu + c];
            for (int c = 0; c < ndw; ++c) Jw[o * N + c] = x[j].d[ndx + ndu + ndp + c];
            if (ndc) Jc[o] = x[j].d[ndx + ndu + ndp + ndw];
        }
    }
    return 0;
}

int fes_oracle_shoot(int kind, int fatigue, int scheme, int N, int ns, int S, double dt, double t_start, double tauc,
                     double r0_km, long long B, const double *params, int params_shared, const double *X,
                     const double *win, int win_shared, const double *U, int u_shared, int nu,
                     int n_pdir, const int *pdir, double *G, double *Jx, double *Ju, double *Jp, double *Jw, int n_threads)
{
    return shoot_core(kind, fatigue, scheme, N, ns, S, dt, t_start, tauc, r0_km, B, params, params_shared, X, win, win_shared,
                      U, u_shared, nu, n_pdir, pdir, G, Jx, Ju, Jp, Jw, n_threads, NULL, 0, NULL);
}

/* The same with the force factor of f_dot_fun (ding2003.py:309-311: F' = (...) * (fl*fv + fp)) given per interval:
 *   Fc [B or 1][ns]   c_k = fl*fv + fp held constant over interval k
 *   Jc [B][ns][nx]    dPhi_k/dc_k (NULL to skip) */
int fes_oracle_shoot_fc(int kind, int fatigue, int scheme, int N, int ns, int S, double dt, double t_start, double tauc,
                        double r0_km, long long B, const double *params, int params_shared, const double *X,
                        const double *win, int win_shared, const double *U, int u_shared, int nu,
                        int n_pdir, const int *pdir, double *G, double *Jx, double *Ju, double *Jp, double *Jw, int n_threads,
                        const double *Fc, int fc_shared, double *Jc)
{
    if (Fc == NULL) return -1;
    return shoot_core(kind, fatigue, scheme, N, ns, S, dt, t_start, tauc, r0_km, B, params, params_shared, X, win, win_shared,
                      U, u_shared, nu, n_pdir, pdir, G, Jx, Ju, Jp, Jw, n_threads, Fc, fc_shared, Jc);
}

int fes_oracle_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
