/*
 * fes_b200.h -- C ABI of the B200-native batched FES muscle-model integrator.
 *
 * This is the drop-in boundary for ONE hot path of Ipuch/cocofest: forward integration (plus forward
 * sensitivities) of the Ding2003 / Ding2007 / Hmed2018 muscle ODEs.  The reference has no FFI for this
 * path -- it is a closed casadi graph -- so each entry point below names the reference object it replaces
 * (paths relative to the cocofest tree):
 *
 *   fes_ivp_*      replaces the node loop of bioptim's solve_ivp_interface
 *                  (external/bioptim/bioptim/interfaces/solve_ivp_interface.py:56-69) driving
 *                  nlp.dynamics[node] (external/bioptim/bioptim/dynamics/integrator.py:82-104, 300-365), as
 *                  reached from IvpFes.integrate (cocofest/integration/ivp_fes.py:282-297), batched over
 *                  independent trajectories, optionally with chained forward sensitivities
 *                  d x(t_n) / d x0 and d x(t_n) / d params.
 *   fes_shoot_*    replaces the continuity-defect evaluation g_k = x_{k+1} - Phi_k(x_k, u_k, p)
 *                  (external/bioptim/bioptim/limits/penalty.py:1056-1105) and casadi's AD of it
 *                  (the Jacobian blocks dPhi/dx_k, dPhi/du_k, dPhi/dp that nlpsol builds in
 *                  external/bioptim/bioptim/interfaces/interface_utils.py:88-95), batched over
 *                  (scenario, interval) pairs.
 *                  With FES_DIR_C the F row carries the factor fl*fv + fp of f_dot_fun
 *                  (cocofest/models/ding2003.py:309-311) that the FES-MSK models evaluate from the joint states
 *                  (cocofest/models/dynamical_model.py:259-310), as an input with its own Jacobian column.
 *   fes_identify_* replaces the force-tracking objective and the parameter search of the identification OCP
 *                  (cocofest/optimization/fes_identification_ocp.py:248-299), one candidate per thread.
 *   model arithmetic: cocofest/models/ding2003.py:200-311, ding2003_with_fatigue.py:138-240,
 *                  ding2007.py:121-188, ding2007_with_fatigue.py:133-241, hmed2018.py:119-180,
 *                  hmed2018_with_fatigue.py:124-229.
 *
 * Conventions
 *   - every function returns int: 0 = OK, < 0 = invalid argument (fes_strerror), > 0 = cudaError_t.
 *   - fes_ivp_* / fes_shoot_* / fes_identify_* take DEVICE pointers owned by the caller; the library allocates
 *     nothing persistent and keeps no global state; calls are re-entrant; work is enqueued on `stream`
 *     (a cudaStream_t passed as void*) and NOT synchronised.  (fes_shoot_static_f64 is the exception: it keeps its
 *     tables in a static device array shared by all calls on a device, see its declaration.)
 *   - fes_*_host_* take HOST pointers (pinned for full speed), stage through stream-ordered temporary
 *     device allocations from a library-owned pool that stays mapped between calls (fes_host_pool_trim
 *     releases it), and return after the results are in the host buffers.
 *   - device layout is structure-of-arrays with the BATCH INDEX INNERMOST (coalesced): an array written
 *     [a][b][B] below stores element (i, j, batch) at ((i*b_dim + j)*B + batch).  "shared" arrays drop
 *     the batch axis entirely.
 *   - state order: Cn, F (, A, Tau1, Km)   (nx = 2 without fatigue, 5 with fatigue).
 */
#ifndef FES_B200_H
#define FES_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FES_ABI_VERSION 1

/* model_kind */
#define FES_DING2003 0 /* DingModelFrequency[WithFatigue]                (ding2003*.py)  nu = 0 */
#define FES_DING2007 1 /* DingModelPulseWidthFrequency[WithFatigue]      (ding2007*.py)  nu = 1 */
#define FES_HMED2018 2 /* DingModelPulseIntensityFrequency[WithFatigue]  (hmed2018*.py)  nu = N */

/* canonical parameter vector, rows of params[FES_NP][B] */
#define FES_NP 14
#define FES_P_A_REST 0 /* a_rest; a_scale for FES_DING2007 (ding2007_with_fatigue.py:108,211) */
#define FES_P_KM_REST 1
#define FES_P_TAU1_REST 2
#define FES_P_TAU2 3
#define FES_P_ALPHA_A 4
#define FES_P_ALPHA_TAU1 5
#define FES_P_ALPHA_KM 6
#define FES_P_TAU_FAT 7
#define FES_P_PD0 8
#define FES_P_PDT 9
#define FES_P_AR 10
#define FES_P_BS 11
#define FES_P_IS 12
#define FES_P_CR 13

/* sensitivity direction groups (FesDesc.dirs is ONE of the supported sets, see fes_num_dirs) */
#define FES_DIR_X0 1  /* d / d initial state of the interval (nx columns)                                  */
#define FES_DIR_U 2   /* d / d control: pulse width (1 column) or FES_TILE pulse intensities from tile_offset */
#define FES_DIR_P 4   /* d / d identifiable parameters: a_rest, km_rest, tau1_rest, tau2 [, pd0, pdt | ar, bs, Is, cr] */
#define FES_DIR_W 8   /* d / d pulse times of the window: FES_TILE columns from tile_offset                 */
#define FES_DIR_PF 16 /* with FES_DIR_P, fatigue models: + alpha_a, alpha_tau1, alpha_km, tau_fat          */
#define FES_DIR_C 32  /* shoot only, with X0|U (FES_HMED2018 also U alone, for the intensity tiles after the first): the force factor c = fl*fv + fp of f_dot_fun (ding2003.py:309-311),
                       * F' = c (A q - F/(tau1 + tau2 q)), is an INPUT -- ctrl carries one more row per node after the model's
                       * own controls, [ns][nu+1][B], holding c of the interval -- and J gets one column d Phi / d c after
                       * the U columns.  The reference evaluates fl, fv, fp from the joint states inside the same ODE
                       * (FES-MSK models); a caller that integrates the joint dynamics itself passes their values here. */
#define FES_TILE 5

/* flags */
#define FES_PARAMS_SHARED 1 /* params is [FES_NP] instead of [FES_NP][B]                  */
#define FES_WIN_SHARED 2    /* stim_win is [rows][N] instead of [rows][N][B]              */
#define FES_CTRL_SHARED 4   /* ctrl is [ns][nu] instead of [ns][nu][B]                    */
#define FES_TABLES_READY 8  /* workspace already holds the tables of this (desc, stim_win) */
#define FES_TARGET_SHARED 16 /* identify: target is [ns+1] instead of [ns+1][B]              */

typedef struct FesDesc {
    int32_t abi_version; /* FES_ABI_VERSION */
    int32_t model_kind;  /* FES_DING2003 | FES_DING2007 | FES_HMED2018 */
    int32_t fatigue;     /* 0 | 1 */
    int32_t scheme;      /* 1 = RK1, 2 = RK2 (midpoint), 4 = RK4  (bioptim integrator.py:324-365) */
    int32_t n_trunc;     /* N = sum_stim_truncation, window length, 1..64 (1..1024 for FES_DING2003 / FES_DING2007 without FES_DIR_W) */
    int32_t n_shooting;  /* ns = number of intervals */
    int32_t n_steps;     /* S = RK steps per interval (n_integration_steps), 1..256 */
    int32_t dirs;        /* sensitivity groups, 0 = none */
    int32_t flags;
    int32_t tile_offset; /* first window slot of the FES_DIR_U (hmed) / FES_DIR_W tile */
    int32_t trace_every; /* ivp only: 0 = no trace, 1 = every RK step (ns*S+1 points), S = nodes (ns+1 points) */
    int32_t reserved0;
    int64_t batch;       /* B: trajectories (ivp) or scenarios (shoot) */
    int64_t trace_batch; /* ivp only: the first trace_batch trajectories are traced */
    double dt;           /* interval length (final_time / n_shooting) */
    double t_start;      /* time of node 0 */
    double tauc;         /* 0.020 (ding2003, hmed2018) / 0.011 (ding2007) */
    double r0_km;        /* r0_km_relationship = 1.04 (ding2003.py:52) */
} FesDesc;

/* Buffers of one ivp call.  T = double (_f64) or float (_f32); stim_win is always double. */
typedef struct FesIvpBuffers {
    const void *params;     /* [FES_NP][B] | [FES_NP]                                         */
    const void *x0;         /* [nx][B]                                                        */
    const double *stim_win; /* [>=ns][N][B] | [>=ns][N]  window of node n = pulses <= t_n, oldest first */
    const void *ctrl;       /* [ns][nu][B] | [ns][nu] | NULL when nu = 0                      */
    void *x_final;          /* [nx][B]                                                        */
    void *trace;            /* [npts][nx][trace_batch] | NULL                                 */
    void *sens_final;       /* [nx][nd][B] | NULL when dirs = 0                               */
    void *sens_trace;       /* [ns+1][nx][nd][trace_batch] at nodes | NULL                    */
    void *workspace;        /* fes_workspace_bytes(desc) bytes of device memory               */
    int64_t workspace_bytes;
} FesIvpBuffers;

/* Buffers of one shoot call.  Columns of J follow the group order X0, U, P(+PF), W of desc.dirs. */
typedef struct FesShootBuffers {
    const void *params;     /* [FES_NP][B] | [FES_NP]                 */
    const void *X;          /* [ns+1][nx][B]  decision states at nodes */
    const double *stim_win; /* [>=ns][N][B] | [>=ns][N]               */
    const void *ctrl;       /* [ns][nu][B] | [ns][nu] | NULL; nu + 1 rows with FES_DIR_C */
    void *G;                /* [ns][nx][B]  g_k = x_{k+1} - Phi_k  | NULL */
    void *Phi;              /* [ns][nx][B]  end point of every interval | NULL */
    void *J;                /* [ns][nx][j_cols][B], this call fills columns j_col0 .. j_col0+nd-1 | NULL when dirs = 0 */
    int32_t j_cols;         /* total number of columns of J */
    int32_t j_col0;         /* first column written by this call */
    void *workspace;
    int64_t workspace_bytes;
} FesShootBuffers;

/* Buffers of one identify call (two-state models only: the reference rejects fatigue models for
 * identification, identification/ding2003_force_parameter_identification.py:196-199).
 * Column d of the parameter group P is a_rest|a_scale, km_rest, tau1_rest, tau2 [, pd0, pdt | ar, bs, Is, cr]. */
typedef struct FesIdentBuffers {
    const void *params;     /* [FES_NP][B] | [FES_NP]  starting point of every candidate (all 14 rows)          */
    const void *x0;         /* [2][B] | NULL = rest (Cn = F = 0)                                               */
    const double *stim_win; /* [>=ns][N][B] | [>=ns][N]                                                        */
    const void *ctrl;       /* [ns][nu][B] | [ns][nu] | NULL when nu = 0                                       */
    const void *target;     /* [ns+1][B] | [ns+1]  force data resampled at the nodes (force_at_node_in_ocp)    */
    const void *weight;     /* [ns+1] node weights | NULL = dt each (Lagrange TRACK_STATE, Node.ALL)           */
    const void *lower;      /* [FES_NP] lower bounds of the canonical parameters | NULL                        */
    const void *upper;      /* [FES_NP] upper bounds | NULL                                                    */
    void *params_out;       /* [FES_NP][B] identified parameters (rows that are not identified are copied)     */
    void *cost;             /* [B]  sum_j w_j (F(t_j) - target_j)^2 at params_out                              */
    void *grad;             /* [np][B]  d cost / d P-column at params_out                                      */
    void *hess;             /* [np(np+1)/2][B]  Gauss-Newton matrix 2 J^T W J, upper triangle row-major        */
    int32_t *iters;         /* [B] number of cost evaluations spent | NULL                                     */
    void *workspace;
    int64_t workspace_bytes;
} FesIdentBuffers;

/* Buffers of one second-order call (fes_shoot_hess_f64).  Directions z = (x_k, u_k [, c_k]): the nx states of node k, the
 * model's controls of interval k (none | pulse width | the N pulse intensities of the window) and, with FES_DIR_C, the force
 * factor; nz = nx + nu (+ 1).  Pairs (i <= j) are numbered along the upper triangle, row-major: p = 0 is (0,0), p = 1 is
 * (0,1), ..., p = nz is (1,1), ...; npairs = nz (nz + 1) / 2. */
typedef struct FesHessBuffers {
    const void *params;      /* [FES_NP][B] | [FES_NP]                                              */
    const void *X;           /* [ns+1][nx][B]  decision states at nodes                             */
    const double *stim_win;  /* [>=ns][N][B] | [>=ns][N]                                            */
    const void *ctrl;        /* [ns][nu][B] | [ns][nu] | NULL; nu + 1 rows with FES_DIR_C           */
    const void *multipliers; /* [ns][nx][B]  lambda of the continuity rows | NULL when H is NULL    */
    void *D2;                /* [ns][nx][npairs][B]  d2 Phi_k[r] / dz_i dz_j | NULL                  */
    void *H;                 /* [ns][npairs][B]  sum_r lambda[k][r] d2 Phi_k[r] / dz_i dz_j | NULL   */
} FesHessBuffers;

typedef struct FesIdentOptions {
    int32_t max_iter; /* Levenberg-Marquardt iterations; 0 = evaluate cost / gradient / Gauss-Newton matrix only */
    int32_t p_mask;   /* bit d set = P-column d is identified (key_parameter_to_identify)                      */
    double lambda0;   /* initial damping (Marquardt scaling: H + lambda diag(H)); e.g. 1e-3                    */
    double ftol;      /* stop when an accepted step lowers the cost by less than ftol * cost                   */
    double xtol;      /* stop when an accepted step moves every parameter by less than xtol (relative)         */
} FesIdentOptions;

int fes_version(void);
const char *fes_strerror(int code);

/* nx of a model; number of sensitivity columns of a direction set (< 0: unsupported set).
 * Supported sets: 0, X0, X0|U, X0|U|C, U, U|C (FES_HMED2018), P, P|PF, W   (ivp: 0, X0, P, P|PF). */
int fes_num_states(int fatigue);
int fes_num_controls(int model_kind, int n_trunc);
int fes_num_dirs(int model_kind, int fatigue, int dirs);

/* device workspace (window tables + stage decay table) needed by a call described by desc */
int64_t fes_workspace_bytes(const FesDesc *desc, int is_f32);

int fes_ivp_f64(const FesDesc *desc, const FesIvpBuffers *buf, void *stream);
int fes_ivp_f32(const FesDesc *desc, const FesIvpBuffers *buf, void *stream);
int fes_shoot_f64(const FesDesc *desc, const FesShootBuffers *buf, void *stream);
int fes_shoot_f32(const FesDesc *desc, const FesShootBuffers *buf, void *stream);

/* fes_shoot_f64 for callers that must not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.  The window
 * tables and the stage table, which fes_shoot_f64 keeps in the caller's workspace, live in a static device array of
 * FES_STATIC_SCRATCH_BYTES that every GPU which loads the library holds until the process exits; buf->workspace and
 * buf->workspace_bytes are ignored, inputs and outputs are the caller's device buffers as in fes_shoot_f64, and the
 * results are bit-identical to fes_shoot_f64's.  The call allocates and frees nothing (no cudaMalloc / cudaFree /
 * pool): it launches the table kernel and the interval kernel on `stream` and returns.
 *   - returns -8 (see fes_strerror) when fes_workspace_bytes(desc, 0) > FES_STATIC_SCRATCH_BYTES = 4 MiB: a shared
 *     window (FES_WIN_SHARED) needs n_shooting * r + 6 n_steps + 1 doubles, r = 2 + 2 n_trunc for FES_HMED2018 or
 *     FES_DIR_W and 2 otherwise; per-scenario windows need batch times the first term;
 *   - the first call on a device resolves the array's device address (cudaGetSymbolAddress, once per device): make one
 *     call before a stream capture begins;
 *   - ALL calls on a device share the array: calls that may overlap on the GPU (different streams, or graphs replayed
 *     concurrently) race on it -- keep them on one stream or order them with events;
 *   - the array keeps the last call's tables; every call rebuilds all it reads before reading it (nothing has to start
 *     at zero), and FES_TABLES_READY is honoured only in that sense: set it only when the previous call through THIS
 *     entry point on this device had the same (desc, stim_win). */
#define FES_STATIC_SCRATCH_BYTES (4 * 1024 * 1024)
int fes_shoot_static_f64(const FesDesc *desc, const FesShootBuffers *buf, void *stream);

/* Second derivatives of the interval map, d2 Phi_k / dz_k dz_k, for an exact-Hessian solver: the reference's Ipopt default
 * is casadi's exact Hessian of the Lagrangian (external/bioptim/bioptim/interfaces/ipopt_options.py:88), i.e. casadi's second
 * derivative of the unrolled integrator inside every continuity row (external/bioptim/bioptim/limits/penalty.py:1056-1105).
 * One thread per (scenario, interval, pair of directions), hyper-dual arithmetic through the same RK stages: the exact
 * second-order tangent of the discrete map.  desc->dirs = FES_DIR_X0 | FES_DIR_U [| FES_DIR_C]; fp64 only; needs no workspace
 * (desc->tile_offset, trace fields and FES_TABLES_READY are ignored) and allocates nothing.  The contribution of the
 * continuity rows g_k = x_{k+1} - Phi_k to the Hessian of the Lagrangian is MINUS buf->H. */
int fes_shoot_hess_num_dirs(int model_kind, int fatigue, int n_trunc, int dirs);
int fes_shoot_hess_f64(const FesDesc *desc, const FesHessBuffers *buf, void *stream);

/* Force-tracking cost (fes_identification_ocp.py:248-266) with its gradient and Gauss-Newton matrix, and the
 * bounded Levenberg-Marquardt identification of the P-group parameters (fes_identification_ocp.py:269-299), one
 * thread per candidate, whole loop on device.  desc->dirs must be FES_DIR_P, desc->fatigue 0. */
int fes_identify_f64(const FesDesc *desc, const FesIdentBuffers *buf, const FesIdentOptions *opt, void *stream);
int fes_identify_f32(const FesDesc *desc, const FesIdentBuffers *buf, const FesIdentOptions *opt, void *stream);

/* Same calls with HOST buffers (workspace fields ignored): H2D, kernels, D2H, stream synchronised. */
int fes_ivp_host_f64(const FesDesc *desc, const FesIvpBuffers *host_buf, void *stream);
int fes_ivp_host_f32(const FesDesc *desc, const FesIvpBuffers *host_buf, void *stream);
int fes_shoot_host_f64(const FesDesc *desc, const FesShootBuffers *host_buf, void *stream);
int fes_shoot_host_f32(const FesDesc *desc, const FesShootBuffers *host_buf, void *stream);
int fes_identify_host_f64(const FesDesc *desc, const FesIdentBuffers *host_buf, const FesIdentOptions *opt, void *stream);
int fes_identify_host_f32(const FesDesc *desc, const FesIdentBuffers *host_buf, const FesIdentOptions *opt, void *stream);
int fes_shoot_hess_host_f64(const FesDesc *desc, const FesHessBuffers *host_buf, void *stream);

/* Roofline denominators measured on the device the caller made current: runs `launches` launches of a
 * register-only FMA kernel on `stream` and returns the best FMA-flop rate (2 flop per FMA) in TFLOP/s. */
int fes_probe_fma_peak(int is_f32, int launches, void *stream, double *tflops_out);

/* Kernel launches issued by this library in the calling process since load (for bench accounting). */
int64_t fes_launch_count(void);

/* Phases of the last fes_*_host_* call of the calling thread, in milliseconds:
 *   ms5[0] allocation of the staging buffers (host clock), ms5[1] host-to-device copies, ms5[2] kernels,
 *   ms5[3] device-to-host copies (CUDA events on `stream`), ms5[4] the whole call (host clock). */
int fes_last_host_timing(double *ms5);

/* The fes_*_host_* calls stage through a memory pool owned by the library (one per device) that keeps its memory
 * mapped between calls; this returns it to the driver (current device). */
int fes_host_pool_trim(void);

#ifdef __cplusplus
}
#endif
#endif /* FES_B200_H */
