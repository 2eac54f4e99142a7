/* abi_client.c -- a plain-C caller of libfes_b200.so (no Python, no torch): what a maintainer binding the
 * library from another host language would write.  It reproduces the reference's own golden case
 * (tests/shard1/test_ivp.py:28-44: DingModelFrequencyWithFatigue, stim [0, 0.1, 0.2], sum_stim_truncation 3,
 * final_time 0.3, RK4 with 10 sub-steps) through fes_ivp_host_f64 and prints the 31 force values, one per line;
 * tests/test_c_abi.py compares them with tests/golden/ivp_golden.json.
 *
 *   abi_client            run the case; exit code 0, or 10 + (library return code != 0)
 *   abi_client --errors   exercise the argument checks only (no device work): exit 0 when every bad call is refused
 *   abi_client --hess     the three intervals through fes_shoot_hess_host_f64 (second derivatives contracted with multipliers of one)
 *   abi_client --factor   the three intervals of the same case through fes_shoot_host_f64 with the force factor of every
 *                         interval as an input (FES_DIR_X0 | FES_DIR_C): prints Phi (3 x 5), then dPhi/dfactor (3 x 5)
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "fes_b200.h"

#define PAD (-10000000.0) /* ding2003.py:396 */

static void fill_desc(FesDesc *d) {
    memset(d, 0, sizeof(*d));
    d->abi_version = FES_ABI_VERSION;
    d->model_kind = FES_DING2003;
    d->fatigue = 1;
    d->scheme = 4;
    d->n_trunc = 3;
    d->n_shooting = 3;
    d->n_steps = 10;
    d->dirs = 0;
    d->flags = FES_PARAMS_SHARED | FES_WIN_SHARED;
    d->trace_every = 1;
    d->batch = 1;
    d->trace_batch = 1;
    d->dt = 0.3 / 3;
    d->t_start = 0.0;
    d->tauc = 0.020;
    d->r0_km = 1.04;
}

int main(int argc, char **argv) {
    /* canonical parameters (ding2003.py:51-66, ding2003_with_fatigue.py:50-59, Python literals 10e-2 / 10e-6 kept) */
    double params[FES_NP] = {3009, 0.103, 0.050957, 0.060, -4.0 * 10e-2, 2.1 * 10e-6, 1.9 * 10e-6, 127,
                             0.000131405, 0.000194138, 0.586, 0.026, 63.1, 0.833};
    double x0[5] = {0, 0, 3009, 0.050957, 0.103};                 /* standard_rest_values */
    /* window of node n = the last 3 pulses <= t_n, left-padded (ding2003.py:394-429) */
    double win[3][3] = {{PAD, PAD, 0.0}, {PAD, 0.0, 0.1}, {0.0, 0.1, 0.2}};
    double x_final[5], trace[31][5];
    FesDesc d;
    FesIvpBuffers b;
    fill_desc(&d);
    memset(&b, 0, sizeof(b));
    b.params = params;
    b.x0 = x0;
    b.stim_win = &win[0][0];
    b.x_final = x_final;
    b.trace = trace;

    if (fes_version() != FES_ABI_VERSION) return 2;
    if (argc > 1 && strcmp(argv[1], "--errors") == 0) {
        FesDesc bad;
        int refused = 0;
        bad = d; bad.abi_version = 99;  refused += fes_ivp_host_f64(&bad, &b, NULL) < 0;
        bad = d; bad.scheme = 3;        refused += fes_ivp_host_f64(&bad, &b, NULL) < 0;
        bad = d; bad.n_trunc = 1025;     refused += fes_ivp_host_f64(&bad, &b, NULL) < 0;
        bad = d; bad.n_steps = 257;     refused += fes_ivp_host_f64(&bad, &b, NULL) < 0;
        bad = d; bad.dirs = FES_DIR_U;  refused += fes_ivp_host_f64(&bad, &b, NULL) < 0;
        bad = d; bad.model_kind = 7;    refused += fes_ivp_host_f64(&bad, &b, NULL) < 0;
        bad = d; bad.batch = 0;         refused += fes_ivp_host_f64(&bad, &b, NULL) < 0;
        refused += fes_ivp_host_f64(&d, NULL, NULL) < 0;
        refused += fes_ivp_host_f64(NULL, &b, NULL) < 0;
        refused += fes_num_dirs(FES_DING2003, 0, FES_DIR_U) < 0;
        /* the force factor rides with the X0 | U sets of the interval entry points only */
        bad = d; bad.dirs = FES_DIR_X0 | FES_DIR_C; refused += fes_ivp_host_f64(&bad, &b, NULL) < 0;
        /* per-scenario windows of 2^20 scenarios do not fit the static scratch: refused with its own code before any device work */
        {
            FesShootBuffers sb;
            memset(&sb, 0, sizeof(sb));
            bad = d; bad.flags = 0; bad.batch = 1 << 20; bad.dirs = FES_DIR_X0; bad.trace_every = 0; bad.trace_batch = 0;
            refused += fes_shoot_static_f64(&bad, &sb, NULL) == -8 && strstr(fes_strerror(-8), "FES_STATIC_SCRATCH_BYTES") != NULL;
        }
        /* the second-order entry point wants all of z = (x, u): a parameter direction set is refused before any device work */
        {
            FesHessBuffers hb;
            double dummy[1];
            memset(&hb, 0, sizeof(hb));
            hb.params = params; hb.X = dummy; hb.stim_win = &win[0][0]; hb.D2 = dummy;
            bad = d; bad.dirs = FES_DIR_P; bad.trace_every = 0; bad.trace_batch = 0;
            refused += fes_shoot_hess_host_f64(&bad, &hb, NULL) == -4;
        }
        printf("refused %d of 13; e.g. \"%s\"\n", refused, fes_strerror(-1));
        return refused == 13 ? 0 : 3;
    }
    if (argc > 1 && strcmp(argv[1], "--factor") == 0) {
        /* node states: the rest state at every node; ctrl = one row per node, the factor fl*fv + fp of the interval */
        double X[4][5], factor[3] = {1.0, 0.8, 1.15}, Phi[3][5], J[3][5][6];
        FesShootBuffers sb;
        for (int k = 0; k < 4; ++k) memcpy(X[k], x0, sizeof(x0));
        memset(&sb, 0, sizeof(sb));
        sb.params = params;
        sb.X = X;
        sb.stim_win = &win[0][0];
        sb.ctrl = factor;
        sb.Phi = Phi;
        sb.J = J;
        sb.j_cols = 6;
        d.dirs = FES_DIR_X0 | FES_DIR_C;
        d.flags |= FES_CTRL_SHARED;
        d.trace_every = 0;
        d.trace_batch = 0;
        int rcf = fes_shoot_host_f64(&d, &sb, NULL);
        if (rcf != 0) {
            fprintf(stderr, "fes_shoot_host_f64 -> %d (%s)\n", rcf, fes_strerror(rcf));
            return 10;
        }
        for (int k = 0; k < 3; ++k) for (int j = 0; j < 5; ++j) printf("%.17g\n", Phi[k][j]);
        for (int k = 0; k < 3; ++k) for (int j = 0; j < 5; ++j) printf("%.17g\n", J[k][j][5]);
        return 0;
    }
    if (argc > 1 && strcmp(argv[1], "--hess") == 0) {
        /* second derivatives of the three interval maps at the rest state, contracted with multipliers of one: 15 pairs each */
        double X[4][5], lam[3][5], H[3][15];
        FesHessBuffers hb;
        for (int k = 0; k < 4; ++k) memcpy(X[k], x0, sizeof(x0));
        for (int k = 0; k < 3; ++k) for (int j = 0; j < 5; ++j) { lam[k][j] = 1.0; X[k][0] = 0.2; X[k][1] = 30.0 + 10.0 * k; }
        memset(&hb, 0, sizeof(hb));
        hb.params = params;
        hb.X = X;
        hb.stim_win = &win[0][0];
        hb.multipliers = lam;
        hb.H = H;
        d.dirs = FES_DIR_X0 | FES_DIR_U;
        d.trace_every = 0;
        d.trace_batch = 0;
        if (fes_shoot_hess_num_dirs(d.model_kind, d.fatigue, d.n_trunc, d.dirs) != 5) return 4;
        int rch = fes_shoot_hess_host_f64(&d, &hb, NULL);
        if (rch != 0) {
            fprintf(stderr, "fes_shoot_hess_host_f64 -> %d (%s)\n", rch, fes_strerror(rch));
            return 10;
        }
        for (int k = 0; k < 3; ++k) for (int p = 0; p < 15; ++p) printf("%.17g\n", H[k][p]);
        return 0;
    }
    int rc = fes_ivp_host_f64(&d, &b, NULL);
    if (rc != 0) {
        fprintf(stderr, "fes_ivp_host_f64 -> %d (%s)\n", rc, fes_strerror(rc));
        return 10;
    }
    for (int i = 0; i < 31; ++i) printf("%.17g\n", trace[i][1]);
    return 0;
}
