/* abi_client.c -- a plain-C caller of libfes_b200.so (no Python, no torch): what a maintainer binding the
 * library from another host language would write.  It reproduces the reference's own golden case
 * (tests/shard1/test_ivp.py:28-44: DingModelFrequencyWithFatigue, stim [0, 0.1, 0.2], sum_stim_truncation 3,
 * final_time 0.3, RK4 with 10 sub-steps) through fes_ivp_host_f64 and prints the 31 force values, one per line;
 * tests/test_c_abi.py compares them with tests/golden/ivp_golden.json.
 *
 *   abi_client            run the case; exit code 0, or 10 + (library return code != 0)
 *   abi_client --errors   exercise the argument checks only (no device work): exit 0 when every bad call is refused
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
        printf("refused %d of 10; e.g. \"%s\"\n", refused, fes_strerror(-1));
        return refused == 10 ? 0 : 3;
    }
    int rc = fes_ivp_host_f64(&d, &b, NULL);
    if (rc != 0) {
        fprintf(stderr, "fes_ivp_host_f64 -> %d (%s)\n", rc, fes_strerror(rc));
        return 10;
    }
    for (int i = 0; i < 31; ++i) printf("%.17g\n", trace[i][1]);
    return 0;
}
