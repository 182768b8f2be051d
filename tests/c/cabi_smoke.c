/* Plain C99 client of the C ABI (include/qpmpc.h, include/tamols.h): proves that the boundary needs neither C++ nor
 * Python.  Without a GPU it checks that the constructors fail loudly (no CPU fallback); with one it runs a small
 * MPPI solve and a small TAMOLS scoring call and prints the results.  Built and run by tests/test_cabi_c_client.py. */
#include <math.h>
#include <stdio.h>
#include <string.h>

#include "qpmpc.h"
#include "tamols.h"

int main(void) {
    printf("%s\n", qpmpc_version());
    qpmpc_config cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.horizon = 12;
    cfg.parametrization = QPMPC_PARAM_LINEAR;
    cfg.num_splines = 2;
    cfg.sampling_method = QPMPC_METHOD_MPPI;
    cfg.num_parallel_computations = 2048;
    cfg.max_robots = 1;
    cfg.mass = 15.019f;
    cfg.inertia[0] = 0.158f; cfg.inertia[4] = 0.469f; cfg.inertia[8] = 0.524f;
    cfg.mu = 0.5f; cfg.f_z_min = 0.f; cfg.f_z_max = 15.019f * 9.81f;
    for (int n = 0; n < 12; ++n) cfg.dts[n] = 0.02f;
    const float q[12] = {0, 0, 1500, 200, 200, 200, 500, 500, 0, 20, 20, 50};
    memcpy(cfg.q_diag, q, sizeof q);
    cfg.duty_factor = 0.65f; cfg.mpc_dt = 0.02f;
    cfg.sigma_random_sampling[0] = 0.2f; cfg.sigma_random_sampling[1] = 3.f; cfg.sigma_random_sampling[2] = 10.f;
    cfg.sigma_mppi = 3.f;
    cfg.cem_elite = 10; cfg.cem_sqrt = 1; cfg.cem_lo = 0.2f; cfg.cem_hi = 5.f;

    qpmpc_handle *h = NULL;
    int rc = qpmpc_create(&cfg, &h);
    if (rc == QPMPC_E_NODEVICE) {
        printf("no device: %s\n", qpmpc_last_error());
        tamols_params tp;
        memset(&tp, 0, sizeof tp);
        tp.gradient_delta = 0.04; tp.slope_threshold = 0.7; tp.l_min = 0.15; tp.l_max = 0.45; tp.h_des = 0.25;
        tp.stance_duration = 0.3; tp.estimated_swing_time = 0.2;
        tamols_handle *th = NULL;
        rc = tamols_create(&tp, 0, 4, &th);
        printf("tamols_create -> %d: %s\n", rc, tamols_last_error());
        return rc == TAMOLS_E_NODEVICE ? 3 : 1;  /* 3 = failed loudly on both, as it must */
    }
    if (rc != QPMPC_OK) { printf("qpmpc_create failed: %s\n", qpmpc_last_error()); return 1; }
    const int P = qpmpc_num_params(h);
    float state[24] = {0}, ref[24] = {0}, contact[48], best[36] = {0}, best_out[36];
    state[2] = ref[2] = 0.27f; ref[3] = 0.3f;
    const float feet[12] = {0.19f, 0.13f, 0, 0.19f, -0.13f, 0, -0.19f, 0.13f, 0, -0.19f, -0.13f, 0};
    memcpy(state + 12, feet, sizeof feet);
    memcpy(ref + 12, feet, sizeof feet);
    for (int k = 0; k < 48; ++k) contact[k] = 1.f;
    qpmpc_result out;
    rc = qpmpc_solve(h, state, ref, contact, best, 42, 1, NULL, 0.f, 0, NULL, NULL, NULL, &out, best_out, NULL, NULL);
    if (rc != QPMPC_OK) { printf("qpmpc_solve failed: %s\n", qpmpc_last_error()); return 1; }
    printf("P=%d best_cost=%g fz=[%g %g %g %g] launches=%lld\n", P, out.best_cost, out.grf[2], out.grf[5], out.grf[8],
           out.grf[11], (long long)qpmpc_launch_count(h));
    /* standing still on four legs: every leg carries about a quarter of the weight */
    const float quarter = 15.019f * 9.81f / 4.f;
    for (int l = 0; l < 4; ++l)
        if (!(fabsf(out.grf[3 * l + 2] - quarter) < 0.35f * quarter)) { printf("implausible fz\n"); return 1; }
    qpmpc_destroy(h);
    return 0;
}
