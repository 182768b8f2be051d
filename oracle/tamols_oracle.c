/*
 * oracle/tamols_oracle.c -- CPU restatement of the reference's TAMOLS candidate loop.  TEST INFRASTRUCTURE: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may call this; the product
 * path never does.
 *
 * Follows /root/reference/quadruped_pympc/helpers/visual_foothold_adaptation.py (float64, scalar, one candidate at a
 * time, operation order of the Python source); each function cites the lines it restates.  The only deliberate
 * deviations: (1) the cKDTree nearest-neighbour query (:29-35) is the analytic nearest cell of the regular grid,
 * clamped to the candidate window; (2) np.linalg.lstsq on the symmetric 3x3 design (:507-510) is its closed form.
 * Both are pinned against the reference itself by tests/golden/tamols_*.npz (oracle/make_golden_tamols.py runs the
 * reference's own file under stubs).
 * Build: oracle/Makefile (-ffp-contract=off so no FMA contraction changes the rounding).
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>

#include "../include/tamols.h"

typedef struct {
    const double *heights; /* [num_maps, rows, pitch] */
    const double *xy;      /* [num_maps, rows, cols, 2] or NULL */
    int rows, cols, pitch;
} maps_t;

typedef struct {
    const tamols_params *p;
    const tamols_leg_job *j;
    const maps_t *m;
    double arx, ary, acx, acy; /* axes divided by their squared length: grid coordinate = d . a */
} ctx_t;

static void ctx_init(ctx_t *c, const tamols_params *p, const tamols_leg_job *j, const maps_t *m) {
    c->p = p;
    c->j = j;
    c->m = m;
    const double r2 = j->row_axis[0] * j->row_axis[0] + j->row_axis[1] * j->row_axis[1];
    const double c2 = j->col_axis[0] * j->col_axis[0] + j->col_axis[1] * j->col_axis[1];
    c->arx = j->row_axis[0] / r2;
    c->ary = j->row_axis[1] / r2;
    c->acx = j->col_axis[0] / c2;
    c->acy = j->col_axis[1] / c2;
}

static int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* FastHeightMap.get_height (:31-35): nearest cell of the window + 0.02 */
static double get_height(const ctx_t *c, double x, double y) {
    const tamols_leg_job *j = c->j;
    const double dx = x - j->origin[0], dy = y - j->origin[1];
    const double gr = dx * c->arx + dy * c->ary;
    const double gc = dx * c->acx + dy * c->acy;
    const int ir = clampi((int)rint(fmin(fmax(gr, -1e9), 1e9)), j->r0, j->r0 + j->nr - 1);
    const int ic = clampi((int)rint(fmin(fmax(gc, -1e9), 1e9)), j->c0, j->c0 + j->nc - 1);
    const double h = c->m->heights[((size_t)j->map * c->m->rows + ir) * c->m->pitch + ic];
    return h + c->p->sensor_offset;
}

/* _check_kinematic_feasibility (:375-395) */
static int kinematic_feasible(const ctx_t *c, const double cand[3]) {
    const tamols_params *p = c->p;
    const tamols_leg_job *j = c->j;
    double d0 = cand[0] - j->hip[0], d1 = cand[1] - j->hip[1], d2 = cand[2] - j->hip[2];
    const double dist_td = sqrt((d0 * d0 + d1 * d1) + d2 * d2);
    if (!(p->l_min <= dist_td && dist_td <= p->l_max)) return 0;
    if (j->flags & TAMOLS_FLAG_HAS_VELOCITY) {
        d0 = cand[0] - (j->hip[0] + j->vel[0] * p->stance_duration);
        d1 = cand[1] - (j->hip[1] + j->vel[1] * p->stance_duration);
        d2 = cand[2] - (j->hip[2] + j->vel[2] * p->stance_duration);
        const double dist_lo = sqrt((d0 * d0 + d1 * d1) + d2 * d2);
        if (!(p->l_min <= dist_lo && dist_lo <= p->l_max)) return 0;
    }
    return 1;
}

/* _check_leg_collision (:397-420) */
static int leg_collision(const ctx_t *c, const double cand[3]) {
    const tamols_params *p = c->p;
    const tamols_leg_job *j = c->j;
    for (int k = 0; k < 5; ++k) {
        const double a = p->collision_alphas[k], b = 1.0 - a;
        const double px = b * j->hip[0] + a * cand[0];
        const double py = b * j->hip[1] + a * cand[1];
        const double pz = b * j->hip[2] + a * cand[2];
        const double h_ground = get_height(c, px, py) - p->sensor_offset;
        if (pz < h_ground + p->collision_margin) return 1;
    }
    return 0;
}

/* _compute_edge_cost (:422-466) */
static double edge_cost(const ctx_t *c, const double cand[3]) {
    const double d = c->p->gradient_delta;
    const double h0 = get_height(c, cand[0] + d, cand[1] + 0.0);
    const double h1 = get_height(c, cand[0] + -d, cand[1] + 0.0);
    const double h2 = get_height(c, cand[0] + 0.0, cand[1] + d);
    const double h3 = get_height(c, cand[0] + 0.0, cand[1] + -d);
    const double gx = fabs(h0 - h1) / (2.0 * d);
    const double gy = fabs(h2 - h3) / (2.0 * d);
    const double g = sqrt(gx * gx + gy * gy);
    return g <= c->p->slope_threshold ? 0.0 : g - c->p->slope_threshold;
}

/* _compute_roughness_cost (:468-521): plane [x, y, 1] least squares on the 3x3 patch, variance of the residual */
static double roughness_cost(const ctx_t *c, const double cand[3]) {
    const double d = c->p->gradient_delta;
    double h[9], sxh = 0.0, syh = 0.0, sh = 0.0;
    for (int i = -1; i <= 1; ++i)
        for (int k = -1; k <= 1; ++k) {
            const double v = get_height(c, cand[0] + i * d, cand[1] + k * d);
            h[(i + 1) * 3 + (k + 1)] = v;
            sxh += (i * d) * v;
            syh += (k * d) * v;
            sh += v;
        }
    const double s2 = 6.0 * (d * d);
    const double a = sxh / s2, b = syh / s2, c0 = sh / 9.0;
    double res[9], mean = 0.0;
    for (int i = -1; i <= 1; ++i)
        for (int k = -1; k <= 1; ++k) {
            const int q = (i + 1) * 3 + (k + 1);
            res[q] = h[q] - (((i * d) * a + (k * d) * b) + c0);
            mean += res[q];
        }
    mean /= 9.0;
    double var = 0.0;
    for (int q = 0; q < 9; ++q) var += (res[q] - mean) * (res[q] - mean);
    return var / 9.0;
}

/* _compute_nominal_kinematic_cost (:523-553) */
static double nominal_kinematic_cost(const ctx_t *c, const double cand[3]) {
    const tamols_leg_job *j = c->j;
    const double d0 = j->hip[0] - (cand[0] - 0.0), d1 = j->hip[1] - (cand[1] - 0.0);
    const double d2 = j->hip[2] - (cand[2] - (-c->p->h_des));
    return (d0 * d0 + d1 * d1) + d2 * d2;
}

/* _compute_reference_tracking_cost (:555-609) */
static double reference_tracking_cost(const ctx_t *c, const double cand[3]) {
    const tamols_leg_job *j = c->j;
    const double dx = cand[0] - j->seed[0];
    if (!(j->flags & TAMOLS_FLAG_HAS_VELOCITY)) return dx < 0.0 ? dx * dx : 0.0;
    const double vm = sqrt(j->vel[0] * j->vel[0] + j->vel[1] * j->vel[1]);
    if (vm < 0.01) return 0.0;
    if (j->vel[0] > 0.0 && dx < 0.0) return dx * dx;
    if (j->vel[0] < 0.0 && dx > 0.0) return dx * dx;
    return 0.0;
}

/* _compute_stability_cost (:611-714) */
static double stability_cost(const ctx_t *c, const double cand[3]) {
    const tamols_params *p = c->p;
    const tamols_leg_job *j = c->j;
    if (!(j->flags & TAMOLS_FLAG_STABILITY)) return 0.0;
    const double comx = j->base_xy[0] + j->vel[0] * p->estimated_swing_time;
    const double comy = j->base_xy[1] + j->vel[1] * p->estimated_swing_time;
    const double vx = j->diag_foot[0] - cand[0], vy = j->diag_foot[1] - cand[1];
    const double wx = comx - cand[0], wy = comy - cand[1];
    const double vv = vx * vx + vy * vy;
    double dist;
    if (vv < 1e-8) {
        dist = sqrt(wx * wx + wy * wy);
    } else {
        double t = (wx * vx + wy * vy) / vv;
        t = t < 0.0 ? 0.0 : (t > 1.0 ? 1.0 : t);
        const double ex = comx - (cand[0] + t * vx), ey = comy - (cand[1] + t * vy);
        dist = sqrt(ex * ex + ey * ey);
    }
    if (dist > p->stability_margin) return (dist - p->stability_margin) * (dist - p->stability_margin);
    return 0.0;
}

/* candidate construction (:188-196) + _compute_tamols_score (:261-373) for window cell (r, cidx) */
static double score_cell(const ctx_t *c, int r, int cidx, double cand[3]) {
    const tamols_params *p = c->p;
    const tamols_leg_job *j = c->j;
    const int R = j->r0 + r, C = j->c0 + cidx;
    if (c->m->xy) {
        const double *q = c->m->xy + (((size_t)j->map * c->m->rows + R) * c->m->cols + C) * 2;
        cand[0] = q[0];
        cand[1] = q[1];
    } else {
        cand[0] = (j->origin[0] + R * j->row_axis[0]) + C * j->col_axis[0];
        cand[1] = (j->origin[1] + R * j->row_axis[1]) + C * j->col_axis[1];
    }
    cand[2] = get_height(c, cand[0], cand[1]) + p->foot_clearance;
    if (!kinematic_feasible(c, cand)) return INFINITY;
    if (leg_collision(c, cand)) return INFINITY;
    const double edge = edge_cost(c, cand) * p->weight_edge_avoidance;
    const double rough = roughness_cost(c, cand) * p->weight_roughness;
    const double e0 = cand[0] - j->seed[0], e1 = cand[1] - j->seed[1], e2 = cand[2] - j->seed[2];
    const double dev = ((e0 * e0 + e1 * e1) + e2 * e2) * p->weight_deviation;
    const double nom = nominal_kinematic_cost(c, cand) * p->weight_nominal_kinematic;
    const double trk = reference_tracking_cost(c, cand) * p->weight_reference_tracking;
    const double stab = stability_cost(c, cand) * p->weight_stability;
    return (((((0.0 + edge) + rough) + dev) + nom) + trk) + stab;
}

/* One leg: the loop at :188-206 and the result packing at :209-222.  scores may be NULL. */
static void run_job(const tamols_params *p, const maps_t *m, const tamols_leg_job *j, tamols_result *out,
                    double *scores) {
    ctx_t c;
    ctx_init(&c, p, j, m);
    double best = INFINITY, bxyz[3] = {0, 0, 0};
    int bi = -1, nfin = 0;
    for (int r = 0; r < j->nr; ++r)
        for (int q = 0; q < j->nc; ++q) {
            double cand[3];
            const double s = score_cell(&c, r, q, cand);
            if (scores) scores[(size_t)r * j->nc + q] = s;
            if (isfinite(s)) ++nfin;
            if (s < best) {
                best = s;
                bi = r * j->nc + q;
                bxyz[0] = cand[0];
                bxyz[1] = cand[1];
                bxyz[2] = cand[2];
            }
        }
    out->best_index = bi;
    out->num_finite = nfin;
    out->best_score = best;
    for (int k = 0; k < 3; ++k) {
        out->best_xyz[k] = bxyz[k];
        out->vertex1[k] = bxyz[k];
        out->vertex2[k] = bxyz[k];
    }
    out->vertex1[0] -= p->constraint_box_dx;
    out->vertex1[1] -= p->constraint_box_dy;
    out->vertex2[0] += p->constraint_box_dx;
    out->vertex2[1] += p->constraint_box_dy;
}

/* n_jobs legs; OpenMP over jobs when built with -fopenmp (the timed CPU baseline). */
int tamols_oracle_score(const tamols_params *p, const double *heights, int32_t num_maps, int32_t rows, int32_t cols,
                        int32_t pitch, const double *xy, const tamols_leg_job *jobs, int32_t n_jobs,
                        tamols_result *results, double *scores, int64_t scores_stride) {
    (void)num_maps;
    maps_t m = {heights, xy, rows, cols, pitch};
#pragma omp parallel for schedule(dynamic, 1)
    for (int k = 0; k < n_jobs; ++k)
        run_job(p, &m, &jobs[k], &results[k], scores ? scores + (size_t)k * scores_stride : NULL);
    return 0;
}

int tamols_oracle_max_threads(void) {
#ifdef _OPENMP
    extern int omp_get_max_threads(void);
    return omp_get_max_threads();
#else
    return 1;
#endif
}
