/*
 * include/tamols.h -- C ABI of the B200 (sm_100a) TAMOLS foothold-scoring hot path.
 *
 * The reference (Magicyw/Quadruped-PyMPC-TAMOLS) has no FFI layer: its boundary for this path is the Python class
 * `VisualFootholdAdaptation` with adaptation_strategy='tamols'
 * (quadruped_pympc/helpers/visual_foothold_adaptation.py:38-72, 86-99, 153-231), whose candidate loop (:188-206) is a
 * Python `for` over every heightmap cell with scipy cKDTree nearest-neighbour height queries (:21-35).
 * The entry points below replace that loop; the Python drop-in class in
 * quadruped_pympc_tamols_b200/helpers/visual_foothold_adaptation.py binds them with ctypes (INTEGRATION.md shows the
 * stub a reference maintainer would add).
 *
 * Data model.  A height map is a regular planar grid: the world xy of cell (r, c) is
 *     origin + r * row_axis + c * col_axis          (orthogonal axes; the grid may be rotated by the sensor yaw)
 * and its height is heights[map][r][c].  This is the layout of gym_quadruped's HeightMap.data[r, c, 0, :] = (x, y, z)
 * as the reference consumes it (:27-28).  The reference's KD-tree nearest-neighbour query is replaced by the analytic
 * nearest cell of that grid (round to nearest in grid coordinates, clamped to the candidate window), which is the same
 * cell except for queries exactly equidistant from two cells (documented tie rule).
 * One "leg job" = one leg of one robot: a candidate window of one map, the seed foothold, the hip, and the robot state
 * the cost terms need.  Every cell of the window is a candidate (:240-243); the winner is the FIRST minimum in
 * row-major order (:203-206).
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success and a negative TAMOLS_E_* code
 * otherwise, with a message from tamols_last_error().  Pointers named d_* are DEVICE pointers, all others HOST.
 * All arithmetic is FP64 like the reference (numpy defaults); heights may be stored as f64 or f32.
 */
#ifndef TAMOLS_H
#define TAMOLS_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { TAMOLS_OK = 0, TAMOLS_E_ARG = -1, TAMOLS_E_CUDA = -2, TAMOLS_E_NODEVICE = -3 };
enum { TAMOLS_F64 = 0, TAMOLS_F32 = 1 };
enum {
    TAMOLS_FLAG_HAS_VELOCITY = 1, /* forward_vel is not None (:388, :567) */
    TAMOLS_FLAG_STABILITY = 2     /* base_position and current_feet_pos given and the leg is in swing (:631-658) */
};

/* cfg.simulation_params['tamols_params'] as read with .get(key, default) on every call (:298-305, :377-378, :438-439,
 * :537, :636, :671, :214-215) plus the literals of the scoring code. */
typedef struct tamols_params {
    double gradient_delta;            /* 'gradient_delta' (0.04) :438,:484 */
    double slope_threshold;           /* 'slope_threshold' (0.7) :439 */
    double weight_edge_avoidance;     /* default 15.0 :298 (config.py:217 sets 10.0) */
    double weight_roughness;          /* :299 */
    double weight_deviation;          /* :300 */
    double weight_nominal_kinematic;  /* :302 */
    double weight_reference_tracking; /* :303 */
    double weight_stability;          /* :305 */
    double stability_margin;          /* :636 */
    double estimated_swing_time;      /* :671 */
    double h_des;                     /* :537 */
    double l_min, l_max;              /* tamols_params['l_min'|'l_max'][robot] :377-378 */
    double stance_duration;           /* literal 0.3 :387 */
    double sensor_offset;             /* literal 0.02 added by FastHeightMap.get_height :35 */
    double foot_clearance;            /* literal 0.005 :192 */
    double collision_margin;          /* literal 0.02 :417 */
    double collision_alphas[5];       /* np.linspace(0.2, 0.8, 5) :402 (host computes them so the doubles are numpy's) */
    double constraint_box_dx;         /* :214 */
    double constraint_box_dy;         /* :215 */
} tamols_params;

/* One leg of one robot (the body of `for leg_id, leg_name in enumerate(legs_order)`, :164-228). */
typedef struct tamols_leg_job {
    int32_t map;             /* which height map */
    int32_t r0, c0, nr, nc;  /* candidate window = the per-leg heightmap the reference receives */
    int32_t flags;           /* TAMOLS_FLAG_* */
    double origin[2];        /* xy of cell (0,0) of the map */
    double row_axis[2];      /* xy step between rows */
    double col_axis[2];      /* xy step between columns */
    double seed[3];          /* reference_footholds[leg] (:165) */
    double hip[3];           /* hip_positions[leg] (:166) */
    double diag_foot[2];     /* current_feet_pos[diagonal leg][:2] (:640-664) */
    double base_xy[2];       /* base_position[:2] (:675) */
    double vel[3];           /* forward_vel (:154) */
} tamols_leg_job;

typedef struct tamols_result {
    int32_t best_index;  /* row-major index inside the window, -1 when no candidate is finite (fallback branch :223-228) */
    int32_t num_finite;  /* number of candidates with a finite score */
    double best_score;
    double best_xyz[3];  /* best_candidate (:206) */
    double vertex1[3];   /* footholds_constraints[leg][0] (:216-218) */
    double vertex2[3];   /* footholds_constraints[leg][1] (:219-221) */
} tamols_result;

typedef struct tamols_handle tamols_handle;

/* VisualFootholdAdaptation.__init__ for 'tamols' (:55-57).  max_jobs bounds the jobs per call. */
int tamols_create(const tamols_params *params, int32_t device, int32_t max_jobs, tamols_handle **out);
int tamols_destroy(tamols_handle *h);
/* self.tamols_params is a live dict upstream; this replaces the handle's copy. */
int tamols_set_params(tamols_handle *h, const tamols_params *params);
const char *tamols_last_error(void);
void *tamols_stream(tamols_handle *h);
int64_t tamols_launch_count(const tamols_handle *h);
/* Which kernel the last scoring launch of this handle ran: "window" (k_tamols_window, one block per job: batched
 * launches of at least 4 jobs per SM), "tile32" / "tile8" (k_tamols_score, a job's window split into 32- or 8-row tiles
 * over many SMs: the latency regime), "" before the first launch.  Measurement hook (bench.py, tests); TAMOLS_KERNEL =
 * tile | window in the environment overrides the choice when the handle is created.  Same result bits either way. */
const char *tamols_last_kernel(const tamols_handle *h);

/*
 * The candidate loop of compute_adaptation (:164-228) for n_jobs legs in ONE launch, device buffers in and out,
 * asynchronous on tamols_stream(h).
 *   d_heights  [num_maps, rows, pitch] of height_dtype; pitch >= cols elements per row.  When pitch * sizeof(elem) is
 *              a multiple of 16 and the base is 16-byte aligned the tile + halo is staged to shared memory by TMA,
 *              otherwise by plain loads.
 *   d_xy       optional [num_maps, rows, cols, 2] f64 candidate coordinates as stored by the sensor (:243); NULL =
 *              computed from the job's grid.
 *   max_nr/nc  upper bounds of the jobs' nr / nc (the jobs live on the device; the bounds size the launch grid)
 *   d_jobs     [n_jobs]; d_results [n_jobs]; d_scores optional [n_jobs, scores_stride] per-candidate scores
 *              (_compute_tamols_score, :261-373; +inf for rejected candidates), NULL to skip.
 */
int tamols_score_device(tamols_handle *h, const void *d_heights, int32_t height_dtype, int32_t num_maps, int32_t rows,
                        int32_t cols, int32_t pitch, const double *d_xy, const tamols_leg_job *d_jobs, int32_t n_jobs,
                        int32_t max_nr, int32_t max_nc, tamols_result *d_results, double *d_scores,
                        int64_t scores_stride);

/*
 * tamols_score_device for callers that cannot tolerate a device allocation inside the call (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments, same kernel and bit-identical results; the only difference is
 * where the kernel's scratch lives.  tamols_score_device keeps the per-tile partial results and the arrival tickets
 * in buffers of the handle and grows the partial buffer with cudaFree + cudaMalloc when a launch needs more; this
 * entry point uses a file-scope `static __device__` array of the library instead (tickets + partials,
 * TAMOLS_STATIC_SCRATCH_BYTES per GPU that loads the library, held until the process exits) and calls no cudaMalloc,
 * cudaFree or other allocating routine.  Limits: n_jobs <= TAMOLS_STATIC_MAX_JOBS and n_jobs * tiles per job <=
 * TAMOLS_STATIC_MAX_PARTIALS (tiles per job = ceil(max_nr / th) * ceil(max_nc / 32), th = 8 or 32 as chosen by the
 * library; always safe: n_jobs * ceil(max_nr / 8) * ceil(max_nc / 32)); anything larger returns TAMOLS_E_ARG with
 * the numbers in tamols_last_error().  The array is shared by EVERY call on that GPU, whatever the handle or stream:
 * the tickets are cleared on the handle's stream before each launch (a memset node, no allocation), the partials are
 * written before they are read, and two such calls that overlap on the GPU (different streams or handles, no
 * ordering between them) race on it -- serialise them.  The handle's own buffers are not touched.
 */
#define TAMOLS_STATIC_MAX_JOBS 4096
#define TAMOLS_STATIC_MAX_PARTIALS 65536
#define TAMOLS_STATIC_SCRATCH_BYTES (((TAMOLS_STATIC_MAX_JOBS + 1) * 4 + 255) / 256 * 256 + TAMOLS_STATIC_MAX_PARTIALS * 16)
int tamols_score_device_static(tamols_handle *h, const void *d_heights, int32_t height_dtype, int32_t num_maps,
                               int32_t rows, int32_t cols, int32_t pitch, const double *d_xy,
                               const tamols_leg_job *d_jobs, int32_t n_jobs, int32_t max_nr, int32_t max_nc,
                               tamols_result *d_results, double *d_scores, int64_t scores_stride);

/* Same with HOST buffers (what the Python class calls): copies heights (f64, [num_maps, rows, cols] dense), optional
 * xy and the jobs to the device, runs the launch, copies results (and scores when not NULL) back, synchronises. */
int tamols_score_host(tamols_handle *h, const double *heights, int32_t num_maps, int32_t rows, int32_t cols,
                      const double *xy, const tamols_leg_job *jobs, int32_t n_jobs, tamols_result *results,
                      double *scores, int64_t scores_stride);

/* Host helper (no GPU work): the regular-grid description of one sensor map's cell coordinates xy[rows, cols, 2]
 * (gym_quadruped HeightMap.data[:, :, 0, :2]): origin = cell (0,0), row_axis / col_axis = mean step between rows /
 * columns.  TAMOLS_E_ARG when the cells are not a regular orthogonal grid to 1e-6 of the cell size -- the analytic
 * nearest-cell lookup that replaces the reference's cKDTree (visual_foothold_adaptation.py:23-35) needs one. */
int tamols_fit_grid(const double *xy, int32_t rows, int32_t cols, double origin[2], double row_axis[2],
                    double col_axis[2]);

/* One leg as VisualFootholdAdaptation.compute_adaptation receives it (visual_foothold_adaptation.py:86-166): the
 * sensor object's cell array plus the per-leg vectors.  `data` = HeightMap.data, float64 [rows, cols, 1, 3] contiguous
 * (x, y, z of every cell). */
typedef struct tamols_sensor_leg {
    const double *data;
    int32_t rows, cols;
    int32_t flags;       /* TAMOLS_FLAG_* */
    int32_t reserved;
    double seed[3];      /* reference_footholds[leg] */
    double hip[3];       /* hip_positions[leg] */
    double diag_foot[2]; /* current_feet_pos[diagonal leg][:2] (TAMOLS_FLAG_STABILITY) */
    double base_xy[2];   /* base_position[:2]                  (TAMOLS_FLAG_STABILITY) */
    double vel[3];       /* forward_vel                        (TAMOLS_FLAG_HAS_VELOCITY) */
} tamols_sensor_leg;

/* The whole candidate loop of compute_adaptation for n_legs legs straight from the sensor arrays: splits every
 * data array into heights and cell coordinates, derives and checks each leg's grid (tamols_fit_grid), assembles
 * the job records and runs tamols_score_host -- the host-side packing in one C call instead of ~100 us of numpy. */
int tamols_score_sensors(tamols_handle *h, const tamols_sensor_leg *legs, int32_t n_legs, tamols_result *results);

/* ---- the step in front of the scoring, for a batch of robots on the device.
 * (a) FootholdReferenceGenerator.compute_footholds_reference (helpers/foothold_reference_generator.py:53-157): hip
 *     positions in the yaw-aligned frame, +-hip_offset, desired-velocity offset stance_time/2 * v_ref clipped to
 *     +-1.5 hip_height, velocity-error compensation sqrt(h/g) (mean of the last `hist_len` velocities - v_ref) clipped
 *     to +-0.05, back to the world frame, manual CoM offset, z = lift-off height.  The moving-average deque is the
 *     ring d_vel_hist / d_hist_count (state, advanced by every call).
 * (b) when d_jobs != NULL: the 4 R tamols_leg_job records of those seeds on ONE global regular height map -- the
 *     candidate window of a leg is win_rows x win_cols cells centred on the cell nearest to its seed (clamped to the
 *     map), replacing the per-leg HeightMap.update_height_map ray-cast grids of wb_interface.py:233-234; flags, hip,
 *     diagonal foot, base and velocity exactly as VisualFootholdAdaptation.compute_adaptation assembles them.
 * All pointers are DEVICE pointers with num_robots leading rows; legs in FL, FR, RL, RR order; float64. */
typedef struct tamols_seed_batch {
    int32_t num_robots;
    int32_t hist_len;                /* vel_moving_average_length (20) */
    const double *d_base_position;   /* [R,3] */
    const double *d_base_euler;      /* [R,3] roll, pitch, yaw */
    const double *d_base_lin_vel;    /* [R,3] world frame */
    const double *d_ref_lin_vel;     /* [R,3] world frame (x, y used) */
    const double *d_hips;            /* [R,4,3] hip positions, world frame */
    const double *d_feet;            /* [R,4,3] current feet positions (job building only), or NULL */
    const uint8_t *d_contact;        /* [R,4] 1 = stance (job building only), or NULL = all swing */
    const double *d_com_height;      /* [R] com_height_nominal */
    const double *d_stance_time;     /* [R] */
    const double *d_lift_off_z;      /* [R,4] lift_off_positions[leg][2] */
    const double *d_com_offset_b;    /* [R,3] com_pos_offset_b, or NULL (zeros) */
    double *d_vel_hist;              /* [R,hist_len,2] state */
    int32_t *d_hist_count;           /* [R] state: velocities pushed so far */
    double hip_height, hip_offset, gravity;
    double *d_ref_feet;              /* out [R,4,3] */
    tamols_leg_job *d_jobs;          /* out [4 R] or NULL */
    int32_t map_rows, map_cols, win_rows, win_cols;
    double map_origin[2], map_row_axis[2], map_col_axis[2];
} tamols_seed_batch;

int tamols_seed_jobs(tamols_handle *h, const tamols_seed_batch *b);

#ifdef __cplusplus
}
#endif
#endif /* TAMOLS_H */
