/* madpose_b200 C ABI -- the drop-in boundary of the B200-native hybrid LO-RANSAC path.
 *
 * The reference (kocurvik/madpose) has no C ABI: its only plugin interface is the
 * pybind11 module built from src/bindings.cpp.  This header is what our own
 * bindings.cpp (madpose_b200/csrc/bindings.cpp, same Python names and fields as
 * /root/reference/src/bindings.cpp:78-174) binds.  Plain pointers and sizes, no
 * torch / Eigen / STL types, caller-owned buffers, int status returns
 * (0 = ok, negative = error, text via mp_last_error()).  Every entry point runs
 * on the GPU; there is no CPU fallback (MP_ERR_NO_DEVICE if no sm_100 device).
 */
#ifndef MADPOSE_B200_H_
#define MADPOSE_B200_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { MP_OK = 0, MP_ERR_ARG = -1, MP_ERR_NO_DEVICE = -2, MP_ERR_CUDA = -3, MP_ERR_ALLOC = -4 };

/* estimator kind: which of the three Python entry points the call serves
 * (bindings.cpp:169-174) */
enum { MP_KIND_CALIBRATED = 0, MP_KIND_SHARED_FOCAL = 1, MP_KIND_TWO_FOCAL = 2,
       /* HybridEstimatePoseAndScale (bindings.cpp:167-168, hybrid_pose_estimator.cpp:37-63): the calibrated estimator
        * without offsets; min_depth is ignored, the returned offsets are 0 */
       MP_KIND_SCALE_ONLY = 3 };

/* madpose.HybridLORansacOptions: ExtendedHybridLORansacOptions
 * (src/hybrid_ransac.h:17-25) over ransac_lib::HybridLORansacOptions
 * (ext/RansacLib/RansacLib/hybrid_ransac.h:47-93); defaults are those. */
typedef struct mp_options {
    uint32_t min_num_iterations;            /* 100   */
    uint32_t max_num_iterations;            /* 10000 */
    uint32_t max_num_iterations_per_solver; /* 10000 */
    uint32_t random_seed;                   /* 0     */
    double success_probability;             /* 0.9999 */
    double squared_inlier_thresholds[2];    /* [reprojection px^2, sampson px^2]; caller must set */
    double data_type_weights[2];            /* caller must set */
    double threshold_multiplier;            /* sqrt(2) */
    int32_t num_lo_steps;                   /* 10 */
    int32_t num_lsq_iterations;             /* 4  */
    int32_t min_sample_multiplicator;       /* 7  */
    int32_t non_min_sample_multiplier;      /* 3  */
    uint32_t lo_starting_iterations;        /* 50 */
    int32_t final_least_squares;            /* 0  */
    int32_t use_ours;                       /* 0; 1: the alternative depth-aware minimal solvers (solver.cpp:536-680,818-984,1045-1153) */
    int32_t use_4p4d;                       /* 0; 1: two-focal estimator only, the 4p4d solver (solver.cpp:1287-1406); use_ours wins */
} mp_options;

/* madpose.EstimatorConfig (src/estimator_config.h:5-33) */
typedef struct mp_config {
    int32_t solver_type; /* 0 HYBRID, 1 EPI_ONLY, 2 MD_ONLY */
    int32_t score_type;
    int32_t lo_type;
    int32_t min_depth_constraint; /* 1 */
    int32_t use_shift;            /* 1 */
    int32_t ceres_use_nonmonotonic_steps; /* 1 */
    int32_t ceres_num_threads;    /* ignored: the LM runs on the device */
    int32_t reserved_;
    double ceres_function_tolerance;  /* 1e-6 */
    double ceres_gradient_tolerance;  /* 1e-8 */
    double ceres_parameter_tolerance; /* 1e-6 */
    double ceres_max_num_iterations;  /* 25 */
} mp_config;

/* One image pair as the Python call receives it (bindings.cpp:169-174):
 * x0,x1 = n x 2 row-major pixel coordinates, d0,d1 = n prior depths.
 * Calibrated uses K0,K1 (row-major 3x3); the focal kinds use pp0,pp1. */
typedef struct mp_pair {
    int32_t n;
    int32_t reserved_;
    const double *x0, *x1, *d0, *d1; /* HOST pointers */
    double min_depth[2];
    double K0[9], K1[9];
    double pp0[2], pp1[2];
} mp_pair;

/* A model = PoseScaleOffset / ...SharedFocal / ...TwoFocal (src/pose.h:7-56):
 * R row-major (9), t (3), scale, offset0, offset1, focal0, focal1
 * (calibrated: focals = 1; shared focal: focal0 == focal1 == focal). */
#define MP_MODEL_DOUBLES 17
typedef struct mp_model { double R[9], t[3], scale, offset0, offset1, focal0, focal1; } mp_model;

/* ransac_lib::HybridRansacStatistics (ext/RansacLib/RansacLib/hybrid_ransac.h:95-104).
 * inlier_indices[t] are returned through the caller's `inliers` buffer. */
typedef struct mp_stats {
    uint32_t num_iterations_total;
    uint32_t num_iterations_per_solver[2];
    int32_t best_num_inliers;
    int32_t best_solver_type;
    int32_t number_lo_iterations;
    int32_t num_inliers[3]; /* lengths of the three inlier lists */
    int32_t reserved_;
    double best_model_score;
    double inlier_ratios[3];
    /* extra counters (not part of the reference struct) */
    uint64_t hypotheses_scored; /* models passed through scoring (minimal + LO) */
    uint64_t residual_evals;    /* sum over scored models of n */
    uint32_t lm_solves, lm_iterations;
    uint64_t evals_fp64;     /* exact fp64 (model x match) evaluations: re-scoring, GetInliers, LO scoring */
    uint64_t lm_block_evals; /* residual blocks evaluated (with Jacobian) over all LM iterations */
    uint64_t evals_fp32;     /* (model x match) evaluations the fused fp32 scoring kernel performed for this pair, including
                              * the speculated iterations that were discarded after a local optimisation */
} mp_stats;

typedef struct mp_ctx mp_ctx;

const char *mp_last_error(void);
int mp_create(int device, mp_ctx **out);
void mp_destroy(mp_ctx *ctx);
/* sm count, max SM clock (kHz), measured register-only FFMA / DFMA peaks (TFLOP/s) */
int mp_device_info(mp_ctx *ctx, int *sm_count, int *sm_clock_khz);
int mp_measure_alu_peaks(mp_ctx *ctx, double *fp32_tflops, double *fp64_tflops);

/* ---- whole path: replaces HybridEstimatePoseScaleOffset{,SharedFocal,TwoFocal}
 * (hybrid_pose_estimator.cpp:8-35, ...shared_focal....cpp:8-51, ...two_focal....cpp:34-75)
 * for a batch of pairs.  out_inliers: caller buffer of sum_p 3*n_p int32; pair p's
 * list for data type t starts at 3*(n_0+...+n_{p-1}) + t*n_p. */
int mp_estimate_batch(mp_ctx *ctx, int kind, int n_pairs, const mp_pair *pairs, const mp_options *opt,
                      const mp_config *cfg, mp_model *out_models, mp_stats *out_stats, int32_t *out_inliers);
/* The same in two steps, for callers that keep a batch resident in HBM: mp_upload_batch packs
 * + copies the pairs and runs the front-end; mp_estimate_resident runs the loop on the resident
 * batch (any number of times, e.g. with different options) and copies the results back. */
int mp_upload_batch(mp_ctx *ctx, int kind, int n_pairs, const mp_pair *pairs, const mp_options *opt);
int mp_estimate_resident(mp_ctx *ctx, const mp_options *opt, const mp_config *cfg, mp_model *out_models,
                         mp_stats *out_stats, int32_t *out_inliers);
/* ---- several GPUs of one node inside the library.  Pairs are independent: the batch is dealt to the
 * devices by decreasing match count (snake order), every device runs its share through mp_estimate_batch on its own
 * context and host thread, results are scattered back in the caller's order (same buffers / layout as
 * mp_estimate_batch).  No inter-GPU traffic.  Contexts are created on first use and kept for the process lifetime.
 * A pair's result does not depend on which device ran it or on n_devices. */
int mp_estimate_batch_multi(int n_devices, const int *devices, int kind, int n_pairs, const mp_pair *pairs,
                            const mp_options *opt, const mp_config *cfg, mp_model *out_models, mp_stats *out_stats,
                            int32_t *out_inliers);
/* The same for a batch that mixes estimator kinds (BASELINE.json configs[4]): kinds[p] in 0..3, opt_by_kind[4] = the
 * options of each kind (entries of absent kinds are not read).  The kinds run concurrently on every device. */
int mp_estimate_mixed_multi(int n_devices, const int *devices, int n_pairs, const int *kinds, const mp_pair *pairs,
                            const mp_options *opt_by_kind, const mp_config *cfg, mp_model *out_models,
                            mp_stats *out_stats, int32_t *out_inliers);
int mp_device_count(void); /* CUDA devices visible to the process (0 without a driver) */
/* per kernel class of the last estimate call: device ms (CUDA events) and launch counts for
 * {sampling, minimal solvers, fp32 scoring, decide (fp64 re-score + inliers + LO/LM)}; rounds */
int mp_last_breakdown(mp_ctx *ctx, double *class_ms4, int *class_launches4, int *rounds);
/* The loop runs the batch as `groups` groups of consecutive pairs, each with its own round pipeline (sampling, minimal
 * solvers, scoring, scan) on its own CUDA stream; local optimisations run on side streams and never hold a round up.
 * 0 = automatic (about 750 pairs per group, at most 8 groups); 1 = everything on one stream (the per-class kernel
 * timings of mp_last_breakdown are exact only then). */
int mp_set_groups(mp_ctx *ctx, int groups);
/* Upper bound of the RANSAC iterations speculated per pair per round (1..256; 0 = default 128).  A pair starts with 64
 * and doubles its batch after every round that found no new best model.  Results do not depend on it. */
int mp_set_batch(mp_ctx *ctx, int iterations);
/* Which form of the point-only minimal solvers that have both runs (6-point shared focal today): 0 = automatic (one
 * warp per sample), 1 = one thread per sample, 2 = one warp per sample.  The two forms agree to rounding (each within
 * the parity tolerances against the oracle); for a given mode results never depend on the schedule. */
int mp_set_minimal_mode(mp_ctx *ctx, int mode);
/* Experiments: in automatic mode the point-solver samples of RANSAC iterations >= `iteration` take the warp-per-sample
 * form and the earlier ones the thread-per-sample form (default 0: all of them the warp form).  -1 restores the default.  A function of the iteration index only, so results
 * do not depend on the schedule; different values give results that agree to rounding. */
int mp_set_warp_from_iteration(mp_ctx *ctx, int iteration);
/* Parity forensics: with events_per_pair > 0 the next estimate calls record, per pair and in chronological order, one
 * {iteration, score} record for every iteration whose best hypothesis was re-scored in fp64 and one {-1, score} record
 * for every ScoreModel inside LocalOptimization / the final least squares (hybrid_ransac.h:383-482, 186-203).
 * mp_read_trace copies pair `pair`'s records (2 doubles each) of the last call; *n_events = events the pair produced
 * (those beyond events_per_pair were dropped).  Results never depend on tracing. */
int mp_set_trace(mp_ctx *ctx, int events_per_pair);
int mp_read_trace(mp_ctx *ctx, int pair, double *out, int cap_events, int *n_events);
/* profiling aid: SM-clock cycles (one thread per pair's CTA, summed over pairs) the decide kernel spent in
 * {sequential scan, GetInliers, fp64 ScoreModel, LM evaluation, LM solves total, shuffles, un-flatten, LO total} */
int mp_last_phase_cycles(mp_ctx *ctx, unsigned long long *cycles8);
/* device time (ms, CUDA events on the launching stream) of the kernels of the last
 * mp_estimate_batch call, and the number of kernel launches it made */
int mp_last_timing(mp_ctx *ctx, double *kernel_ms, double *total_ms, int *launches);

/* ---- per-subsystem entry points (same kernels' device code, used by the parity tests) */
/* ScoreModel (hybrid_ransac.h:265-282) over many models of one pair.
 * scores_f32: fused fp32 scoring kernel; scores_f64: exact fp64 evaluation. Either may be NULL. */
int mp_score_batch(mp_ctx *ctx, int kind, const mp_pair *pair, const mp_options *opt, const mp_config *cfg,
                   int n_models, const mp_model *models, float *scores_f32, double *scores_f64);
/* The fp64 leg of mp_score_batch WITHOUT any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  The scratch the scoring path needs (the pair's constants and fp64 match records) lives in one
 * fixed file-scope device array of the library (~770 KB per GPU, held until the process exits), so a pair may have at
 * most MP_NOALLOC_MAX_MATCHES matches: larger pairs are rejected with MP_ERR_ARG (use mp_score_batch).
 * pair_dev: an mp_pair whose x0, x1, d0, d1 are DEVICE pointers on `device`; d_models (n_models) and d_scores64
 * (n_models, output) are DEVICE buffers of the caller; stream: a cudaStream_t of `device` (NULL = default stream).
 * The call only enqueues work on `stream` and returns: synchronise before reading d_scores64.  It calls no
 * cudaMalloc / cudaFree / allocating routine.  All calls on a device share the scratch array, which keeps the last
 * call's contents (the call clears and rewrites what it uses): calls that overlap on the GPU race -- keep them on one
 * stream.  Scores are bit-identical to mp_score_batch's scores_f64. */
#define MP_NOALLOC_MAX_MATCHES 16384
int mp_score_models_noalloc(int device, int kind, const mp_pair *pair_dev, const mp_options *opt, const mp_config *cfg,
                            int n_models, const mp_model *d_models, double *d_scores64, void *stream);
/* The step before the path: madpose.utils.get_depths (madpose/utils.py:4-22) -- nearest-neighbour
 * lookup of the prior depth of n keypoints in a depth map of another resolution: scale the keypoints by
 * (map_w / img_w, map_h / img_h), round half to even (numpy.round), clip to the map, gather.
 * depth_map: HOST buffer, row-major map_h x map_w, elem_size 4 (float) or 8 (double); mkpts: n x 2 doubles (x, y);
 * out: n doubles.  Keeps the gather on the device for callers whose depth maps come from a network. */
int mp_get_depths(mp_ctx *ctx, const void *depth_map, int elem_size, int map_h, int map_w, int img_h, int img_w,
                  const double *mkpts, int n, double *out);
/* The step after the path: madpose.utils.compute_pose_error (madpose/utils.py:59-78, with
 * angle_error_vec :52-56 and angle_error_mat :46-49) for n estimated poses against n ground truths.
 * T_0to1: n x 16 (row-major 4x4), R: n x 9 (row-major), t: n x 3; t_thres < 0 means None (utils.py:66).
 * err_t / err_R: n doubles each, in degrees (err_t folded to [0, 90]: the sign ambiguity of E). */
int mp_pose_errors(mp_ctx *ctx, int n, const double *T_0to1, const double *R, const double *t, double t_thres,
                   double *err_t, double *err_R);
/* GetInliers (hybrid_ransac.h:289-349): ascending indices with r^2 < thr[t] (thr in
 * the front-end's rescaled units, 3 entries); out_idx has 3*n entries (type t at t*n). */
int mp_inliers(mp_ctx *ctx, int kind, const mp_pair *pair, const mp_options *opt, const mp_config *cfg,
               const mp_model *model, const double *thr3, int32_t *out_idx, int32_t *out_counts3);
/* MinimalSolver (hybrid_pose_estimator.cpp:65-187 etc.) on explicit samples:
 * sample_idx = n_samples x k (k = 3/4/4 for solver 0, 5/6/7 for solver 1);
 * out_models = n_samples x max_models (4/8/4 or 10/15/3); models are in the
 * front-end's internal (pp-subtracted, normalised) units. */
int mp_minimal_solve_batch(mp_ctx *ctx, int kind, int solver_idx, const mp_pair *pair, const mp_options *opt,
                           const mp_config *cfg, int n_samples, const int32_t *sample_idx, mp_model *out_models,
                           int32_t *out_counts);
/* The stand-alone solver functions of the Python module (bindings.cpp:157-166) on n_samples explicit samples:
 * solve_scale_and_shift{,_shared_focal,_two_focal} (solver.cpp:35-125, 127-293, 295-480) -> out_sols, and
 * solve_scale_shift_pose{,_shared_focal,_two_focal} (solver.cpp:482-534, 682-739, 986-1043, scale_on_x = false)
 * -> out_models.  kind 0: 3 points, kinds 1/2: 4 points (NP).  x_homo / y_homo: per sample a 3 x NP row-major matrix
 * whose COLUMNS are homogeneous points (the reference's Eigen::Matrix3d / Matrix3x4d as numpy passes it);
 * depth_x / depth_y: per sample NP depths.  out_sols: per sample 8 rows of 6 doubles (1, b1, a2, b2*a2, f1, f2)
 * (calibrated: f = 1; shared focal: f1 == f2), out_nsols rows valid.  out_models: per sample 8 models with
 * scale = a2, offset0 = b1, offset1 = b2*a2 (the wrappers' convention, no min-depth filter), out_nmodels valid.
 * Any of the four outputs may be NULL. */
int mp_solve_affine(mp_ctx *ctx, int kind, int n_samples, const double *x_homo, const double *y_homo,
                    const double *depth_x, const double *depth_y, double *out_sols, int32_t *out_nsols,
                    mp_model *out_models, int32_t *out_nmodels);
/* estimate_scale_and_pose (solver.cpp:5-33, bindings.cpp:156): weighted Umeyama Y ~ scale R X + t.
 * X, Y: 3 x n row-major (columns = points), W: n weights.  out: R, t, scale (offsets 0, focals 1). */
int mp_estimate_scale_and_pose(mp_ctx *ctx, int n, const double *X, const double *Y, const double *W, mp_model *out);
/* NonMinimalSolver / LeastSquares (optimizer.h:9-131,247-499): one LM problem per
 * entry; problem j uses indices idx[off[3j+t] .. off[3j+t+1]) for data type t.
 * out_info: per problem {iterations, termination, initial_cost, final_cost} as doubles.
 * nonminimal: 1 = the call is NonMinimalSolver, 0 = LeastSquares; they differ only in how EstimatorConfig.use_shift
 * reaches the optimiser (calibrated: both; shared focal: NonMinimalSolver only,
 * hybrid_pose_shared_focal_estimator.cpp:149; two focal: never). */
int mp_lm_batch(mp_ctx *ctx, int kind, const mp_pair *pair, const mp_options *opt, const mp_config *cfg,
                int n_problems, const mp_model *start, const int32_t *idx, const int32_t *off, mp_model *out_models,
                double *out_info, int nonminimal);

#ifdef __cplusplus
}
#endif
#endif
