/*
 * ba_engine.h — C-ABI of the B200 bundle-adjustment engine.
 *
 * This is the drop-in boundary for the reference's src/adjuster optimisation path.
 * The reference has no FFI: its "operator API" is the C++ functor
 *     BundleAdjuster::BundleAdjuster()                       src/adjuster/BundleAdjuster.h:167-174
 *     void BundleAdjuster::operator()(WorldStructure::Ptr&)  src/adjuster/BundleAdjuster.h:176-186
 * which gathers parameters from the map (loadDataFromWorld, :82-98), hands raw double*
 * blocks to the solver (bundleAdjustment, :100-141; ceres::Solve at :126) and scatters
 * the result back (updateWorld, :143-156).  The boundary sits exactly at :126 — the
 * functions below replace that call (plus the SoA form of the gather/scatter either side).
 * The host adapter that keeps the functor's shape lives in
 * 3dreconstruction_b200/host/adjuster/BundleAdjuster.h and calls only this header.
 *
 * Two shared libraries implement this header with identical symbols:
 *   3dreconstruction_b200/csrc/libba_engine.so   sm_100a CUDA engine (the product; no CPU fallback)
 *   oracle/libba_oracle.so                       CPU restatement of the reference's solver (test infrastructure only)
 *
 * Conventions: every function returns 0 on success and a negative
 * ba_status on error; ba_last_error() gives the message; no C++ exception crosses the
 * boundary; host arrays are borrowed for the duration of a call only; a handle is used
 * by one host thread at a time.  All reals are IEEE double (the reference's arithmetic
 * type: cv::Matx..d, BundleAdjuster.h:74-76); all indices are int32.
 */
#ifndef BA_ENGINE_H
#define BA_ENGINE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BA_ABI_VERSION 1

typedef struct ba_engine ba_engine; /* opaque */

typedef enum ba_status {
    BA_OK = 0,
    BA_ERR_INVALID_ARGUMENT = -1,
    BA_ERR_NO_PROBLEM = -2,
    BA_ERR_CUDA = -3,        /* CUDA runtime/driver failure, or no usable device (the engine never falls back to the CPU) */
    BA_ERR_NCCL = -4,
    BA_ERR_UNSUPPORTED = -5, /* option combination this library does not implement */
    BA_ERR_EVAL_FAILED = -6, /* non-finite residual/Jacobian at the initial point (Ceres: "initial evaluation failed") */
    BA_ERR_INTERNAL = -7
} ba_status;

/* Robust kernel applied to s = ||r||^2 of each 2-vector residual.
 * Reference: one shared ceres::HuberLoss(4), BundleAdjuster.h:109.  Cauchy is the north star's second kernel. */
typedef enum ba_robust_kind { BA_ROBUST_NONE = 0, BA_ROBUST_HUBER = 1, BA_ROBUST_CAUCHY = 2 } ba_robust_kind;

/* Residual model. PINHOLE = ReprojectCost (BundleAdjuster.h:38-68): +z forward, intrinsics group = (fx,fy,cx,cy).
 * SNAVELY = SnavelyReprojectionError (SnavelyReprojectionError.h:16-48): -z, intrinsics group = (f,l1,l2). */
typedef enum ba_model { BA_MODEL_PINHOLE = 0, BA_MODEL_SNAVELY = 1 } ba_model;

/* How the reduced camera system is solved.
 * EXACT is what the reference configures (SPARSE_SCHUR + EIGEN_SPARSE, BundleAdjuster.h:172-173): a direct
 * factorisation of S.  The engine forms S densely and runs a blocked Cholesky on the GPU; that needs one GPU,
 * at most 8192 reduced unknowns and, with free intrinsics, one shared Camera group — ba_set_problem fails with
 * BA_ERR_UNSUPPORTED otherwise.  PCG is block-preconditioned conjugate gradients on the implicit Schur complement
 * (what Ceres calls ITERATIVE_SCHUR; the reference even sets preconditioner_type = JACOBI at :171, which only
 * ITERATIVE_SCHUR would read).  AUTO (engine default) takes EXACT when it applies and the reduced system has
 * at most 2048 unknowns (341 poses: the reference's own SequentialActuator regime), PCG otherwise; the oracle
 * reads AUTO as EXACT. */
typedef enum ba_linear_solver { BA_LINEAR_EXACT = 0, BA_LINEAR_PCG = 1, BA_LINEAR_AUTO = 2 } ba_linear_solver;
typedef enum ba_preconditioner { BA_PRECOND_JACOBI = 0 /* blockdiag(Hcc + Dc^2) */, BA_PRECOND_SCHUR_JACOBI = 1 /* blockdiag(S) */ } ba_preconditioner;
/* SCHUR_JACOBI costs one more by-camera pass per linear solve (k_schur_diag) and saves PCG iterations; engine and oracle implement both. */

/* Which rule selects the first-order branch of AngleAxisRotatePoint. */
typedef enum ba_taylor_rule { BA_TAYLOR_THETA2_LE_EPS = 0 /* theta^2 <= DBL_EPSILON */, BA_TAYLOR_THETA_IS_ZERO = 1 /* hypot()==0 */ } ba_taylor_rule;

/* Which minimizer policy drives the outer loop.
 * CERES: trust-region LM exactly as the reference configures it (BundleAdjuster.h:167-174) — the parity target.
 * G2O:   g2o's OptimizationAlgorithmLevenberg + VertexSE3Expmap/EdgeSE3ProjectXYZ semantics, for the north star's literal wording
 *        (the reference finds g2o with CMake but never calls it): lambda0 = 1e-5 max diag(H), uniform damping H + lambda I, no
 *        Jacobi scaling, gain ratio (chi2 - chi2+) / (dx.(lambda dx + b) + 1e-3), accept iff > 0, lambda *= clamp(1-(2rho-1)^3, 1/3, 2/3)
 *        on success and *= nu (nu doubles) on failure, <= 10 trials per iteration, stop after max_num_iterations iterations;
 *        poses move by the SE3 exponential map on the left, T <- exp(omega, upsilon) T; intrinsics are constant. */
typedef enum ba_lm_policy { BA_POLICY_CERES = 0, BA_POLICY_G2O = 1 } ba_lm_policy;

typedef enum ba_termination {
    BA_TERM_CONVERGENCE = 0,    /* function / parameter / gradient tolerance or minimum radius reached */
    BA_TERM_NO_CONVERGENCE = 1, /* iteration cap */
    BA_TERM_FAILURE = 2,        /* too many consecutive invalid steps, or evaluation failure */
    BA_TERM_TARGET_REACHED = 3  /* chi2 <= target_chi2 (bench-only stop rule, off by default) */
} ba_termination;

/* Options. Call ba_options_default() first; the defaults are the reference's configuration:
 * Ceres 2.2 trust-region LM defaults + what BundleAdjuster() sets (BundleAdjuster.h:167-174), see SURVEY §8 (a-LM). */
typedef struct ba_options {
    int32_t abi_version;                 /* BA_ABI_VERSION */
    /* --- problem semantics --- */
    int32_t model;                       /* ba_model */
    int32_t robust_kind;                 /* ba_robust_kind; reference: HUBER */
    double  robust_a;                    /* reference: 4.0 */
    int32_t intrinsics_free;             /* 1: every intrinsics group is a free block (reference, :106-107); 0: held constant (north star's 2x6/2x3 form) */
    int32_t taylor_rule;                 /* ba_taylor_rule */
    /* --- Levenberg-Marquardt (Ceres names) --- */
    int32_t max_num_iterations;          /* 50 */
    double  initial_trust_region_radius; /* 1e4 */
    double  max_trust_region_radius;     /* 1e16 */
    double  min_trust_region_radius;     /* 1e-32 */
    double  min_relative_decrease;       /* 1e-3 */
    double  min_lm_diagonal;             /* 1e-6 */
    double  max_lm_diagonal;             /* 1e32 */
    int32_t max_num_consecutive_invalid_steps; /* 5 */
    double  function_tolerance;          /* 1e-6 */
    double  gradient_tolerance;          /* 1e-10 */
    double  parameter_tolerance;         /* 1e-8 */
    int32_t jacobi_scaling;              /* 1 */
    /* --- reduced-system solve --- */
    int32_t linear_solver;               /* ba_linear_solver; engine default AUTO, oracle default EXACT */
    int32_t preconditioner;              /* ba_preconditioner */
    int32_t pcg_max_iterations;          /* parity mode 2000 */
    double  pcg_rel_tolerance;           /* stop when sqrt(r'M^-1 r / b'M^-1 b) <= tol; parity mode 1e-12 */
    int32_t pcg_forcing;                 /* 0: fixed tolerance above; 1: Ceres inexact-Newton rule (eta, Q-criterion, r_tolerance off) */
    double  eta;                         /* 0.1 */
    /* --- extra stop rule for time-to-converge measurements --- */
    double  target_chi2;                 /* >0: stop as soon as chi2 (= 2*cost) <= target_chi2; 0: off */
    /* --- execution --- */
    int32_t device;                      /* CUDA ordinal (engine) */
    int32_t num_threads;                 /* oracle: OpenMP threads; reference sets 1 (:170) */
    int32_t rank;                        /* multi-GPU: this process' rank, one process per GPU */
    int32_t world_size;                  /* 1 = single GPU */
    const void* nccl_unique_id;          /* world_size>1: 128-byte ncclUniqueId from ba_nccl_unique_id() on rank 0, broadcast by the caller */
    void*   stream;                      /* engine: cudaStream_t to run on (NULL: the engine creates its own) */
    int32_t scatter_mode;                /* engine camera-side scatter of K4/K5: 0 = default (4), 1 = red.global.add.f64, 4 = partial rows + by-camera gather,
                                            one CTA per tile, 6 = the same rows from the persistent cp.async.bulk streaming kernel (experimental opt-in:
                                            bit-identical to 4 on the small shapes, slower per PCG iteration and not reproducible on the Final-13682 shape) */
    int32_t p2p_allreduce;               /* multi-GPU all-reduce of the PCG matvec: 0 = auto (one-shot NVLink peer-memory all-reduce fused into the
                                            K5b gather when every peer buffer can be mapped with CUDA IPC, else NCCL), 1 = NCCL only, 2 = peer memory required */
    int32_t lm_policy;                   /* ba_lm_policy; default CERES */
    int32_t warm_start_radius;           /* 1: ba_solve starts from the final trust-region radius of the previous solve on this handle (incremental use); 0 (default): always initial_trust_region_radius, as the reference (a fresh ceres::Solver per call) */
    int32_t reserved[4];
} ba_options;

/* Problem in canonical indexing: observations sorted by point id (non-decreasing obs_pt),
 * point id = rank of WorldPoint::idx_, pose id = rank of Image::idx_, group id = Camera object. */
typedef struct ba_problem {
    int32_t num_cameras;      /* Nc poses: 6 doubles each (angle-axis, translation) — image_extrinsic_, BundleAdjuster.h:93-94 */
    int32_t num_groups;       /* Ng intrinsics groups — camera_intrinsics_, :95-96 */
    int32_t num_points;       /* Np — world_points_pos_, :84-86 */
    int32_t num_observations; /* No residual blocks, :110-121 */
    const double*  pose;      /* [Nc][6] */
    const double*  intr;      /* [Ng][4] (pinhole: fx,fy,cx,cy; snavely: f,l1,l2,unused) */
    const int32_t* cam_group; /* [Nc] */
    const uint8_t* cam_fixed; /* [Nc] 1 = SetParameterBlockConstant (:105); any subset */
    const double*  pt;        /* [Np][3] */
    const int32_t* obs_cam;   /* [No] */
    const int32_t* obs_pt;    /* [No] non-decreasing */
    const double*  obs_uv;    /* [No][2] */
} ba_problem;

typedef struct ba_summary {
    int32_t termination;       /* ba_termination */
    int32_t usable;            /* mirrors Summary::IsSolutionUsable(), BundleAdjuster.h:128 */
    int32_t num_iterations;    /* records in the trajectory, including iteration 0 */
    int32_t num_successful_steps;
    int32_t num_unsuccessful_steps;
    int32_t num_residuals;     /* 2*No */
    int64_t total_pcg_iterations;
    double  initial_cost;      /* cost = 1/2 sum rho(||r||^2); chi2 = 2*cost */
    double  final_cost;
    double  solve_seconds;     /* engine: device time of ba_solve from CUDA events on its stream; oracle: wall clock */
    int64_t kernel_launches;   /* engine: CUDA kernels launched by this ba_solve; oracle: 0 */
    char    message[160];
} ba_summary;

/* One record per LM iteration — the parity artefact of SURVEY §8 (a-LM). */
typedef struct ba_iter_record {
    int32_t iteration;
    int32_t step_is_valid;
    int32_t step_is_successful;
    int32_t pcg_iterations;
    double  cost;               /* cost of the current x after this iteration */
    double  candidate_cost;     /* cost at x+delta (DBL_MAX if evaluation failed; 0 for iteration 0) */
    double  model_cost_change;
    double  relative_decrease;  /* rho_k */
    double  radius_used;        /* trust-region radius the step was computed with */
    double  radius;             /* radius after accept/reject */
    double  step_norm;
    double  gradient_max_norm;
    double  pcg_rel_residual;
} ba_iter_record;

/* Per-kernel device timing for the roofline report. */
typedef struct ba_kernel_stat {
    char    name[32];
    int64_t launches;
    double  total_ms;           /* CUDA-event time summed over launches (only when profiling is enabled) */
    double  bytes_per_launch;   /* algorithmic bytes, DESIGN.md formula for this kernel's formulation */
} ba_kernel_stat;

int  ba_options_default(ba_options* opt);
int  ba_create(const ba_options* opt, ba_engine** out);
int  ba_set_problem(ba_engine* e, const ba_problem* p);       /* copies (engine: host->device) and builds index structure */
int  ba_reset_state(ba_engine* e);                            /* restore the parameters given to ba_set_problem (engine: device-to-device) */

/* Incremental use.  The reference runs a full bundle adjustment after every image it adds to a growing map and
 * rebuilds the whole ceres::Problem from the map each time (actuator/SequentialActuator.h:138-229 -> BundleAdjuster.h:158-164,
 * 176-186).  ba_append grows the problem held by the handle instead: the state already on the device (the result of the last
 * ba_solve) is kept, new poses / Camera groups / points are appended after the existing ones (ids continue), new observations
 * may reference old and new ids in any order and are merged into point order (within a point: existing observations first, then
 * the new ones in the order given).  The merged state becomes the one ba_reset_state restores.  Single GPU only.
 * The delta is validated before anything changes (a rejected delta leaves the handle as it was); if the rebuild itself fails afterwards
 * (out of device memory, BA_LINEAR_EXACT limits exceeded by the grown problem, a CUDA error) the handle holds NO problem any more:
 * ba_solve then returns BA_ERR_NO_PROBLEM and the caller starts again with ba_set_problem (IncrementalBundleAdjuster does). */
typedef struct ba_delta {
    int32_t num_new_cameras;      const double* pose;  const int32_t* cam_group;  const uint8_t* cam_fixed;
    int32_t num_new_groups;       const double* intr;
    int32_t num_new_points;       const double* pt;
    int32_t num_new_observations; const int32_t* obs_cam;  const int32_t* obs_pt;  const double* obs_uv;
} ba_delta;
int  ba_append(ba_engine* e, const ba_delta* d);
/* Overwrite parameters held by the handle, both the current state and the one ba_reset_state restores: kind 0 = poses (6 doubles
 * each), 1 = intrinsics groups (4), 2 = points (3); ids [first, first + count).  For callers that keep a handle alive while the
 * map is edited outside bundle adjustment (a re-triangulated point, a pose re-estimated by PnP).  Single GPU only. */
int  ba_update_parameters(ba_engine* e, int32_t kind, int32_t first, int32_t count, const double* values);
/* The observation list in the handle's order (what per-observation outputs of ba_evaluate are indexed by); any pointer may be NULL. */
int  ba_get_observations(ba_engine* e, int32_t* obs_cam, int32_t* obs_pt, double* obs_uv, int32_t cap, int32_t* n);
int  ba_solve(ba_engine* e, ba_summary* out);                 /* blocking; whole LM loop */
int  ba_get_state(ba_engine* e, double* pose, double* intr, double* pt); /* any may be NULL */
int  ba_get_trajectory(ba_engine* e, ba_iter_record* out, int32_t cap, int32_t* n);
/* Index structure for the bit-exact check: perm_by_cam[No] (stable sort of observation ids by camera),
 * cam_off[Nc+1], pt_off[Np+1], schur_pairs[2*n_pairs] (i<j, both free, co-observing >=1 point; lexicographic).
 * Any pointer may be NULL; *n_pairs is always set when non-NULL. */
int  ba_get_structure(ba_engine* e, int32_t* perm_by_cam, int32_t* cam_off, int32_t* pt_off,
                      int32_t* schur_pairs, int64_t pairs_cap, int64_t* n_pairs);
/* One evaluation at the current parameters without solving: cost, residuals r[No][2] (robustified),
 * Jacobian blocks Jc[No][2][6], Jp[No][2][3], Ji[No][2][4] (robustified, NOT column-scaled), gradient
 * g[6Nc+4Ng+3Np] (pose, intrinsics, point order).  Any output may be NULL. For the K1/K2 parity tests. */
int  ba_evaluate(ba_engine* e, double* cost, double* r, double* Jc, double* Jp, double* Ji, double* g);
int  ba_set_profiling(ba_engine* e, int32_t mask);             /* per-kernel CUDA-event timing: 0 off (default), 1 or -1 all kernels, else bit k = kernel k of ba_get_kernel_stats */
int  ba_set_profiling_stride(ba_engine* e, int32_t stride);    /* time only every stride-th launch of the PCG-loop kernels (K5, PCG vector part): an event
                                                                 * between two launches serialises them, which would hide the programmatic dependent launch
                                                                 * of the chain from the very measurement; ba_get_kernel_stats scales total_ms to all launches */
int  ba_get_kernel_stats(ba_engine* e, ba_kernel_stat* out, int32_t cap, int32_t* n);
int  ba_linear_solver_in_use(ba_engine* e, int32_t* solver);  /* after ba_set_problem: BA_LINEAR_EXACT or BA_LINEAR_PCG, what AUTO resolved to for this problem */
/* Execution facts for reports (bench.py, tools): info[0] = linear solver in use (as ba_linear_solver_in_use), [1] = scatter mode in effect for
 * the current problem (1, 4 or 6), [2] = all-reduce of the PCG matvec (0: one GPU, 1: NCCL, 2: one-shot NVLink peer memory), [3] = host
 * synchronisations of the last ba_solve (Scalars read-backs), [4] = tiles and [5] = partial rows of this rank's shard, [6] = 1 when the PCG
 * matvec runs in its pose-factored form (streams the point columns only; BA_POSE_FACTORED=0 in the environment selects the explicit pose
 * columns), [7] = CTAs of the PCG tail's thread-block cluster (8, or 16 for >= 32768 reduced unknowns where the device can place one).
 * *n = entries the library has; at most cap are written.  The oracle reports zeros. */
int  ba_get_execution_info(ba_engine* e, int64_t* info, int32_t cap, int32_t* n);
int  ba_get_io_bytes(int64_t* h2d, int64_t* d2h);
/* ba_cost_noalloc has NO counterpart in the reference: it was added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  It evaluates cost = 1/2 sum rho(||r||^2)
 * (the candidate-cost step of the LM loop: camera precompute + residual / robust-loss sweep) of a problem whose arrays are ALREADY ON THE
 * DEVICE, and allocates nothing: no cudaMalloc / cudaFree / allocating library call, no engine handle.  Its scratch (per-camera records,
 * per-block partial sums) is one file-scope static device array of 672 KB that every GPU loading this library holds until exit; hence
 * the limit num_cameras <= BA_NOALLOC_MAX_CAMERAS (larger problems are rejected with BA_ERR_INVALID_ARGUMENT and a message in
 * ba_last_error(NULL); use a ba_engine handle for them).  Observations and points are not limited.
 * All d_* pointers are device pointers on `device`: pose[Nc][6], intr[Ng][4], cam_group[Nc], pt[Np][3], obs_cam[No], obs_pt[No], obs_uv[No][2]
 * (16-byte aligned), outputs d_cost[1] and d_bad[1] (1 if any residual was non-finite).  The work is enqueued on `stream` (a cudaStream_t, NULL =
 * the default stream) and NOT synchronised: the outputs are valid once the stream has reached this point.  All calls on a device share the
 * array, which keeps the last call's contents (what must start at zero is cleared by every call): calls that overlap on the GPU — two
 * streams, two host threads, two processes' worth of graphs replayed concurrently on one device — race and give wrong results.  The first
 * call on a device looks up the array's device address (cudaGetSymbolAddress); make that call outside stream capture. */
#define BA_NOALLOC_MAX_CAMERAS 4096
int  ba_cost_noalloc(int32_t device, void* stream, int32_t model, int32_t robust_kind, double robust_a, int32_t taylor_rule,
                     int32_t num_cameras, int32_t num_groups, int32_t num_observations, const double* d_pose, const double* d_intr,
                     const int32_t* d_cam_group, const double* d_pt, const int32_t* d_obs_cam, const int32_t* d_obs_pt,
                     const double* d_obs_uv, double* d_cost, int32_t* d_bad);              /* cumulative host->device / device->host bytes moved by this library in this process */
/* Host-only (no device needed): the point ranges of the multi-GPU sharding. Rank r owns points
 * [shard_pt[r], shard_pt[r+1]) and all their observations; ranges are contiguous, cut at point boundaries and
 * balanced by observation count.  obs_pt[No] non-decreasing; shard_pt has world_size+1 entries. */
int  ba_shard_points(int32_t num_points, int32_t num_observations, const int32_t* obs_pt, int32_t world_size, int32_t* shard_pt);
int  ba_nccl_unique_id(void* out128);                          /* rank 0 only; engine library only */
const char* ba_last_error(const ba_engine* e);                /* e may be NULL: last error of a failed ba_create on this thread */
const char* ba_backend(void);                                 /* "cuda-sm100a" or "cpu-oracle" */
void ba_destroy(ba_engine* e);

#ifdef __cplusplus
}
#endif
#endif /* BA_ENGINE_H */
