/*
 * ba_b200.h — C ABI of the B200-native optimisation inner loop of ba::BundleAdjuster.
 *
 * This is the drop-in boundary: everything the reference does inside
 *   BundleAdjuster::Solve            (reference src/BundleAdjuster.cpp:278-705)
 *   BundleAdjuster::BuildProblem     (reference src/BundleAdjuster.cpp:1166-1803)
 *   BundleAdjuster::SolveInternal    (reference src/BundleAdjuster.cpp:838-1161)
 *   BundleAdjuster::CalculateGn      (reference src/BundleAdjuster.cpp:748-833)
 *   BundleAdjuster::GetLandmarkDelta (reference src/BundleAdjuster.cpp:709-744)
 *   BundleAdjuster::ApplyUpdate      (reference src/BundleAdjuster.cpp:21-140)
 *   BundleAdjuster::EvaluateResiduals(reference src/BundleAdjuster.cpp:144-274)
 * happens behind these entry points, on the GPU.  The host template class
 * (include/ba/BundleAdjuster.h in this repo, mirroring reference
 * include/ba/BundleAdjuster.h:111-753) only does bookkeeping and calls this ABI.
 *
 * Plain C: pointers and sizes only.  All floating point arrays are double (the
 * reference's default REAL_TYPE, reference CMakeLists.txt:16-20) and the device computes in FP64;
 * bagpu_dims.scalar_bytes names the caller's REAL_TYPE (the float instantiation converts on the host side).
 *
 * Conventions
 *   SE3 = 7 doubles: qx qy qz qw tx ty tz   (Hamilton quaternion, Eigen coefficient order)
 *   ids are uint32_t, identical to the ids the reference's Add* calls return.
 *   all functions return bagpu_status (0 = ok); bagpu_last_error() gives a message.
 */
#ifndef BA_B200_H
#define BA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum bagpu_status {
  BAGPU_OK = 0,
  BAGPU_ERR_INVALID = 1,     /* bad argument / bad ids (reference asserts, BundleAdjuster.h:332,383,433,465,521) */
  BAGPU_ERR_NO_DEVICE = 2,   /* no CUDA device: the product path never falls back to the CPU */
  BAGPU_ERR_CUDA = 3,
  BAGPU_ERR_NCCL = 4,
  BAGPU_ERR_UNSUPPORTED = 5  /* template instantiation outside this build (e.g. CalibSize>0) */
} bagpu_status;

/* reference include/ba/BundleAdjuster.h:38-46 (same numeric values) */
typedef enum bagpu_result {
  BAGPU_Success = 0,
  BAGPU_ErrorIncreased = 1,
  BAGPU_ErrorChangeBelowThreshold = 2,
  BAGPU_ParamChangeBelowThreshold = 3,
  BAGPU_FactorizationError = 4,
  BAGPU_SolverError = 5
} bagpu_result;

/* Template parameters of ba::BundleAdjuster<Scalar,LmSize,PoseSize,CalibSize,DoTvs>
 * (reference include/ba/BundleAdjuster.h:111-134). */
typedef struct bagpu_dims {
  int32_t scalar_bytes; /* sizeof(REAL_TYPE) of the caller (reference CMakeLists.txt:16-20): 8, or 4 for the float
                           instantiation, whose host class converts at this boundary; every array here is double and
                           the device computes in FP64 either way */
  int32_t lm_dim;       /* LmSize   : 0, 1 (inverse depth) or 3 */
  int32_t pose_dim;     /* PoseSize : 6, 9 or 15 */
  int32_t calib_dim;    /* CalibSize: 0 in this build */
  int32_t do_tvs;       /* DoTvs    : 0 in this build */
} bagpu_dims;

/* POD mirror of ba::Options<Scalar> (reference include/ba/BundleAdjuster.h:72-107)
 * plus gravity (SetGravity, BundleAdjuster.h:243-252) and the solver choice. */
typedef struct bagpu_options {
  double trust_region_size;            /* -1 = kTrustRegionAuto (BundleAdjuster.h:30) */
  double gyro_sigma, accel_sigma, gyro_bias_sigma, accel_bias_sigma;
  double projection_outlier_threshold;
  double error_change_threshold;
  double param_change_threshold;
  uint32_t dogleg_max_inner_iterations;
  int32_t apply_results;
  int32_t use_dogleg;
  int32_t use_triangular_matrices;
  int32_t use_sparse_solver;
  int32_t write_reduced_camera_matrix;
  int32_t calculate_calibration_marginals;
  int32_t use_per_pose_cam_params;
  int32_t regularize_biases_in_batch;
  int32_t enable_auto_regularization;
  int32_t use_robust_norm_for_proj_residuals;
  int32_t use_robust_norm_for_inertial_residuals;
  int32_t calculate_inertial_covariance_once;
  double gravity[3];                   /* imu_.g_vec */
  uint32_t root_pose_id;               /* SetRootPoseId, BundleAdjuster.h:553 */
  /* extensions (not in the reference): */
  int32_t reduced_solver;              /* 0 = tiled Cholesky (default), 1 = block-Jacobi PCG */
  int32_t record_deltas;               /* keep per-iteration delta_p / delta_l for parity tests */
  double pcg_tolerance;                /* relative residual for solver 1 */
  uint32_t pcg_max_iterations;
  int32_t keep_on_device;              /* skip the end-of-Solve() state hand-over (pose read-back, world-frame landmark
                                          positions): for timing the iterations alone.  A following Solve() continues
                                          from the refined poses but re-derives the landmarks from the positions of the
                                          last hand-over; the next solve without this flag, or bagpu_sync_state, hands
                                          the refined state over */
} bagpu_options;

/* reference ba::SolutionSummary<Scalar> (BundleAdjuster.h:48-70) + the private error
 * members GetErrors returns (BundleAdjuster.h:593-602) + trust_region_size_ (:720). */
typedef struct bagpu_summary {
  uint32_t num_proj_residuals, num_inertial_residuals;
  uint32_t num_cond_proj_residuals, num_cond_inertial_residuals;
  double cond_proj_error, cond_inertial_error;
  double proj_error, inertial_error, unary_error, binary_error;
  double delta_norm, pre_solve_norm, post_solve_norm;
  double trust_region_size;
  int32_t result;            /* bagpu_result */
  uint32_t iterations;       /* outer iterations executed */
} bagpu_summary;

/* One record per outer iteration; this is what the parity tests compare. */
typedef struct bagpu_iter_trace {
  double proj_error_build, binary_error_build, unary_error_build, inertial_error_build;
  double median_maha;        /* nth_element value, src/BundleAdjuster.cpp:1356-1358 */
  double sd_numerator, sd_denominator, sd_norm;   /* src/BundleAdjuster.cpp:858-929 */
  double gn_norm;            /* src/BundleAdjuster.cpp:971-973 (0 if GN not computed) */
  double trust_region_after;
  double pre_solve_norm, post_solve_norm, delta_norm;
  double delta_p_norm, delta_l_norm;
  uint32_t inner_iterations;
  int32_t result;
} bagpu_iter_trace;

typedef struct bagpu_handle_s* bagpu_handle;

/* ---- lifetime ------------------------------------------------------------------- */
int bagpu_create(const bagpu_dims* dims, bagpu_handle* out);
void bagpu_destroy(bagpu_handle h);
const char* bagpu_last_error(void);
int bagpu_set_options(bagpu_handle h, const bagpu_options* opt);
void bagpu_default_options(bagpu_options* opt);   /* values of BundleAdjuster.h:75-106 */

/* Page-locked host buffers for the arrays handed to bagpu_set_* / bagpu_get_*: copies from and to them run as direct
 * DMA instead of being staged by the driver.  Optional (any host pointer is accepted by the setters); without a CUDA
 * device the buffer is ordinary memory.  ba/BundleAdjuster.h keeps its structure-of-arrays observation buffers in
 * them (the reference's std::vector storage, BundleAdjuster.h:738-745). */
int bagpu_host_alloc(size_t bytes, void** out);
int bagpu_host_free(void* p);

/* ---- problem upload (host -> HBM).  Replaces the AoS vectors BundleAdjuster.h:738-745 -- */
/* rig: calibu linear camera params fx fy u0 v0 and T_vs per camera (BundleAdjuster.h:259) */
int bagpu_set_cameras(bagpu_handle h, uint32_t n, const double* params4, const double* t_vs7);
/* AddPose (BundleAdjuster.h:267-323). mask: pose_dim bytes per pose, 1 = parameter free
 * (PoseT::param_mask, Types.h:47); mask_used = PoseT::is_param_mask_used. May be NULL. */
int bagpu_set_poses(bagpu_handle h, uint32_t n, const double* t_wp7, const double* v_w3,
                    const double* b6, const uint8_t* is_active,
                    const uint8_t* mask_used, const uint8_t* mask);
/* AddLandmark (BundleAdjuster.h:326-367) */
int bagpu_set_landmarks(bagpu_handle h, uint32_t n, const double* x_w4,
                        const uint32_t* ref_pose_id, const uint32_t* ref_cam_id,
                        const uint8_t* is_active);
/* AddProjectionResidual (BundleAdjuster.h:459-513): only residuals that call kept. */
int bagpu_set_proj_residuals(bagpu_handle h, uint32_t n, const double* z2,
                             const uint32_t* meas_pose_id, const uint32_t* landmark_id,
                             const uint32_t* cam_id, const double* orig_weight);
/* AddUnaryConstraint (BundleAdjuster.h:377-407): cov_inv = covariance^-1, 36 doubles row-major */
int bagpu_set_unary_residuals(bagpu_handle h, uint32_t n, const uint32_t* pose_id,
                              const double* t_wp7, const double* cov_inv36,
                              const uint8_t* use_rotation);
/* AddBinaryConstraint (BundleAdjuster.h:410-456) */
int bagpu_set_binary_residuals(bagpu_handle h, uint32_t n, const uint32_t* pose1_id,
                               const uint32_t* pose2_id, const double* t_12_7,
                               const double* cov_inv36, const double* weight,
                               const uint8_t* use_rotation);
/* AddImuResidual (BundleAdjuster.h:516-546): CSR over measurements,
 * meas7 = wx wy wz ax ay az time (Types.h:222-244) */
int bagpu_set_imu_residuals(bagpu_handle h, uint32_t n, const uint32_t* pose1_id,
                            const uint32_t* pose2_id, const uint32_t* meas_start /* n+1 */,
                            const double* meas7, const double* weight);

/* ---- the hot path ------------------------------------------------------------- */
/* BundleAdjuster::Solve(uMaxIter, gn_damping, error_increase_allowed), BundleAdjuster.h:549 */
int bagpu_solve(bagpu_handle h, uint32_t max_iter, double gn_damping,
                int error_increase_allowed, bagpu_summary* out);

/* One BuildProblem + Schur assembly at the current state without stepping: the
 * "Schur-build" phase of the metric (reference timers _BuildProblem_ + _steup_problem_,
 * src/BundleAdjuster.cpp:300-302,309,629).  ms_out (may be NULL) = device time. */
int bagpu_build(bagpu_handle h, float* ms_out);

/* ---- results (HBM -> host). Buffers may be NULL to skip. -------------------------
 * After a solve only the poses are copied back eagerly; the refined landmarks stay in HBM and
 * bagpu_get_landmarks copies them straight into the caller's arrays (page-locked ones, see
 * bagpu_host_alloc, take the copy at full PCIe rate).  The library's own host copies of the
 * landmark state are refreshed by bagpu_sync_state. */
int bagpu_get_poses(bagpu_handle h, double* t_wp7, double* v_w3, double* b6,
                    uint8_t* mask_used, uint8_t* mask);
int bagpu_get_landmarks(bagpu_handle h, double* x_w4, double* x_s4, uint8_t* is_reliable,
                        uint32_t* num_outlier_residuals);
int bagpu_get_proj_residuals(bagpu_handle h, double* residual2, double* weight, double* maha);
/* What Solve() leaves in the unary / binary / inertial residual structs (reference src/BundleAdjuster.cpp:190-256:
 * res.residual and res.mahalanobis_distance of the last EvaluateResiduals; cpp:1469,1526: res.cov_inv, multiplied by the
 * Huber weight in place).  residual: 6 doubles per unary / binary residual, PoseSize per inertial residual; cov_inv:
 * 36 / PoseSize^2 row-major.  Any pointer may be NULL.  On several GPUs every rank must call it (one small reduction). */
int bagpu_get_other_residuals(bagpu_handle h, double* un_residual6, double* un_maha, double* un_cov_inv36, double* bin_residual6,
                              double* bin_maha, double* imu_residualD, double* imu_maha, double* imu_cov_invDD);
int bagpu_get_trace(bagpu_handle h, uint32_t max_records, bagpu_iter_trace* out, uint32_t* n_out);
/* delta of outer iteration `iter` (needs options.record_deltas) */
int bagpu_get_deltas(bagpu_handle h, uint32_t iter, double* delta_p, double* delta_l);

/* ---- inspection of the assembled system (parity of block structure and values) --- */
typedef enum bagpu_pattern { BAGPU_PATTERN_U = 0, BAGPU_PATTERN_W = 1, BAGPU_PATTERN_S = 2 } bagpu_pattern;
/* block-CSC, column pointers (ncols+1) and ascending row indices, as the reference's
 * SparseBlockMatrix stores them (SparseBlockMatrix.h:108-112). First call with NULLs to size. */
int bagpu_get_pattern(bagpu_handle h, int which, uint32_t* ncols, uint32_t* nnzb,
                      uint32_t* colptr, uint32_t* rowidx);
/* dense upper-triangular S ((P*d)^2, column-major like Eigen MatrixXt, BundleAdjuster.h:718),
 * rhs_p_sc, rhs_p_, rhs_l_, vi_ (L*L per active landmark) of the last build */
int bagpu_get_system(bagpu_handle h, double* s_dense, double* rhs_sc, double* rhs_p,
                     double* rhs_l, double* v_inv);
/* s.txt / rhs.txt of the last build in the reference's CSV format (src/BundleAdjuster.cpp:600-606, Eigen kLongCsvFmt of
 * include/ba/Utils.h:66: ", " between coefficients, "\n" between rows; %.17g digits).  With
 * options.write_reduced_camera_matrix set, bagpu_solve writes both files into the working directory at every outer
 * iteration, as the reference does.  j_pr.txt / r_pr.txt / j_l.txt are not written: the build is matrix-free and never
 * forms J.  ba_b200/dumpio.py reads the files and converts them to the compact binary twin. */
int bagpu_dump(bagpu_handle h, const char* dir);

/* ---- multi-GPU (landmark sharding) --------------------------------- */
/* nccl_unique_id: 128 bytes from bagpu_nccl_unique_id() on rank 0, broadcast by the caller.
 * bagpu_comm_init is collective: it creates the NCCL communicator, connects its channels (one small
 * and one bulk warm-up all-reduce) and maps the peers' windows for the small reductions over NVLink
 * peer memory (CUDA IPC; all ranks fall back to NCCL together if a window cannot be mapped;
 * BAGPU_P2P=0 in the environment forces NCCL). */
int bagpu_nccl_unique_id(void* out128);
int bagpu_comm_init(bagpu_handle h, int rank, int world, const void* nccl_unique_id128);
/* per-phase device timings of the last solve/build in ms: see bagpu_phase */
typedef enum bagpu_phase {
  BAGPU_PH_RESIDUAL = 0, BAGPU_PH_MEDIAN = 1, BAGPU_PH_ACCUMULATE = 2, BAGPU_PH_OTHER_RES = 3, BAGPU_PH_SCHUR = 4,
  BAGPU_PH_ALLREDUCE = 5, BAGPU_PH_SOLVE = 6, BAGPU_PH_STEP = 7, BAGPU_PH_TOTAL = 8, BAGPU_PH_COUNT = 9
} bagpu_phase;
int bagpu_get_phase_ms(bagpu_handle h, float* ms /* BAGPU_PH_COUNT */, uint64_t* kernel_launches);
/* bring the device state back into the host copies after a solve with options.keep_on_device */
int bagpu_sync_state(bagpu_handle h);

/* Host-only symbolic step of the reduced solve, exposed for tests (no device needed): from the upper pattern of an
 * nT x nT tile matrix given as n_edges pairs (a, b), the elimination order (nested dissection; replaces the AMD
 * ordering inside Eigen::SimplicialLDLT, reference src/BundleAdjuster.cpp:753-761) and the filled pattern in that
 * order as CSR rows of the strict upper triangle.  perm[natural tile] = position; row_ptr has nT + 1 entries;
 * row_col takes up to cap entries (BAGPU_ERR_INVALID if it does not fit). */
/* Host-only, for tests: the work-item table of the Schur reduction.  Replaces the column walks of the reference's sparse
 * block products (include/ba/SparseBlockMatrixOps.h:182-254) as the unit of work.  pose_rec_start[0..n_active_poses]:
 * pose-major record slots; block_key (sorted) = (row * n_active_poses + col) * 2 + kind bit, block_count = entries of
 * each run in the pair list.  Items are ten uint32 each: begin, end, row, col, kind (0 diagonal over records, 1
 * diagonal over reference-pose records, 2 off-diagonal, 3 same-pose pairs, 4 measurement/reference entries), mfirst,
 * mcount, mself, mgroup, pad.  *n_items / *n_sym: capacity in, count out. */
int bagpu_build_work_items(uint32_t n_active_poses, int lm_dim, const uint32_t* pose_rec_start, const uint32_t* pose_lmrec_start,
                           uint32_t n_blocks, const uint64_t* block_key, const uint32_t* block_count, uint32_t* items10,
                           uint32_t* n_items, uint32_t* sym10, uint32_t* n_sym, uint32_t* n_partial, uint32_t* n_groups);
/* Host-only, for tests: the warp tasks of the landmark-major kernels for the landmark range [lm_lo, lm_hi) of a
 * landmark-major observation stream with CSR bounds starts[0..n_lms] (the per-landmark residual lists of the
 * reference, include/ba/Types.h:186-217, flattened).  A task is a run of whole landmarks with at most 32 observations
 * in total: (first observation relative to starts[lm_lo], count); landmarks with more than 32 observations go to
 * long_lms.  *n_tasks / *n_long: capacity in, count out. */
int bagpu_pack_warp_tasks(uint32_t n_lms, const uint32_t* starts, uint32_t lm_lo, uint32_t lm_hi, uint32_t* task_first,
                          uint32_t* task_count, uint32_t* n_tasks, uint32_t* long_lms, uint32_t* n_long);
int bagpu_symbolic_tile_fill(uint32_t nT, uint32_t n_edges, const uint32_t* edge_a, const uint32_t* edge_b,
                             uint32_t* perm, uint32_t* row_ptr, uint32_t* row_col, uint32_t cap);
/* Host-only, for tests: the task list of the dataflow factorisation + substitution kernel for a filled tile pattern
 * (row_ptr / row_col as bagpu_symbolic_tile_fill returns them; tile_dim = 8 x PoseSize = 48, 72 or 120).  Replaces the
 * elimination schedule inside Eigen::SimplicialLDLT (reference src/BundleAdjuster.cpp:748-833).  The CTAs of the kernel
 * claim tasks in list order from one counter and wait on flags of other tasks' outputs, so the ORDER is what keeps the
 * kernel free of deadlock; tests/test_symbolic_cpu.py replays it with a few workers.  Task = eight int32: kind (0 pivot
 * tile: factor + invert, 1 off-diagonal tile: update + triangular solve, 2 backward tile product U_ij x_j, 4 backward
 * chain of the dependent rows j .. i solved by one CTA), i, j, tile slot, first / end index of its update terms, 0, 0.
 * Term = three int32: slots of U_ki and U_kj, row k.  Slots: row by row, the diagonal tile first (slot(k, k) = k +
 * row_ptr[k]).  *n_tasks / *n_terms: capacity in, count out (BAGPU_ERR_INVALID if it does not fit). */
int bagpu_chol_task_plan(int tile_dim, uint32_t nT, const uint32_t* row_ptr, const uint32_t* row_col, int32_t* tasks8,
                         uint32_t* n_tasks, int32_t* terms3, uint32_t* n_terms);

/* Allocation-free exact k-th order statistic (0-based; k = n / 2 is the value std::nth_element leaves at
 * errors_.begin() + n / 2, i.e. what the build calls the median).  NO COUNTERPART IN THE REFERENCE: added
 * for callers that may not call cudaMalloc (CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  d_x: n non-negative finite doubles (sign bit clear) in DEVICE memory of the current device; d_out: one
 * double in device memory; stream: a cudaStream_t (NULL = the default stream).  The call only queues kernels on `stream`
 * (13 launches, no synchronisation, no host round trip): *d_out holds the value once that work has run.  The radix-select
 * state lives in a 16640-byte static __device__ array inside the library (held by every GPU that loads it until exit),
 * shared by all calls: it is cleared on `stream` at the start of every call, and calls that overlap on the GPU race.
 * The first call on a device loads the module — make it before a graph capture begins.
 * n > BAGPU_SELECT_STATIC_MAX_N (32-bit histogram bins), n == 0, k >= n or a null pointer: BAGPU_ERR_INVALID. */
#define BAGPU_SELECT_STATIC_MAX_N (1ull << 31)
int bagpu_select_kth_static(const double* d_x, uint64_t n, uint64_t k, double* d_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* BA_B200_H */
