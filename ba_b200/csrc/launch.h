// launch.h — host-side launchers of the CUDA kernels (one translation unit per kernel family).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "ba_gpu.h"

namespace bagpu {

constexpr int kSelCandCap = 8191;   // candidates the final single-CTA selection takes (= one rank's share of a peer gather)
struct SelectState {
  unsigned int hist[4096];
  unsigned long long prefix, k_rem, prefix_next, k_rem_next;
  unsigned long long bucket_count;   // values (all ranks) in the bucket the last scan chose
  unsigned int cand_n, all_n;        // local candidates written by k_select_compact; candidates of all ranks after the gather
};

// unary / binary / inertial residual records in HBM (SoA)
struct OtherResiduals {
  int n_unary, n_binary, n_imu;
  // unary (reference Types.h:255-266)
  const uint32_t* un_pose; const double* un_t;   // 7 per residual
  double* un_cov_inv;                            // 36, multiplied by the Huber weight every build (Q6)
  const uint8_t* un_use_rot;
  double* un_maha;                               // scratch: r^T cov_inv r
  // per-residual outputs of the last EvaluateResiduals (cpp:190-256: res.residual, res.mahalanobis_distance)
  double* un_out;                                // 7 per unary residual: residual (6) | mahalanobis_distance
  double* bin_out;                               // 7 per binary residual
  double* imu_out;                               // D + 1 per inertial residual
  // binary (Types.h:268-280)
  const uint32_t *bin_p1, *bin_p2; const double* bin_t12; const double* bin_cov_inv; const double* bin_weight;
  const uint8_t* bin_use_rot;
  // inertial (Types.h:300-321)
  const uint32_t *imu_p1, *imu_p2, *imu_meas_start; const double* imu_meas; const double* imu_weight;
  double* imu_cov_inv;      // D*D per residual, of the last build (weighted)
  double* imu_J1; double* imu_J2;   // D*D per residual (row-major), kept for the SD denominator
  double* imu_res;          // D per residual
  // options.calculate_inertial_covariance_once (parallel_algos.h:190-205): dintegration_db (60) and c_integration (100) of
  // the first build are kept per residual and reused by every later one
  int cov_once; double* imu_once; uint8_t* imu_once_done;
  double gravity[3];
  double imu_r[6];          // sensor noise diagonal (imu_.r)
  double imu_rb[6];         // bias random walk (imu_.r_b)
};

// ---- pose-pair-major reduction (schur_kernels.cu / symbolic.cu) -------------------------------
// pair list entry: (a, b) observation indices (positions in the landmark-major stream), a belongs to the block's row
// pose.  LmSize 1 also lists one entry per observation for the (measurement pose, reference pose) block, tagged in b.
constexpr uint32_t kPairTagMeasRow = 0xffffffffu;   // row pose = measurement pose: += At^T Q
constexpr uint32_t kPairTagRefRow = 0xfffffffeu;    // row pose = reference pose:   += Q^T At
enum { kItemDiagObs = 0, kItemDiagLm = 1, kItemOff = 2, kItemSym = 3, kItemOffRef = 4 };
struct ReduceItem {        // one warp's share of one S block
  uint32_t begin, end;     // range of the pose's observation / landmark list (diagonal kinds) or pair range
  uint32_t row, col;       // pose opt ids, row <= col
  uint32_t kind;
  uint32_t mfirst, mcount; // chunks of the same block: partial-sum slots mfirst .. mfirst + mcount
  uint32_t mself, mgroup;  // this chunk's slot, the block's arrival counter
  uint32_t pad;
};

// Builds the pair list on the device (symbolic.cu): *pairs_out is allocated with dev_alloc (caller frees) and holds three
// arrays of n_pairs entries [observation of the row pose | observation of the column pose or tag | landmark], sorted by
// key = (row * n_active_poses + col) * 2 + (1 for the LmSize-1 measurement/reference entries), stable in landmark
// order; block_key / block_count = the run table of that key.
int build_pair_lists(const DeviceProblem& P, cudaStream_t st, uint32_t** pairs_out, size_t* n_pairs_out,
                     std::vector<unsigned long long>& block_key, std::vector<uint32_t>& block_count, std::string& err);

// ---- device-side symbolic set-up (setup_device.cu) ------------------------------------------------
int sort_positions_by_key(const uint32_t* keys_in, uint32_t* keys_out, uint32_t* vals_out, uint32_t n, uint32_t max_key, cudaStream_t st,
                          std::string& err);
void launch_key_starts(const uint32_t* keys_sorted, uint32_t n, uint32_t n_keys, uint32_t* starts, cudaStream_t st);
void launch_validate_obs(const uint32_t* pose, const uint32_t* lm, const uint32_t* cam, uint32_t n, uint32_t n_poses, uint32_t n_lms, int L,
                         const uint32_t* lm_ref_pose, const uint32_t* lm_ref_cam, int* flag2, cudaStream_t st);
void launch_obs_flags(const uint32_t* pose, const uint32_t* lm, uint32_t n, int L, const uint32_t* lm_ref_pose, const int* pose_opt,
                      uint8_t* has_res, unsigned int* n_cond, cudaStream_t st);
void launch_gather_obs(const uint32_t* perm, uint32_t n, const double2* z, const uint32_t* pose, const uint32_t* cam, const double* w0,
                       double2* oz, uint32_t* opose, uint32_t* ocam, double* ow0, uint32_t* operm, cudaStream_t st);
void launch_local_starts(const uint32_t* gstarts, uint32_t n_lms, uint32_t s_lo, uint32_t s_hi, uint32_t* out, cudaStream_t st);
void launch_slot_keys(const uint32_t* obs_pose, const uint32_t* obs_lm, uint32_t n, const uint32_t* lm_ref_pose, const int* pose_opt,
                      uint32_t n_ap, uint32_t* keys, uint8_t* lm_any, cudaStream_t st);
void launch_pose_lists(const uint32_t* vals_sorted, uint32_t n, const uint32_t* obs_lm, uint32_t* pose_obs, uint32_t* pose_obs_lm, cudaStream_t st);
void launch_rslot_keys(const uint8_t* lm_any, const uint32_t* lm_ref_pose, const int* pose_opt, uint32_t n_lms, uint32_t n_ap, uint32_t* keys,
                       cudaStream_t st);
// dense nT x nT tile table (-1 = structurally empty) from the CSR rows: tiles numbered row by row, diagonal tile first
void launch_fill_tile_map(const int* row_ptr, const int* row_col, int nT, int* tile_map, cudaStream_t st);
void launch_tile_pattern(const uint32_t* starts, const uint32_t* perm, const uint32_t* pose, const uint32_t* lm_ref_pose, const int* pose_opt,
                         uint32_t n_lms, int L, int nT, uint8_t* present, cudaStream_t st);
void launch_cond_error(const uint32_t* obs_pose, const uint32_t* obs_lm, const uint32_t* obs_perm, uint32_t n, const uint32_t* lm_ref_pose,
                       const int* pose_opt, const double* maha_by_id, const double* w, double* out, cudaStream_t st);
void launch_scatter_by_perm(const uint32_t* perm, const double* in, uint32_t n, double* out, cudaStream_t st);
void launch_iota_i32(int* v, uint32_t n, cudaStream_t st);   // v[i] = i
void launch_reliable_merge(uint8_t* rel, unsigned int* unrel, uint32_t n, int phase, cudaStream_t st);

// block-Jacobi PCG on the tiled S (pcg.cu)
size_t pcg_work_doubles(int nT, int T, int D);
int launch_pcg_solve(int D, const double* S, const int* tile_map, const int* row_ptr, const int3* ent, int nT, const double* b, double* x_out,
                     double* work, double* red_partial, unsigned int* red_counter, double tol, int max_iter, int* iters_out, uint64_t* launches,
                     cudaStream_t st);

void set_sm_count(int n);
void launch_pose_frames(const DeviceProblem& P, cudaStream_t st);
void launch_prepare_xs(const DeviceProblem& P, const double* lm_xw, cudaStream_t st);
void launch_finish_xw(const DeviceProblem& P, double* lm_xw, cudaStream_t st);
void launch_residual(const DeviceProblem& P, int G, cudaStream_t st);
// packed tasks: (first observation, number of observations <= 32) of runs of whole landmarks; long_lms: landmarks with > 32 observations
void launch_linearize(const DeviceProblem& P, const uint2* tasks, int n_tasks, const uint32_t* long_lms, int n_long, cudaStream_t st);
void launch_reduce_blocks(const DeviceProblem& P, const ReduceItem* items, int n_items, const uint32_t* pair_a, const uint32_t* pair_b,
                          const uint32_t* pair_lm, double* partial, unsigned int* counters, unsigned int* next_item, cudaStream_t st);
void launch_sd_denominator(const DeviceProblem& P, int G, cudaStream_t st);
// per residual id: r sqrt(w) | dz_dx_meas sqrt(w) | dz_dx_ref sqrt(w) | dz_dlm sqrt(w) | meas / ref opt ids (28 + 2 L doubles)
void launch_jacobian_rows(const DeviceProblem& P, double* out, cudaStream_t st);
void launch_back_substitute(const DeviceProblem& P, const uint2* tasks, int n_tasks, const uint32_t* long_lms, int n_long, const double* delta_p,
                            double* delta_l, cudaStream_t st);
void launch_evaluate(const DeviceProblem& P, int G, double* out_sum, double* res_out, double* maha_out, int write_maha0, int count_outliers,
                     cudaStream_t st);
// CTAs of `kernel` (block threads, smem dynamic shared bytes) that are resident on one SM at a time (occupancy API, cached).
// The grid-stride kernels are launched as ONE full wave of resident CTAs: a grid of "8 per SM" of a kernel of which 6 fit
// runs its last 2 per SM alone at a quarter of the occupancy, for as long as the first 6 took.
int resident_ctas_per_sm(const void* kernel, int block, size_t smem = 0);

// the arrays the dogleg loop restores a rejected step from
struct StateBackup { double *q, *t, *v, *b, *lm_x; uint8_t* rel; uint32_t* outliers; };
void launch_blend_diff(const DeviceProblem& P, double* tmp_p, const double* gn_p, const double* sd_p, size_t np, double* tmp_l, const double* gn_l,
                       const double* sd_l, size_t nl, double f, double* out4, cudaStream_t st);
void launch_blend_step(const DeviceProblem& P, double* dl_p, const double* tmp_p, const double* sd_p, size_t np, double* dl_l, const double* tmp_l,
                       const double* sd_l, size_t nl, double beta, double f, double* out2, cudaStream_t st);
void launch_update_poses(const DeviceProblem& P, const double* delta_p, double coef, cudaStream_t st, const StateBackup* bk = nullptr);
void launch_update_landmarks(const DeviceProblem& P, const double* delta_l, double coef, cudaStream_t st, const StateBackup* bk = nullptr);
void launch_mask_diagonal(const DeviceProblem& P, const uint8_t* mask_used, int global_flag, cudaStream_t st);
void launch_dot(const DeviceProblem& P, const double* a, const double* b, size_t n, double* out, cudaStream_t st);   // overwrites *out
void launch_axpby(double* out, double alpha, const double* x, double beta, const double* y, size_t n, cudaStream_t st);
void launch_compact_bl(const DeviceProblem& P, double* out, cudaStream_t st);
void launch_permute_tiles(double* dst, const double* src, const int* perm, int T, int nT, int forward, cudaStream_t st);

int select_num_passes();
void launch_select_init(SelectState* st, unsigned long long k, cudaStream_t s);
void launch_select_hist(const double* x, size_t n, SelectState* st, int pass, int sms, cudaStream_t s);
void launch_select_scan(SelectState* st, int pass, double* out, cudaStream_t s);
// keys whose top 12 * passes_done bits equal the prefix -> cand[0 .. st->cand_n)
void launch_select_compact(const double* x, size_t n, SelectState* st, int passes_done, unsigned long long* cand, int sms, cudaStream_t s);
// the st->k_rem-th smallest of all[0 .. *all_n) (at most kSelCandCap keys) -> *out
void launch_select_final(const unsigned long long* all, const unsigned int* all_n, SelectState* st, double* out, cudaStream_t s);

// reduced solve (chol.cu): tile tasks of the persistent dataflow kernel, built on the host from the tile pattern
struct CholTask { int kind, i, j, slot, lb, le, pad0, pad1; };
struct CholTerm { int sa, sb, k; };   // slots of U_ki, U_kj and the row k of one left-looking update term
void chol_build_plan(int T, int nT, const std::vector<int>& row_ptr, const std::vector<int>& row_col,
                     std::vector<CholTask>& tasks, std::vector<CholTerm>& terms);
size_t chol_sync_ints(int n_tiles, int nT);
// sync_ints[1] holds the first failing pivot + 1 (0 = ok) after the solve; returns a cudaError_t
int launch_cholesky_solve(int T, double* tiles, double* uinv, double* work /* 3*nT*T + 2*n_tiles*T */, double* rhs,
                          const CholTask* tasks, const CholTerm* terms, const int* row_ptr, const int* row_col, int n_tasks, int* sync_ints,
                          int n_tiles, int nT, int sms,
                          cudaStream_t st, unsigned long long* trace = nullptr);

// unary / binary / inertial residuals (other_kernels.cu)
void launch_unary_maha(const DeviceProblem& P, const OtherResiduals& R, cudaStream_t st);
void launch_other_build(const DeviceProblem& P, const OtherResiduals& R, cudaStream_t st, uint64_t* launches);
void launch_other_sd(const DeviceProblem& P, const OtherResiduals& R, cudaStream_t st, uint64_t* launches);
void launch_other_evaluate(const DeviceProblem& P, const OtherResiduals& R, cudaStream_t st, uint64_t* launches);

}  // namespace bagpu
