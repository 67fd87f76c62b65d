/* TEST INFRASTRUCTURE — C entry points of the CPU oracle (PARITY UNPINNED, see minimath.h).
 * Same signatures as include/ba_b200.h with the prefix baoracle_, so the parity tests feed
 * identical arrays to the oracle and to the CUDA path.  Never linked into the product. */
#ifndef BA_ORACLE_H
#define BA_ORACLE_H
#include "../include/ba_b200.h"
#ifdef __cplusplus
extern "C" {
#endif
typedef struct baoracle_handle_s* baoracle_handle;
int baoracle_create(const bagpu_dims* dims, baoracle_handle* out);
void baoracle_destroy(baoracle_handle h);
const char* baoracle_last_error(void);
void baoracle_default_options(bagpu_options* opt);
int baoracle_set_options(baoracle_handle h, const bagpu_options* opt);
int baoracle_set_cameras(baoracle_handle h, uint32_t n, const double* params4, const double* t_vs7);
int baoracle_set_poses(baoracle_handle h, uint32_t n, const double* t_wp7, const double* v_w3, const double* b6,
                       const uint8_t* is_active, const uint8_t* mask_used, const uint8_t* mask);
int baoracle_set_landmarks(baoracle_handle h, uint32_t n, const double* x_w4, const uint32_t* ref_pose_id,
                           const uint32_t* ref_cam_id, const uint8_t* is_active);
int baoracle_set_proj_residuals(baoracle_handle h, uint32_t n, const double* z2, const uint32_t* meas_pose_id,
                                const uint32_t* landmark_id, const uint32_t* cam_id, const double* orig_weight);
int baoracle_set_unary_residuals(baoracle_handle h, uint32_t n, const uint32_t* pose_id, const double* t_wp7,
                                 const double* cov_inv36, const uint8_t* use_rotation);
int baoracle_set_binary_residuals(baoracle_handle h, uint32_t n, const uint32_t* pose1_id, const uint32_t* pose2_id,
                                  const double* t_12_7, const double* cov_inv36, const double* weight,
                                  const uint8_t* use_rotation);
int baoracle_set_imu_residuals(baoracle_handle h, uint32_t n, const uint32_t* pose1_id, const uint32_t* pose2_id,
                               const uint32_t* meas_start, const double* meas7, const double* weight);
int baoracle_solve(baoracle_handle h, uint32_t max_iter, double gn_damping, int error_increase_allowed, bagpu_summary* out);
int baoracle_build(baoracle_handle h, float* ms_out);
int baoracle_get_poses(baoracle_handle h, double* t_wp7, double* v_w3, double* b6, uint8_t* mask_used, uint8_t* mask);
int baoracle_get_landmarks(baoracle_handle h, double* x_w4, double* x_s4, uint8_t* is_reliable, uint32_t* num_outlier_residuals);
int baoracle_get_proj_residuals(baoracle_handle h, double* residual2, double* weight, double* maha);
int baoracle_get_other_residuals(baoracle_handle h, double* un_residual6, double* un_maha, double* un_cov_inv36, double* bin_residual6,
                                 double* bin_maha, double* imu_residualD, double* imu_maha, double* imu_cov_invDD);
int baoracle_get_trace(baoracle_handle h, uint32_t max_records, bagpu_iter_trace* out, uint32_t* n_out);
int baoracle_get_deltas(baoracle_handle h, uint32_t iter, double* delta_p, double* delta_l);
int baoracle_get_pattern(baoracle_handle h, int which, uint32_t* ncols, uint32_t* nnzb, uint32_t* colptr, uint32_t* rowidx);
int baoracle_get_system(baoracle_handle h, double* s_dense, double* rhs_sc, double* rhs_p, double* rhs_l, double* v_inv);
/* s.txt rhs.txt j_pr.txt r_pr.txt j_l.txt of the last build, reference src/BundleAdjuster.cpp:600-620 */
int baoracle_dump(baoracle_handle h, const char* dir);
int baoracle_sync_state(baoracle_handle h);
int baoracle_get_phase_ms(baoracle_handle h, float* ms, uint64_t* kernel_launches);
#ifdef __cplusplus
}
#endif
#endif
