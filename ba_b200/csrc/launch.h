// launch.h — host-side launchers of the CUDA kernels (one translation unit per kernel family).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "ba_gpu.h"

namespace bagpu {

struct SelectState {
  unsigned int hist[4096];
  unsigned long long prefix, k_rem, prefix_next, k_rem_next;
};

// unary / binary / inertial residual records in HBM (SoA)
struct OtherResiduals {
  int n_unary, n_binary, n_imu;
  // unary (reference Types.h:255-266)
  const uint32_t* un_pose; const double* un_t;   // 7 per residual
  double* un_cov_inv;                            // 36, multiplied by the Huber weight every build (Q6)
  const uint8_t* un_use_rot;
  double* un_maha;                               // scratch: r^T cov_inv r
  // binary (Types.h:268-280)
  const uint32_t *bin_p1, *bin_p2; const double* bin_t12; const double* bin_cov_inv; const double* bin_weight;
  const uint8_t* bin_use_rot;
  // inertial (Types.h:300-321)
  const uint32_t *imu_p1, *imu_p2, *imu_meas_start; const double* imu_meas; const double* imu_weight;
  double* imu_cov_inv;      // D*D per residual, of the last build (weighted)
  double* imu_J1; double* imu_J2;   // D*D per residual (row-major), kept for the SD denominator
  double* imu_res;          // D per residual
  double gravity[3];
  double imu_r[6];          // sensor noise diagonal (imu_.r)
  double imu_rb[6];         // bias random walk (imu_.r_b)
};

void set_sm_count(int n);
void launch_pose_frames(const DeviceProblem& P, cudaStream_t st);
void launch_prepare_xs(const DeviceProblem& P, const double* lm_xw, cudaStream_t st);
void launch_finish_xw(const DeviceProblem& P, double* lm_xw, cudaStream_t st);
void launch_residual(const DeviceProblem& P, int G, cudaStream_t st);
void launch_accumulate(const DeviceProblem& P, int G, cudaStream_t st);
void launch_schur(const DeviceProblem& P, cudaStream_t st);
void launch_sd_denominator(const DeviceProblem& P, int G, cudaStream_t st);
void launch_back_substitute(const DeviceProblem& P, int G, const double* delta_p, double* delta_l, cudaStream_t st);
void launch_evaluate(const DeviceProblem& P, int G, double* out_sum, double* res_out, double* maha_out, cudaStream_t st);
void launch_update_poses(const DeviceProblem& P, const double* delta_p, double coef, cudaStream_t st);
void launch_update_landmarks(const DeviceProblem& P, const double* delta_l, double coef, cudaStream_t st);
void launch_mask_diagonal(const DeviceProblem& P, const uint8_t* mask_used, int global_flag, cudaStream_t st);
void launch_dot(const double* a, const double* b, size_t n, double* out, cudaStream_t st);
void launch_axpby(double* out, double alpha, const double* x, double beta, const double* y, size_t n, cudaStream_t st);
void launch_compact_bl(const DeviceProblem& P, double* out, cudaStream_t st);

int select_num_passes();
void launch_select_init(SelectState* st, unsigned long long k, cudaStream_t s);
void launch_select_hist(const double* x, size_t n, SelectState* st, int pass, int sms, cudaStream_t s);
void launch_select_scan(SelectState* st, int pass, double* out, cudaStream_t s);

void launch_cholesky_solve(int T, double* tiles, double* uinv, double* work /* 2*nT*T */, double* rhs, const int* tile_map,
                           const int* row_ptr_dev, const int* row_col_dev, const int* row_ptr_host, int nT, int* fail,
                           cudaStream_t st, uint64_t* launches);

// unary / binary / inertial residuals (other_kernels.cu)
void launch_unary_maha(const DeviceProblem& P, const OtherResiduals& R, cudaStream_t st);
void launch_other_build(const DeviceProblem& P, const OtherResiduals& R, cudaStream_t st, uint64_t* launches);
void launch_other_sd(const DeviceProblem& P, const OtherResiduals& R, cudaStream_t st, uint64_t* launches);
void launch_other_evaluate(const DeviceProblem& P, const OtherResiduals& R, cudaStream_t st, uint64_t* launches);

}  // namespace bagpu
