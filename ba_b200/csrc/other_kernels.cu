// other_kernels.cu — unary, binary and inertial residuals of the hot path (FP64).
//
// Replaces, per outer iteration:
//   binary residuals   reference src/BundleAdjuster.cpp:1392-1429 (+ J insertion :1646-1669, U/b_p :357-371)
//   unary residuals    reference src/BundleAdjuster.cpp:1431-1484 (+ :1671-1690, :374-386)
//   inertial residuals reference include/ba/parallel_algos.h:178-358, src/BundleAdjuster.cpp:1486-1541
//                      (+ :1692-1726, :389-401), integrator reference include/ba/Types.h:324-738
// and their parts of EvaluateResiduals (src/BundleAdjuster.cpp:190-255) and of the
// steepest-descent denominator (src/BundleAdjuster.cpp:889-899).
//
// The reference whitens residuals and Jacobians with sqrtm(cov_inv); only the products
// J^T cov_inv J, J^T cov_inv r and r^T cov_inv r enter the normal equations, so the GPU path
// uses cov_inv directly and never forms a matrix square root.  The SE3 Jacobians are the
// closed forms of the reference's quaternion chains (inverse SO3 Jacobians), see DESIGN.md §3.
#include "ba_gpu.h"
#include "device_math.cuh"
#include "launch.h"
#include "proj_kernels.cuh"
#include "imu_device.cuh"

namespace bagpu {

BA_D void load_pose(const DeviceProblem& P, int p, double* q, double* t) {
  q[0] = P.pose_q[4 * (size_t)p]; q[1] = P.pose_q[4 * (size_t)p + 1]; q[2] = P.pose_q[4 * (size_t)p + 2]; q[3] = P.pose_q[4 * (size_t)p + 3];
  t[0] = P.pose_t[3 * (size_t)p]; t[1] = P.pose_t[3 * (size_t)p + 1]; t[2] = P.pose_t[3 * (size_t)p + 2];
}

BA_D void mask_cols6(double* J /*6x6 row-major*/, uint32_t fr) {
  if ((fr & 63u) == 63u) return;
  for (int k = 0; k < 6; ++k)
    if (!((fr >> k) & 1u)) for (int r = 0; r < 6; ++r) J[6 * r + k] = 0.0;
}

// U(i,j) += alpha * Ja^T C Jb into the top-left 6x6 of the D x D pose block (i <= j); upper only if i == j
BA_D void add_jtcj(const DeviceProblem& P, int i, int j, const double* Ja, const double* C, const double* Jb, double alpha) {
  double CJ[36];
  for (int r = 0; r < 6; ++r)
    for (int c = 0; c < 6; ++c) {
      double s = 0.0;
      for (int k = 0; k < 6; ++k) s += C[6 * r + k] * Jb[6 * k + c];
      CJ[6 * r + c] = s;
    }
  const SBlock blk = s_block(P, i, j);
  for (int c = 0; c < 6; ++c)
    for (int r = 0; r < 6; ++r) {
      if (i == j && r > c) continue;
      double s = 0.0;
      for (int k = 0; k < 6; ++k) s += Ja[6 * k + r] * CJ[6 * k + c];
      red_add(blk.p + (size_t)r * blk.rs + (size_t)c * blk.cs, alpha * s);
    }
}

// ---- binary -------------------------------------------------------------------------------------
struct BinEval { double r[6]; double J1[36]; double J2[36]; int o1, o2; };

BA_D void binary_eval(const DeviceProblem& P, const OtherResiduals& R, int i, BinEval& e, bool jac) {
  const int p1 = (int)R.bin_p1[i], p2 = (int)R.bin_p2[i];
  double q1[4], t1[3], q2[4], t2[3];
  load_pose(P, p1, q1, t1); load_pose(P, p2, q2, t2);
  const double* tm = R.bin_t12 + 7 * (size_t)i;
  const double q1c[4] = {-q1[0], -q1[1], -q1[2], q1[3]};
  double R1[9];
  quat_to_R(q1, R1);
  const double dt[3] = {t2[0] - t1[0], t2[1] - t1[1], t2[2] - t1[2]};
  double t12[3];
  mat3T_vec(R1, dt, t12);
  double q12[4];
  quat_mul(q1c, q2, q12); quat_normalize(q12);
  const double qmc[4] = {-tm[0], -tm[1], -tm[2], tm[3]};
  double qQ[4];
  quat_mul(q12, qmc, qQ); quat_normalize(qQ);
  double phi[3];
  so3_log(qQ, phi);
  e.r[0] = t12[0] - tm[4]; e.r[1] = t12[1] - tm[5]; e.r[2] = t12[2] - tm[6];
  e.r[3] = phi[0]; e.r[4] = phi[1]; e.r[5] = phi[2];
  const bool rot = R.bin_use_rot[i] != 0;
  if (!rot) { e.r[3] = e.r[4] = e.r[5] = 0.0; }
  e.o1 = P.pose_opt[p1]; e.o2 = P.pose_opt[p2];
  if (!jac) return;
  for (int k = 0; k < 36; ++k) { e.J1[k] = 0.0; e.J2[k] = 0.0; }
  // d r / d x1 = [ -R1^T  [t12]x ; 0  -Jl^-1(phi) ],  d r / d x2 = [ R1^T  0 ; 0  Jr^-1(phi) R_m ]
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) { e.J1[6 * r + c] = -R1[3 * c + r]; e.J2[6 * r + c] = R1[3 * c + r]; }
  e.J1[6 * 0 + 4] = -t12[2]; e.J1[6 * 0 + 5] = t12[1];
  e.J1[6 * 1 + 3] = t12[2];  e.J1[6 * 1 + 5] = -t12[0];
  e.J1[6 * 2 + 3] = -t12[1]; e.J1[6 * 2 + 4] = t12[0];
  if (rot) {
    double Jl[9], Jr[9], Rm[9], JrRm[9];
    so3_jinv(phi, -0.5, Jl);
    so3_jinv(phi, 0.5, Jr);
    quat_to_R(tm, Rm);
    mat3_mul(Jr, Rm, JrRm);
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) { e.J1[6 * (3 + r) + 3 + c] = -Jl[3 * r + c]; e.J2[6 * (3 + r) + 3 + c] = JrRm[3 * r + c]; }
  }
  if (e.o1 >= 0) mask_cols6(e.J1, P.pose_free[p1]);
  if (e.o2 >= 0) mask_cols6(e.J2, P.pose_free[p2]);
}

// Scalar sums of these kernels (errors, evaluation sums, the steepest-descent denominator) go through grid_sum_fixed:
// fixed summation order, so every rank and every run gets the same bits (the step control branches on them).
__global__ void __launch_bounds__(64) k_binary_build(DeviceProblem P, OtherResiduals R) {
  double acc[1] = {0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < R.n_binary; i += gridDim.x * blockDim.x) {
  BinEval e;
  binary_eval(P, R, i, e, true);
  const double* C = R.bin_cov_inv + 36 * (size_t)i;
  const double wgt = R.bin_weight[i];
  double Cr[6];
  for (int r = 0; r < 6; ++r) { double s = 0.0; for (int k = 0; k < 6; ++k) s += C[6 * r + k] * e.r[k]; Cr[r] = s; }
  // binary_error_ += (r_w^T cov_inv r_w) * weight with r_w = sqrtm(cov_inv) r  (quirk Q7) = |cov_inv r|^2
  double m = 0.0;
  for (int r = 0; r < 6; ++r) m += Cr[r] * Cr[r];
  acc[0] += m * wgt;
  if (e.o1 >= 0) {
    add_jtcj(P, e.o1, e.o1, e.J1, C, e.J1, wgt);
    for (int c = 0; c < 6; ++c) { double s = 0.0; for (int k = 0; k < 6; ++k) s += e.J1[6 * k + c] * Cr[k]; red_add(P.b_p + (size_t)e.o1 * P.D + c, wgt * s); }
  }
  if (e.o2 >= 0) {
    add_jtcj(P, e.o2, e.o2, e.J2, C, e.J2, wgt);
    for (int c = 0; c < 6; ++c) { double s = 0.0; for (int k = 0; k < 6; ++k) s += e.J2[6 * k + c] * Cr[k]; red_add(P.b_p + (size_t)e.o2 * P.D + c, wgt * s); }
  }
  if (e.o1 >= 0 && e.o2 >= 0 && e.o1 != e.o2) {
    if (e.o1 < e.o2) add_jtcj(P, e.o1, e.o2, e.J1, C, e.J2, wgt);
    else add_jtcj(P, e.o2, e.o1, e.J2, C, e.J1, wgt);
  }
  }
  grid_sum_fixed<1>(acc, P.red_partial, P.red_counter, &P.sc->binary_error, true);
}

__global__ void __launch_bounds__(64) k_binary_sd(DeviceProblem P, OtherResiduals R) {
  double acc[1] = {0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < R.n_binary; i += gridDim.x * blockDim.x) {
    BinEval e;
    binary_eval(P, R, i, e, true);
    double v[6] = {0, 0, 0, 0, 0, 0};
    if (e.o1 >= 0) for (int r = 0; r < 6; ++r) for (int k = 0; k < 6; ++k) v[r] += e.J1[6 * r + k] * P.b_p[(size_t)e.o1 * P.D + k];
    if (e.o2 >= 0) for (int r = 0; r < 6; ++r) for (int k = 0; k < 6; ++k) v[r] += e.J2[6 * r + k] * P.b_p[(size_t)e.o2 * P.D + k];
    const double* C = R.bin_cov_inv + 36 * (size_t)i;
    double s = 0.0;
    for (int r = 0; r < 6; ++r) for (int k = 0; k < 6; ++k) s += v[r] * C[6 * r + k] * v[k];
    acc[0] += s;
  }
  grid_sum_fixed<1>(acc, P.red_partial, P.red_counter, &P.sc->sd_den_other, true);
}

__global__ void __launch_bounds__(64) k_binary_eval(DeviceProblem P, OtherResiduals R) {
  double acc[1] = {0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < R.n_binary; i += gridDim.x * blockDim.x) {
    BinEval e;
    binary_eval(P, R, i, e, false);
    double s = 0.0;
    for (int r = 0; r < 6; ++r) s += e.r[r] * e.r[r];
    acc[0] += s * R.bin_weight[i];
    for (int r = 0; r < 6; ++r) R.bin_out[7 * (size_t)i + r] = e.r[r];
    R.bin_out[7 * (size_t)i + 6] = s * R.bin_weight[i];
  }
  grid_sum_fixed<1>(acc, P.red_partial, P.red_counter, &P.sc->eval_binary, true);
}

// ---- unary --------------------------------------------------------------------------------------
struct UnEval { double r[6]; double J[36]; int o; };

BA_D void unary_eval(const DeviceProblem& P, const OtherResiduals& R, int i, UnEval& e, bool jac) {
  const int p = (int)R.un_pose[i];
  double q[4], t[3];
  load_pose(P, p, q, t);
  const double* tp = R.un_t + 7 * (size_t)i;
  const double qpc[4] = {-tp[0], -tp[1], -tp[2], tp[3]};
  double qQ[4];
  quat_mul(q, qpc, qQ); quat_normalize(qQ);
  double phi[3];
  so3_log(qQ, phi);
  e.r[0] = t[0] - tp[4]; e.r[1] = t[1] - tp[5]; e.r[2] = t[2] - tp[6];
  e.r[3] = phi[0]; e.r[4] = phi[1]; e.r[5] = phi[2];
  const bool rot = R.un_use_rot[i] != 0;
  if (!rot) { e.r[3] = e.r[4] = e.r[5] = 0.0; }
  e.o = P.pose_opt[p];
  if (!jac) return;
  for (int k = 0; k < 36; ++k) e.J[k] = 0.0;
  e.J[0] = e.J[7] = e.J[14] = 1.0;
  if (rot) {
    double Jr[9], Rp[9], JrRp[9];
    so3_jinv(phi, 0.5, Jr);
    quat_to_R(tp, Rp);
    mat3_mul(Jr, Rp, JrRp);
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) e.J[6 * (3 + r) + 3 + c] = JrRp[3 * r + c];
  }
  if (e.o >= 0) mask_cols6(e.J, P.pose_free[p]);
}

__global__ void k_unary_maha(DeviceProblem P, OtherResiduals R) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= R.n_unary) return;
  UnEval e;
  unary_eval(P, R, i, e, false);
  const double* C = R.un_cov_inv + 36 * (size_t)i;
  double s = 0.0;
  for (int r = 0; r < 6; ++r) for (int k = 0; k < 6; ++k) s += e.r[r] * C[6 * r + k] * e.r[k];
  R.un_maha[i] = s;
}

__global__ void __launch_bounds__(64) k_unary_build(DeviceProblem P, OtherResiduals R) {
  double acc[1] = {0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < R.n_unary; i += gridDim.x * blockDim.x) {
    UnEval e;
    unary_eval(P, R, i, e, true);
    const double c_huber = 1.2107 * sqrt(P.sc->unary_median);
    const double em = sqrt(R.un_maha[i]);
    const double weight = (em > c_huber) ? c_huber / em : 1.0;
    double* C = R.un_cov_inv + 36 * (size_t)i;
    for (int k = 0; k < 36; ++k) C[k] *= weight;   // persists across iterations (quirk Q6)
    double Cr[6];
    for (int r = 0; r < 6; ++r) { double s = 0.0; for (int k = 0; k < 6; ++k) s += C[6 * r + k] * e.r[k]; Cr[r] = s; }
    double m = 0.0;
    for (int r = 0; r < 6; ++r) m += e.r[r] * Cr[r];
    acc[0] += m;
    if (e.o >= 0) {
      add_jtcj(P, e.o, e.o, e.J, C, e.J, 1.0);
      for (int c = 0; c < 6; ++c) { double s = 0.0; for (int k = 0; k < 6; ++k) s += e.J[6 * k + c] * Cr[k]; red_add(P.b_p + (size_t)e.o * P.D + c, s); }
    }
  }
  grid_sum_fixed<1>(acc, P.red_partial, P.red_counter, &P.sc->unary_error, true);
}

__global__ void __launch_bounds__(64) k_unary_sd(DeviceProblem P, OtherResiduals R) {
  double acc[1] = {0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < R.n_unary; i += gridDim.x * blockDim.x) {
    UnEval e;
    unary_eval(P, R, i, e, true);
    if (e.o < 0) continue;
    double v[6] = {0, 0, 0, 0, 0, 0};
    for (int r = 0; r < 6; ++r) for (int k = 0; k < 6; ++k) v[r] += e.J[6 * r + k] * P.b_p[(size_t)e.o * P.D + k];
    const double* C = R.un_cov_inv + 36 * (size_t)i;
    double s = 0.0;
    for (int r = 0; r < 6; ++r) for (int k = 0; k < 6; ++k) s += v[r] * C[6 * r + k] * v[k];
    acc[0] += s;
  }
  grid_sum_fixed<1>(acc, P.red_partial, P.red_counter, &P.sc->sd_den_other, true);
}

__global__ void __launch_bounds__(64) k_unary_evalk(DeviceProblem P, OtherResiduals R) {
  double acc[1] = {0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < R.n_unary; i += gridDim.x * blockDim.x) {
    UnEval e;
    unary_eval(P, R, i, e, false);
    const double* C = R.un_cov_inv + 36 * (size_t)i;
    double s = 0.0;
    for (int r = 0; r < 6; ++r) for (int k = 0; k < 6; ++k) s += e.r[r] * C[6 * r + k] * e.r[k];
    acc[0] += s;
    for (int r = 0; r < 6; ++r) R.un_out[7 * (size_t)i + r] = e.r[r];
    R.un_out[7 * (size_t)i + 6] = s;
  }
  grid_sum_fixed<1>(acc, P.red_partial, P.red_counter, &P.sc->eval_unary, true);
}

// ---- inertial -----------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(64) k_imu_build(DeviceProblem P, OtherResiduals R) {
  double acc[1] = {0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < R.n_imu; i += gridDim.x * blockDim.x) acc[0] += imu_build_one<D>(P, R, i);
  grid_sum_fixed<1>(acc, P.red_partial, P.red_counter, &P.sc->inertial_error, true);
}
template <int D>
__global__ void __launch_bounds__(64) k_imu_sd(DeviceProblem P, OtherResiduals R) {
  double acc[1] = {0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < R.n_imu; i += gridDim.x * blockDim.x) acc[0] += imu_sd_one<D>(P, R, i);
  grid_sum_fixed<1>(acc, P.red_partial, P.red_counter, &P.sc->sd_den_other, true);
}
template <int D>
__global__ void __launch_bounds__(64) k_imu_eval(DeviceProblem P, OtherResiduals R) {
  double acc[1] = {0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < R.n_imu; i += gridDim.x * blockDim.x) acc[0] += imu_eval_one<D>(P, R, i);
  grid_sum_fixed<1>(acc, P.red_partial, P.red_counter, &P.sc->eval_inertial, true);
}

// ---- launchers ----------------------------------------------------------------------------------
static inline int og(int n) { const int b = (n + 63) / 64; return b > kRedMaxGrid ? kRedMaxGrid : (b < 1 ? 1 : b); }   // grid_sum_fixed: at most kRedMaxGrid CTAs
void launch_unary_maha(const DeviceProblem& P, const OtherResiduals& R, cudaStream_t st) {
  if (R.n_unary) k_unary_maha<<<(R.n_unary + 63) / 64, 64, 0, st>>>(P, R);
}
void launch_other_build(const DeviceProblem& P, const OtherResiduals& R, cudaStream_t st, uint64_t* launches) {
  if (R.n_binary) { k_binary_build<<<og(R.n_binary), 64, 0, st>>>(P, R); ++*launches; }
  if (R.n_unary) { k_unary_build<<<og(R.n_unary), 64, 0, st>>>(P, R); ++*launches; }
  if (R.n_imu) {
    if (P.D == 9) k_imu_build<9><<<og(R.n_imu), 64, 0, st>>>(P, R);
    else if (P.D == 15) k_imu_build<15><<<og(R.n_imu), 64, 0, st>>>(P, R);
    ++*launches;
  }
}
void launch_other_sd(const DeviceProblem& P, const OtherResiduals& R, cudaStream_t st, uint64_t* launches) {
  if (R.n_binary) { k_binary_sd<<<og(R.n_binary), 64, 0, st>>>(P, R); ++*launches; }
  if (R.n_unary) { k_unary_sd<<<og(R.n_unary), 64, 0, st>>>(P, R); ++*launches; }
  if (R.n_imu) {
    if (P.D == 9) k_imu_sd<9><<<og(R.n_imu), 64, 0, st>>>(P, R);
    else if (P.D == 15) k_imu_sd<15><<<og(R.n_imu), 64, 0, st>>>(P, R);
    ++*launches;
  }
}
void launch_other_evaluate(const DeviceProblem& P, const OtherResiduals& R, cudaStream_t st, uint64_t* launches) {
  if (R.n_binary) { k_binary_eval<<<og(R.n_binary), 64, 0, st>>>(P, R); ++*launches; }
  if (R.n_unary) { k_unary_evalk<<<og(R.n_unary), 64, 0, st>>>(P, R); ++*launches; }
  if (R.n_imu) {
    if (P.D == 9) k_imu_eval<9><<<og(R.n_imu), 64, 0, st>>>(P, R);
    else if (P.D == 15) k_imu_eval<15><<<og(R.n_imu), 64, 0, st>>>(P, R);
    ++*launches;
  }
}

}  // namespace bagpu
