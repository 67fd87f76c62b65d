// ba_gpu.h — internal device-side data layout of the B200 BundleAdjuster path.
//
// HBM layout (all structure-of-arrays unless noted, doubles unless noted):
//   poses        q[4P] t[3P] v[3P] b[6P], opt id (int32, -1 inactive), free-parameter bit mask
//   pose frames  per (pose, camera): T_sw = [R_sw(9) t_sw(3)], T_ws = [R_ws(9) t_ws(3)], plus R_wp(9)
//                (replaces the PoseT::t_sw cache, reference include/ba/Types.h:59-70)
//   landmarks    x[4N] (x_s for LmSize 1, x_w for LmSize 3), ref pose/cam, opt id, reliable flag
//   observations sorted by landmark id, ties by residual id: z[2] (double2), meas pose (u32),
//                cam (u32), orig weight; CSR lm_obs_start[N+1]; perm[] back to residual ids
//   per-obs      maha0 (median input), weight (Huber)
//   linearisation (rewritten by k_linearize every build).  With B = d project / d X_world (2 x 3) every projection
//   Jacobian factors as  d z / d x_pose = B K_pose,  d z / d landmark = -B (LmSize 3) or -B u (LmSize 1, u = baseline
//   between the reference and the measuring sensor),  d z / d x_ref = -B K_ref,  K_p(X, rho) = [rho I | -[X - rho t_wp]x] D_p,
//   D_p = blkdiag(I, R_wp) with the masked parameter columns zeroed.  So an observation enters EVERY block of the
//   reduced system only through
//     per observation  G = w B^T B (3 x 3 symmetric, 6 doubles in a 64-byte record, obs_G) and h = w B^T r (32-byte record, obs_h)
//                      LmSize 1 also gv = [G u (3) | V^-1 of the landmark] (32-byte record, obs_gv): all an off-diagonal
//                      block needs of an observation, since there  M_e = -V^-1 (G_a u_a)(G_b u_b)^T
//     per landmark     LmRec (128 bytes): LmSize 3: X (3) | rho | V^-1 upper (6) | dhat = V^-1 b_l (3)
//                                         LmSize 1: X (3) | rho | t_ws of the reference sensor (3) | V^-1 | dhat | gbar = sum G u (3)
//                      LmSize 1 only, LmRecC (96 bytes): Gbar = sum G (6) | hbar = sum h (3), over the observations the
//                      reference pose takes part in
//   and each S block is a sum of  K~_i^T M_e K~_j  over its list entries, M_e a 3 x 3 built from those records
//   (schur_kernels.cu); D_i^T . D_j is applied once per block.  No Jacobian is stored and none is re-evaluated.
//   pose-major   pose_obs: the observation indices (landmark-major positions) of every optimised pose, and for
//                LmSize 1 pose_lms: the landmarks referenced to every optimised pose.
//   reduced S    upper block-tile storage: tiles of T x T doubles (T = 8 poses x PoseSize),
//                column-major inside a tile, tile_map[tI * nT + tJ] -> tile slot or -1
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/ba_b200.h"

namespace bagpu {

constexpr int kTilePoses = 8;   // poses per S tile
constexpr uint32_t kNoSlot = 0xffffffffu;
constexpr int kLmRec = 16, kLmRecC = 12;   // doubles per landmark record (128 B) / LmSize-1 extension (96 B)
constexpr int kObsG = 8, kObsH = 4;         // doubles per observation record: G (6 + 2 pad, 64 B), h (3 + 1 pad, 32 B)
constexpr int kObsGv = 4;                   // LmSize 1: G u (3) | V^-1 (32 B)

struct Scalars {       // device-resident scalar block (one per handle)
  double median;       // nth_element value of maha0
  double proj_error;   // sum maha with Huber weights (BuildProblem)
  double eval_proj, eval_binary, eval_unary, eval_inertial;
  double binary_error, unary_error, inertial_error;
  double sd_den_proj, sd_den_other;
  double norm2_bp, norm2_bl;
  double dot[8];       // scratch for vector reductions
  double unary_median;
  int chol_fail;       // first non-positive pivot (tile column + 1) or 0
  int pad;
};

struct DeviceProblem {
  // sizes
  int L, D;                 // LmSize, PoseSize
  int n_cams, n_poses, n_active_poses, n_lms, n_active_lms, n_obs;
  int n_unary, n_binary, n_imu;
  int cam_identity;          // every camera has T_vs = identity (the usual case: one branch skips the extrinsics products)
  // cameras
  const double* cam_params;  // 4 per cam
  const double* cam_Rvs;     // 9 per cam
  const double* cam_tvs;     // 3 per cam
  // poses
  double *pose_q, *pose_t, *pose_v, *pose_b;
  const int* pose_opt;
  const uint32_t* pose_free;   // bit k set = parameter k free (active poses)
  double* Tsw;   // [P * n_cams * 12]
  double* Tws;   // [P * n_cams * 12]
  double* Rwp;   // [P * 9]
  double* kd_pack;   // [P * 12] per-pose operands of K_pose delta for the current delta (k_pose_delta_pack)
  // landmarks
  double* lm_x;             // 4 per lm: x_s (L=1) or x_w (L=3)
  const uint32_t* lm_ref_pose;
  const uint32_t* lm_ref_cam;
  const int* lm_opt;        // -1 inactive
  uint8_t* lm_reliable;
  uint32_t* lm_outliers;
  const uint32_t* lm_obs_start;  // n_lms + 1
  // observations (landmark-sorted)
  const double2* obs_z;
  const uint32_t* obs_pose;
  const uint32_t* obs_cam;   // may be null when n_cams == 1
  const double* obs_w0;
  const uint32_t* obs_perm;  // sorted index -> residual id
  const uint32_t* obs_lm;    // sorted index -> landmark id (non-decreasing)
  double* obs_maha0;
  double* obs_w;
  double* obs_G;             // kObsG doubles per observation: w B^T B upper (00 01 02 11 12 22)
  double* obs_h;             // kObsH doubles per observation: w B^T r
  double* obs_gv;            // LmSize 1, kObsGv doubles per observation: G u | V^-1 of its landmark
  const uint32_t* pose_obs;  // observation indices grouped by optimised measurement pose (pose-major), n_pose_obs entries
  const uint32_t* pose_obs_lm;  // landmark of pose_obs[k]
  const uint32_t* pose_lms;  // L == 1: landmark ids grouped by optimised reference pose
  const int* opt_to_pose;    // opt id -> pose id
  // per landmark outputs
  double* lm_bl;             // L  (rhs_l_, indexed by landmark id; compacted on read-back)
  double* lm_rec;            // kLmRec doubles per landmark (layout above)
  double* lm_recC;           // L == 1: kLmRecC doubles per landmark
  double* red_partial;       // per-CTA partial sums of the scalar reductions (summed in index order by the last CTA)
  unsigned int* red_counter; // arrival counter of the running reduction (self-resetting)
  // reduced system
  double* S_tiles;
  const int* tile_map;       // [nT * nT] in PERMUTED tile coordinates -> slot or -1
  const int* tile_perm;      // natural tile (opt id / 8) -> position in the elimination order (nested dissection)
  int nT, T;
  double* b_p;               // rhs_p_   [n_active_poses * D] padded to nT*T
  double* rhs_sc;            // rhs_p_sc [nT*T]
  Scalars* sc;
  // options
  double c_huber_factor;     // 1.2107
  int use_robust;
  double outlier_threshold;
};

// S element address helpers -------------------------------------------------------------
__host__ __device__ inline long long tile_elem_offset(int T, int slot, int r, int c) {
  return (long long)slot * T * T + (long long)c * T + r;
}

}  // namespace bagpu
