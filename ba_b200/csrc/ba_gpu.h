// ba_gpu.h — internal device-side data layout of the B200 BundleAdjuster path.
//
// HBM layout (all structure-of-arrays, doubles unless noted):
//   poses        q[4P] t[3P] v[3P] b[6P], opt id (int32, -1 inactive), free-parameter bit mask
//   pose frames  per (pose, camera): T_sw = [R_sw(9) t_sw(3)], T_ws = [R_ws(9) t_ws(3)], plus R_wp(9)
//                (replaces the PoseT::t_sw cache, reference include/ba/Types.h:59-70)
//   landmarks    x[4N] (x_s for LmSize 1, x_w for LmSize 3), ref pose/cam, opt id, reliable flag
//   observations sorted by landmark id, ties by residual id: z[2] (double2), meas pose (u32),
//                cam (u32), orig weight; CSR lm_obs_start[N+1]; perm[] back to residual ids
//   per-obs      maha0 (median input), weight (Huber)
//   records      one per observation whose measurement pose is optimised, stored POSE-MAJOR
//                (slot = obs_slot[obs]; pose p owns slots pose_rec_start[p] .. [p+1]):
//                  LmSize 3 (28 doubles): C = Bt V^-1 (2x3) | At (2x6) | Bt (2x3) | rt (2) | rho (2)
//                  LmSize 1 (40 doubles): E = W V^-1 (6) | W = At^T Bt (6) | At (2x6) | Q (2x6) | rt (2) | rho (2)
//                with At = sqrt(w) J_meas, Bt = sqrt(w) J_lm, rt = sqrt(w) r, rho = rt - Bt V^-1 b_l,
//                Q = At_ref - Bt V^-1 W_ref^T (the reference-pose Jacobian with the landmark projected out)
//   per-landmark V^-1 (L x L), b_l (L), dhat = V^-1 b_l, W_ref (6, LmSize 1 only), and for LmSize 1 a
//                reference-pose record (36 doubles, stored by reference pose): H_rr - W_r V^-1 W_r^T
//                (upper, 21) | g_r (6) | g_r - W_r dhat (6)
//   reduced S    upper block-tile storage: tiles of T x T doubles (T = 8 poses x PoseSize),
//                column-major inside a tile, tile_map[tI * nT + tJ] -> tile slot or -1
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/ba_b200.h"

namespace bagpu {

constexpr int kTilePoses = 8;   // poses per S tile
constexpr uint32_t kNoSlot = 0xffffffffu;
constexpr int kRec3 = 28, kRec1 = 40, kLmRec = 36;

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
  const uint32_t* obs_slot;  // pose-major record slot of the observation, kNoSlot if its pose takes no part
  double* rec;               // records, rec_stride doubles each
  int rec_stride;            // 28 (LmSize 3) or 40 (LmSize 1)
  // per landmark outputs
  double* lm_Vinv;           // L*L
  double* lm_bl;             // L  (rhs_l_, indexed by landmark id; compacted on read-back)
  double* lm_dhat;           // L  V^-1 b_l
  double* lm_Wref;           // 6 (L == 1)
  const uint32_t* lm_rslot;  // L == 1: slot of the landmark's reference-pose record, kNoSlot if none
  double* lm_rec;            // L == 1: 36 doubles per slot
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
