// TEST INFRASTRUCTURE — CPU oracle for the B200 BundleAdjuster path.  PARITY UNPINNED
// (see minimath.h: the reference cannot be built here and has no golden vectors).
//
// Literal single-threaded restatement of the reference optimisation loop:
//   Solve              reference src/BundleAdjuster.cpp:278-705
//   BuildProblem       reference src/BundleAdjuster.cpp:1166-1803
//   projection functor reference include/ba/parallel_algos.h:35-152  (called serially, cpp:1347)
//   inertial functor   reference include/ba/parallel_algos.h:178-358
//   SolveInternal      reference src/BundleAdjuster.cpp:838-1161
//   CalculateGn        reference src/BundleAdjuster.cpp:748-833
//   GetLandmarkDelta   reference src/BundleAdjuster.cpp:709-744
//   ApplyUpdate        reference src/BundleAdjuster.cpp:21-140
//   EvaluateResiduals  reference src/BundleAdjuster.cpp:144-274
// It keeps the reference's structure (AoS residual records with stored Jacobians,
// block-CSC J / J^T matrices, four sparse block products for the Schur complement,
// LDL^T of the reduced system) so that it can double as the timed CPU baseline.
// Deviations, all value-preserving: S is kept in upper skyline storage instead of a
// dense MatrixXt (block_sparse.h), CalibSize = 0 / DoTvs = false only, and the
// LmSize = 3 pose Jacobian reads lm.x_w where the reference reads the never
// initialised lm.x_s.
//
// Exposes the same C entry points as include/ba_b200.h with the prefix baoracle_.
#include "ba_types.h"
#include "block_sparse.h"
#include "ba_oracle.h"
#include <chrono>
#include <climits>
#include <cstdio>
#include <fstream>
#include <string>
#include <thread>

namespace orc {

static double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

struct OracleBase {
  virtual ~OracleBase() {}
  bagpu_options opt;
  bagpu_summary summary;
  std::vector<bagpu_iter_trace> trace;
  std::vector<VecX> rec_delta_p, rec_delta_l;
  double phase_s[BAGPU_PH_COUNT] = {0};
  virtual int set_cameras(uint32_t n, const double* p4, const double* tvs7) = 0;
  virtual int set_poses(uint32_t n, const double* t, const double* v, const double* b, const uint8_t* act, const uint8_t* mu, const uint8_t* m) = 0;
  virtual int set_landmarks(uint32_t n, const double* xw, const uint32_t* rp, const uint32_t* rc, const uint8_t* act) = 0;
  virtual int set_proj(uint32_t n, const double* z, const uint32_t* mp, const uint32_t* lm, const uint32_t* cam, const double* w) = 0;
  virtual int set_unary(uint32_t n, const uint32_t* pid, const double* t, const double* ci, const uint8_t* ur) = 0;
  virtual int set_binary(uint32_t n, const uint32_t* p1, const uint32_t* p2, const double* t, const double* ci, const double* w, const uint8_t* ur) = 0;
  virtual int set_imu(uint32_t n, const uint32_t* p1, const uint32_t* p2, const uint32_t* ms, const double* m7, const double* w) = 0;
  virtual void init_after_options() = 0;
  virtual void Solve(uint32_t it, double damping, bool allow) = 0;
  virtual void BuildOnly() = 0;
  virtual int get_poses(double* t, double* v, double* b, uint8_t* mu, uint8_t* m) = 0;
  virtual int get_landmarks(double* xw, double* xs, uint8_t* rel, uint32_t* nout) = 0;
  virtual int get_proj(double* r, double* w, double* m) = 0;
  virtual int get_other(double* ur, double* um, double* uc, double* br, double* bm, double* ir, double* im, double* ic) = 0;
  virtual int get_pattern(int which, uint32_t* ncols, uint32_t* nnzb, uint32_t* colptr, uint32_t* rowidx) = 0;
  virtual int get_system(double* s, double* rsc, double* rp, double* rl, double* vi) = 0;
  virtual int dump(const char* dir) = 0;
};

template <int LM, int D>
struct BundleAdjuster : OracleBase {
  static constexpr int kPrPoseDim = 6;
  static constexpr int kLmDim = LM;
  static constexpr int LMS = LM > 0 ? LM : 1;
  static constexpr int kPoseDim = D;
  static constexpr bool kVelInState = (D >= 9);
  static constexpr bool kBiasInState = (D >= 15);
  static constexpr int kTrustRegionAuto = -1;  // BundleAdjuster.h:30 (constexpr int = -1.0)
  typedef orc::Landmark<LM> LandmarkT;
  typedef orc::ProjectionResidual<LM> ProjRes;
  typedef orc::ImuResidual<D> ImuRes;
  struct Delta { VecX delta_p, delta_k, delta_l; };
  struct ImuCalibration { Vec3 b_g, b_a, g_vec; Diag6 r; Vec6 r_b; } imu_;

  std::vector<LinearCamera> rig_;
  std::vector<Pose> poses_;
  std::vector<LandmarkT> landmarks_;
  std::vector<ProjRes> proj_residuals_;
  std::vector<uint32_t> conditioning_proj_residuals_, conditioning_inertial_residuals_;
  std::vector<BinaryResidual> binary_residuals_;
  std::vector<UnaryResidual> unary_residuals_;
  std::vector<ImuRes> inertial_residuals_;
  std::vector<double> errors_;

  BlockMat<2, 6> j_pr_;  BlockMat<6, 2> jt_pr;
  BlockMat<2, LMS> j_l_;
  VecX r_pr_;
  BlockMat<6, D> j_pp_;  BlockMat<D, 6> jt_pp_;  VecX r_pp_;
  BlockMat<6, D> j_u_;   BlockMat<D, 6> jt_u_;   VecX r_u_;
  BlockMat<D, D> j_i_;   BlockMat<D, D> jt_i_;   VecX r_i_;
  BlockMat<D, D> u_;
  BlockMat<LMS, LMS> vi_;
  BlockMat<LMS, 6> jt_l_j_pr_;
  BlockMat<6, LMS> jt_pr_j_l_;
  BlockMat<6, 6> schur_term_;  // jt_pr_j_l_vi_jt_l_j_pr of the last build (kept for pattern queries)
  VecX rhs_p_, rhs_k_, rhs_l_, rhs_p_sc_;
  Skyline s_, s_factor_;
  double trust_region_size_ = -1;
  bool is_param_mask_used_ = false;
  double proj_error_ = 0, binary_error_ = 0, unary_error_ = 0, inertial_error_ = 0;
  double last_median_ = 0;
  uint32_t root_pose_id_ = 0, num_active_poses_ = 0, num_active_landmarks_ = 0;
  uint32_t binary_residual_offset_ = 0, unary_residual_offset_ = 0, proj_residual_offset = 0, inertial_residual_offset_ = 0;

  // ---- Init (BundleAdjuster.h:177-237) ------------------------------------------
  void init_after_options() override {
    trust_region_size_ = opt.trust_region_size;
    root_pose_id_ = opt.root_pose_id;
    imu_.r.setZero();
    for (int i = 0; i < 3; ++i) { imu_.r(i, i) = powi(opt.gyro_sigma, 2); imu_.r(3 + i, 3 + i) = powi(opt.accel_sigma, 2); }
    for (int i = 0; i < 3; ++i) { imu_.r_b[i] = powi(opt.gyro_bias_sigma, 2); imu_.r_b[3 + i] = powi(opt.accel_bias_sigma, 2); }
    for (int i = 0; i < 3; ++i) imu_.g_vec[i] = opt.gravity[i];
    memset(&summary, 0, sizeof(summary));
  }

  int set_cameras(uint32_t n, const double* p4, const double* tvs7) override {
    rig_.resize(n);
    for (uint32_t i = 0; i < n; ++i) {
      rig_[i].fx = p4[4 * i]; rig_[i].fy = p4[4 * i + 1]; rig_[i].u0 = p4[4 * i + 2]; rig_[i].v0 = p4[4 * i + 3];
      rig_[i].t_vs = SE3::from7(tvs7 + 7 * i);
    }
    return 0;
  }
  // AddPose, BundleAdjuster.h:292-323
  int set_poses(uint32_t n, const double* t, const double* v, const double* b, const uint8_t* act, const uint8_t* mu, const uint8_t* m) override {
    poses_.clear(); num_active_poses_ = 0;
    for (uint32_t i = 0; i < n; ++i) {
      Pose pose;
      pose.t_wp = SE3::from7(t + 7 * i);
      if (v) for (int k = 0; k < 3; ++k) pose.v_w[k] = v[3 * i + k];
      if (b) for (int k = 0; k < 6; ++k) pose.b[k] = b[6 * i + k];
      pose.is_active = act ? act[i] != 0 : true;
      pose.is_param_mask_used = mu ? mu[i] != 0 : false;
      if (pose.is_param_mask_used && m) { pose.param_mask.resize(D); for (int k = 0; k < D; ++k) pose.param_mask[k] = m[D * i + k] != 0; }
      pose.id = i;
      if (pose.is_active) { pose.opt_id = num_active_poses_; num_active_poses_++; } else pose.opt_id = UINT_MAX;
      poses_.push_back(pose);
    }
    return 0;
  }
  // AddLandmark, BundleAdjuster.h:326-367
  int set_landmarks(uint32_t n, const double* xw, const uint32_t* rp, const uint32_t* rc, const uint8_t* act) override {
    landmarks_.clear(); num_active_landmarks_ = 0;
    for (uint32_t i = 0; i < n; ++i) {
      if (rp[i] >= poses_.size()) return BAGPU_ERR_INVALID;
      LandmarkT lm;
      for (int k = 0; k < 4; ++k) lm.x_w[k] = xw[4 * i + k];
      lm.ref_pose_id = rp[i]; lm.ref_cam_id = rc ? rc[i] : 0;
      lm.is_active = act ? act[i] != 0 : true; lm.is_reliable = true; lm.id = i;
      if (lm.is_active) { lm.opt_id = num_active_landmarks_; num_active_landmarks_++; } else lm.opt_id = UINT_MAX;
      landmarks_.push_back(lm);
    }
    return 0;
  }
  // AddProjectionResidual, BundleAdjuster.h:459-513
  int set_proj(uint32_t n, const double* z, const uint32_t* mp, const uint32_t* lmid, const uint32_t* cam, const double* w) override {
    proj_residuals_.clear(); conditioning_proj_residuals_.clear(); proj_residual_offset = 0;
    for (auto& p : poses_) p.proj_residuals.clear();
    for (auto& l : landmarks_) l.proj_residuals.clear();
    proj_residuals_.reserve(n);
    for (uint32_t i = 0; i < n; ++i) {
      if (lmid[i] >= landmarks_.size() || mp[i] >= poses_.size()) return BAGPU_ERR_INVALID;
      ProjRes residual;
      residual.orig_weight = w ? w[i] : 1.0;
      residual.landmark_id = lmid[i];
      residual.x_meas_id = mp[i];
      residual.x_ref_id = landmarks_[lmid[i]].ref_pose_id;
      residual.z[0] = z[2 * i]; residual.z[1] = z[2 * i + 1];
      residual.cam_id = cam ? cam[i] : 0;
      residual.residual_id = proj_residuals_.size();
      residual.residual_offset = proj_residual_offset;
      LandmarkT& lm = landmarks_[lmid[i]];
      if (residual.x_meas_id == residual.x_ref_id && residual.cam_id == lm.ref_cam_id) lm.z_ref = residual.z;
      const uint32_t res_id = residual.residual_id;
      const bool diff_poses = residual.x_meas_id != residual.x_ref_id;
      if (diff_poses || residual.cam_id != lm.ref_cam_id || LM != 1) {
        lm.proj_residuals.push_back(res_id);
        if (diff_poses || LM != 1) {
          poses_[residual.x_meas_id].proj_residuals.push_back(res_id);
          if (LM == 1) poses_[residual.x_ref_id].proj_residuals.push_back(res_id);
        }
      } else {
        return BAGPU_ERR_INVALID;  // the reference returns -1 and does not store it
      }
      proj_residuals_.push_back(residual);   // is_conditioning stays false in the stored copy (Q2)
      proj_residual_offset += 2;
      if (poses_[residual.x_ref_id].is_active == false && poses_[residual.x_meas_id].is_active == true)
        conditioning_proj_residuals_.push_back(residual.residual_id);
    }
    return 0;
  }
  // AddUnaryConstraint, BundleAdjuster.h:377-407 (cov_inv = covariance.inverse() done by the caller)
  int set_unary(uint32_t n, const uint32_t* pid, const double* t, const double* ci, const uint8_t* ur) override {
    unary_residuals_.clear(); unary_residual_offset_ = 0;
    for (auto& p : poses_) p.unary_residuals.clear();
    for (uint32_t i = 0; i < n; ++i) {
      if (pid[i] >= poses_.size()) return BAGPU_ERR_INVALID;
      UnaryResidual r;
      r.orig_weight = 1.0; r.pose_id = pid[i]; r.residual_id = unary_residuals_.size();
      r.residual_offset = unary_residual_offset_; r.t_wp = SE3::from7(t + 7 * i);
      r.use_rotation = ur ? ur[i] != 0 : true;
      for (int k = 0; k < 36; ++k) r.cov_inv.a[k] = ci[36 * i + k];
      r.cov_inv_sqrt = sqrtm_spd(r.cov_inv);
      unary_residuals_.push_back(r); unary_residual_offset_ += 6;
      poses_[pid[i]].unary_residuals.push_back(r.residual_id);
    }
    return 0;
  }
  // AddBinaryConstraint, BundleAdjuster.h:425-456
  int set_binary(uint32_t n, const uint32_t* p1, const uint32_t* p2, const double* t, const double* ci, const double* w, const uint8_t* ur) override {
    binary_residuals_.clear(); binary_residual_offset_ = 0;
    for (auto& p : poses_) p.binary_residuals.clear();
    for (uint32_t i = 0; i < n; ++i) {
      if (p1[i] >= poses_.size() || p2[i] >= poses_.size()) return BAGPU_ERR_INVALID;
      BinaryResidual r;
      r.orig_weight = w ? w[i] : 1.0; r.x1_id = p1[i]; r.x2_id = p2[i];
      r.residual_id = binary_residuals_.size(); r.residual_offset = binary_residual_offset_;
      r.t_12 = SE3::from7(t + 7 * i);
      for (int k = 0; k < 36; ++k) r.cov_inv.a[k] = ci[36 * i + k];
      r.cov_inv_sqrt = sqrtm_spd(r.cov_inv);
      r.use_rotation = ur ? ur[i] != 0 : true;
      binary_residuals_.push_back(r); binary_residual_offset_ += 6;
      poses_[p1[i]].binary_residuals.push_back(r.residual_id);
      poses_[p2[i]].binary_residuals.push_back(r.residual_id);
    }
    return 0;
  }
  // AddImuResidual, BundleAdjuster.h:516-546
  int set_imu(uint32_t n, const uint32_t* p1, const uint32_t* p2, const uint32_t* ms, const double* m7, const double* w) override {
    inertial_residuals_.clear(); conditioning_inertial_residuals_.clear(); inertial_residual_offset_ = 0;
    for (auto& p : poses_) p.inertial_residuals.clear();
    if (n > 0 && D < 9) return BAGPU_ERR_INVALID;
    for (uint32_t i = 0; i < n; ++i) {
      if (p1[i] >= poses_.size() || p2[i] >= poses_.size()) return BAGPU_ERR_INVALID;
      ImuRes r;
      r.orig_weight = w ? w[i] : 1.0; r.pose1_id = p1[i]; r.pose2_id = p2[i];
      for (uint32_t k = ms[i]; k < ms[i + 1]; ++k) {
        ImuMeasurement m;
        for (int c = 0; c < 3; ++c) { m.w[c] = m7[7 * k + c]; m.a[c] = m7[7 * k + 3 + c]; }
        m.time = m7[7 * k + 6];
        r.measurements.push_back(m);
      }
      r.residual_id = inertial_residuals_.size(); r.residual_offset = inertial_residual_offset_;
      inertial_residuals_.push_back(r); inertial_residual_offset_ += D;
      poses_[p1[i]].inertial_residuals.push_back(r.residual_id);
      poses_[p2[i]].inertial_residuals.push_back(r.residual_id);
      if (poses_[p1[i]].is_active == false && poses_[p2[i]].is_active == true)
        conditioning_inertial_residuals_.push_back(r.residual_id);
    }
    return 0;
  }

  // BundleAdjuster.h:634-652
  uint32_t GetGravityRegularizationDimension(uint32_t pose_id) {
    Pose& pose = poses_[pose_id];
    const Vec3 gravity = imu_.g_vec;
    const Mat3 rot = pose.t_wp.rotationMatrix();
    double max_dot = 0; uint32_t max_dim = 0;
    for (uint32_t ii = 0; ii < 3; ++ii) {
      const double d = std::fabs(rot(0, ii) * gravity[0] + rot(1, ii) * gravity[1] + rot(2, ii) * gravity[2]);
      if (d > max_dot) { max_dot = d; max_dim = ii; }
    }
    return max_dim + 3;
  }

  // ---- ApplyUpdate, src/BundleAdjuster.cpp:21-140 ------------------------------
  void ApplyUpdate(const Delta& delta, const bool do_rollback, const double damping = 1.0) {
    summary.delta_norm = std::sqrt(squaredNorm(delta.delta_l)) + std::sqrt(squaredNorm(delta.delta_p));
    const double coef = (do_rollback == true ? -1.0 : 1.0) * damping;
    for (size_t ii = 0; ii < poses_.size(); ++ii) {
      if (poses_[ii].is_active) {
        const uint32_t p_offset = poses_[ii].opt_id * kPoseDim;
        Vec6 p_update;
        for (int k = 0; k < 6; ++k) p_update[k] = -delta.delta_p[p_offset + k] * coef;
        poses_[ii].t_wp = exp_decoupled(poses_[ii].t_wp, p_update);
        if (kVelInState) for (int k = 0; k < 3; ++k) poses_[ii].v_w[k] -= delta.delta_p[p_offset + 6 + k] * coef;
        if (kBiasInState) for (int k = 0; k < 6; ++k) poses_[ii].b[k] -= delta.delta_p[p_offset + 9 + k] * coef;
      }
      poses_[ii].t_sw.clear();
    }
    for (size_t ii = 0; ii < landmarks_.size(); ++ii) {
      if (landmarks_[ii].is_active) {
        if (kLmDim == 1) {
          const double lm_delta = delta.delta_l[landmarks_[ii].opt_id * LMS] * coef;
          landmarks_[ii].x_s[3] -= lm_delta;
          if (landmarks_[ii].x_s[3] < 0) { landmarks_[ii].x_s[3] += lm_delta; landmarks_[ii].is_reliable = false; }
        } else {
          for (int k = 0; k < LM; ++k) landmarks_[ii].x_w[k] -= delta.delta_l[landmarks_[ii].opt_id * LMS + k] * coef;
        }
      }
    }
  }

  // ---- EvaluateResiduals, src/BundleAdjuster.cpp:144-274 -------------------------
  void EvaluateResiduals(double* proj_error, double* binary_error, double* unary_error, double* inertial_error) {
    if (proj_error) {
      for (LandmarkT& lm : landmarks_) lm.num_outlier_residuals = 0;
      *proj_error = 0;
      for (ProjRes& res : proj_residuals_) {
        LandmarkT& lm = landmarks_[res.landmark_id];
        Pose& pose = poses_[res.x_meas_id];
        Pose& ref_pose = poses_[res.x_ref_id];
        const SE3 t_sw_m = pose.GetTsw(res.cam_id, rig_);
        const SE3 t_ws_r = ref_pose.GetTsw(lm.ref_cam_id, rig_).inverse();
        Vec3 xw3, xs3;
        for (int k = 0; k < 3; ++k) { xw3[k] = lm.x_w[k]; xs3[k] = lm.x_s[k]; }
        const Vec2 p = kLmDim == 3 ? rig_[res.cam_id].Transfer3d(t_sw_m, xw3, lm.x_w[3])
                                   : rig_[res.cam_id].Transfer3d(t_sw_m * t_ws_r, xs3, lm.x_s[3]);
        res.residual = res.z - p;
        res.mahalanobis_distance = res.residual.squaredNorm() * res.weight;
        *proj_error += res.mahalanobis_distance;
        if (res.residual.norm() > opt.projection_outlier_threshold) landmarks_[res.landmark_id].num_outlier_residuals++;
      }
    }
    if (unary_error) {
      *unary_error = 0;
      for (UnaryResidual& res : unary_residuals_) {
        const Pose& pose = poses_[res.pose_id];
        res.residual = log_decoupled(pose.t_wp, res.t_wp);
        if (!res.use_rotation) for (int k = 3; k < 6; ++k) res.residual[k] = 0;
        res.mahalanobis_distance = (res.residual.transpose() * res.cov_inv * res.residual)(0, 0);
        *unary_error += res.mahalanobis_distance;
      }
    }
    if (binary_error) {
      *binary_error = 0;
      for (BinaryResidual& res : binary_residuals_) {
        const Pose& pose1 = poses_[res.x1_id];
        const Pose& pose2 = poses_[res.x2_id];
        res.residual = log_decoupled(pose1.t_wp.inverse() * pose2.t_wp, res.t_12);
        if (!res.use_rotation) for (int k = 3; k < 6; ++k) res.residual[k] = 0;
        res.mahalanobis_distance = res.residual.squaredNorm() * res.weight;
        *binary_error += res.mahalanobis_distance;
      }
    }
    if (inertial_error) {
      *inertial_error = 0;
      if constexpr (D >= 9)
      for (ImuRes& res : inertial_residuals_) {
        const Vec3 gravity = imu_.g_vec;
        const Pose& pose1 = poses_[res.pose1_id];
        const Pose& pose2 = poses_[res.pose2_id];
        Vec3 bg, ba;
        for (int k = 0; k < 3; ++k) { bg[k] = pose1.b[k]; ba[k] = pose1.b[3 + k]; }
        const ImuPose imu_pose = ImuRes::IntegrateResidual(ImuPose(pose1.t_wp, pose1.v_w, Vec3(), 0), res.measurements, bg, ba, gravity, res.poses);
        const SE3& t_wb = pose2.t_wp;
        res.residual.setZero();
        const Vec6 l = log_decoupled(imu_pose.t_wp, t_wb);
        for (int k = 0; k < 6; ++k) res.residual[k] = l[k];
        for (int k = 0; k < 3; ++k) res.residual[6 + k] = imu_pose.v_w[k] - pose2.v_w[k];
        if (kBiasInState) for (int k = 0; k < 6; ++k) res.residual[(D >= 15 ? 9 : 0) + k] = pose1.b[k] - pose2.b[k];
        res.mahalanobis_distance = (res.residual.transpose() * res.cov_inv * res.residual)(0, 0);
        *inertial_error += res.mahalanobis_distance;
      }
    }
  }

  // ---- projection functor, parallel_algos.h:35-152 ---------------------------------
  void ProjectionResiduals(std::vector<double>& errors, std::vector<double>& cond_errors) {
    for (size_t ii = 0; ii < proj_residuals_.size(); ii++) {
      ProjRes& res = proj_residuals_[ii];
      LandmarkT& lm = landmarks_[res.landmark_id];
      Pose& pose = poses_[res.x_meas_id];
      Pose& ref_pose = poses_[res.x_ref_id];
      const LinearCamera& cam = rig_[res.cam_id];
      const SE3& t_vs_m = rig_[res.cam_id].t_vs;
      const SE3& t_vs_r = rig_[lm.ref_cam_id].t_vs;
      const SE3 t_sw_m = pose.GetTsw(res.cam_id, rig_);
      const SE3 t_ws_r = ref_pose.GetTsw(lm.ref_cam_id, rig_).inverse();
      Vec3 xw3, xs3;
      for (int k = 0; k < 3; ++k) { xw3[k] = lm.x_w[k]; xs3[k] = lm.x_s[k]; }
      const Vec2 p = kLmDim == 3 ? cam.Transfer3d(t_sw_m, xw3, lm.x_w[3]) : cam.Transfer3d(t_sw_m * t_ws_r, xs3, lm.x_s[3]);
      res.residual = res.z - p;
      const Vec4 x_s_m = kLmDim == 1 ? MultHomogeneous(t_sw_m * t_ws_r, lm.x_s) : MultHomogeneous(t_sw_m, lm.x_w);
      Vec3 xsm3; for (int k = 0; k < 3; ++k) xsm3[k] = x_s_m[k];
      const Mat<2, 4> dt_dp_m = cam.dTransfer3d_dray(SE3(), xsm3, x_s_m[3]);
      const Mat<2, 4> dt_dp_s = kLmDim == 3 ? dt_dp_m * t_sw_m.matrix() : dt_dp_m * (t_sw_m * t_ws_r).matrix();
      if (lm.is_active) {
        for (int r = 0; r < 2; ++r) for (int c = 0; c < LMS; ++c) res.dz_dlm(r, c) = -dt_dp_s(r, (kLmDim == 3 ? 0 : 3) + c);
      }
      const bool diff_poses = res.x_ref_id != res.x_meas_id;
      if (pose.is_active || ref_pose.is_active) {
        if (diff_poses) {
          // LmSize==3: the reference reads lm.x_s here, which is never initialised for
          // LmSize 3; the intended world point is lm.x_w.
          const Vec4 x_world = kLmDim == 1 ? Vec4(t_ws_r.matrix() * lm.x_s) : lm.x_w;
          res.dz_dx_meas = -dt_dp_m * dt_x_dt(t_sw_m, x_world) * dt1_t2_dt2(t_vs_m.inverse()) * dinv_exp_decoupled_dx(pose.t_wp);
        } else {
          res.dz_dx_meas.setZero();
        }
        if (kLmDim == 1) {
          if (diff_poses) {
            res.dz_dx_ref = -dt_dp_m * dt_x_dt(t_sw_m * ref_pose.t_wp, Vec4(t_vs_r.matrix() * lm.x_s)) * dt1_t2_dt2(t_sw_m) * dexp_decoupled_dx(ref_pose.t_wp);
          } else {
            res.dz_dx_ref.setZero();
          }
        }
      }
      res.weight = res.orig_weight;
      res.mahalanobis_distance = res.residual.squaredNorm() * res.weight;
      if (res.is_conditioning) cond_errors.push_back(res.mahalanobis_distance);
      else errors.push_back(res.mahalanobis_distance);
    }
  }

  // ---- inertial functor, parallel_algos.h:178-358 ----------------------------------
  // The reference runs this functor under tbb::parallel_reduce (src/BundleAdjuster.cpp:1494); the residuals are
  // independent, so a static split over std::threads gives the same numbers, and the error list is filled in
  // residual order afterwards (what the reduce's join produces).  imu_threads_used reports the split for the
  // cpu_baseline's core count; BAORACLE_THREADS=1 forces the serial loop.
  int imu_threads_used = 1;
  void InertialResiduals(std::vector<double>& errors) {
    const size_t n = inertial_residuals_.size();
    size_t nt = std::thread::hardware_concurrency();
    if (const char* e = getenv("BAORACLE_THREADS")) nt = (size_t)std::max(1, atoi(e));
    nt = std::max<size_t>(1, std::min(nt, n / 8));
    imu_threads_used = (int)nt;
    if (nt <= 1) {
      for (size_t ii = 0; ii < n; ii++) InertialResidual(ii);
    } else {
      std::vector<std::thread> pool;
      for (size_t t = 0; t < nt; ++t)
        pool.emplace_back([this, t, nt, n]() { for (size_t ii = n * t / nt; ii < n * (t + 1) / nt; ++ii) InertialResidual(ii); });
      for (std::thread& th : pool) th.join();
    }
    if constexpr (D >= 9)
      for (size_t ii = 0; ii < n; ii++) errors.push_back(inertial_residuals_[ii].mahalanobis_distance);
  }
  void InertialResidual(const size_t ii) {
    if constexpr (D >= 9)
    {
      ImuRes& res = inertial_residuals_[ii];
      const Vec3 gravity = imu_.g_vec;
      const Pose& pose1 = poses_[res.pose1_id];
      const Pose& pose2 = poses_[res.pose2_id];
      bool compute_covarince = !opt.calculate_inertial_covariance_once || !res.covariance_computed;
      if (compute_covarince) res.c_integration.setZero();
      Vec3 bg, ba;
      for (int k = 0; k < 3; ++k) { bg[k] = pose1.b[k]; ba[k] = pose1.b[3 + k]; }
      ImuPose imu_pose = ImuRes::IntegrateResidual(ImuPose(pose1.t_wp, pose1.v_w, Vec3(), 0), res.measurements, bg, ba, gravity, res.poses,
                                                   compute_covarince ? &res.dintegration_db : nullptr, nullptr,
                                                   compute_covarince ? &res.c_integration : nullptr, compute_covarince ? &imu_.r : nullptr);
      res.covariance_computed = true;
      double total_dt = res.measurements.back().time - res.measurements.front().time;
      const SE3& t_w1 = pose1.t_wp;
      const SE3& t_w2 = pose2.t_wp;
      SE3 t_12_0 = imu_pose.t_wp;
      t_12_0.translation() -= (-gravity * 0.5 * powi(total_dt, 2) + pose1.v_w * total_dt);
      t_12_0 = pose1.t_wp.inverse() * t_12_0;
      Vec3 v_12_0 = imu_pose.v_w - pose1.v_w;
      v_12_0 += gravity * total_dt;
      v_12_0 = pose1.t_wp.so3().inverse() * v_12_0;
      res.residual.setZero(); res.dz_dx1.setZero(); res.dz_dx2.setZero(); res.dz_db.setZero();
      res.dz_dx1.setBlock(0, 6, Mat3::Identity() * total_dt);
      for (int k = 0; k < 3; ++k) {
        Vec3 c = t_w1.so3().matrix() * SO3::generator(k) * v_12_0;
        res.dz_dx1.setBlock(6, 3 + k, c);
      }
      res.dz_dx1.setBlock(6, 6, Mat3::Identity());
      Mat<6, 6> d1 = dLog_decoupled_dt1(imu_pose.t_wp, t_w2) * dt1_t2_dt1(t_w1, t_12_0) * dexp_decoupled_dx(t_w1);
      res.dz_dx1.setBlock(0, 0, d1);
      Mat<6, 6> d2 = dlog_decoupled_dt2(imu_pose.t_wp, t_w2) * dexp_decoupled_dx(t_w2);
      res.dz_dx2.setBlock(0, 0, d2);
      res.dz_dx2.setBlock(6, 6, -Mat3::Identity());
      res.weight = res.orig_weight;
      const Vec6 l = log_decoupled(imu_pose.t_wp, t_w2);
      for (int k = 0; k < 6; ++k) res.residual[k] = l[k];
      for (int k = 0; k < 3; ++k) res.residual[6 + k] = imu_pose.v_w[k] - pose2.v_w[k];
      const Mat<6, 7> dlogt1t2_dt1 = dLog_decoupled_dt1(imu_pose.t_wp, t_w2);
      Mat<9, 10> dse3t1t2v_dt1;
      dse3t1t2v_dt1.setBlock(0, 0, dlogt1t2_dt1);
      dse3t1t2v_dt1.setBlock(6, 7, Mat3::Identity());
      res.cov_inv.setZero();
      Mat<D, 1> sigmas; for (int k = 0; k < D; ++k) sigmas[k] = 1.0;
      if (kBiasInState) for (int k = 0; k < 6; ++k) sigmas[(D >= 15 ? 9 : 0) + k] = imu_.r_b[k] * total_dt;
      for (int k = 0; k < D; ++k) res.cov_inv(k, k) = sigmas[k];
      Mat<9, 9> c9 = dse3t1t2v_dt1 * res.c_integration * dse3t1t2v_dt1.transpose();
      res.cov_inv.setBlock(0, 0, c9);
      res.cov = res.cov_inv;
      res.cov_inv = inverse(res.cov_inv);
      if (kBiasInState) {
        constexpr int BO = (D >= 15 ? 9 : 0);
        Mat<6, 6> db = dlogt1t2_dt1 * res.dintegration_db.template block<7, 6>(0, 0);
        res.dz_db.setBlock(0, 0, db);
        res.dz_db.setBlock(6, 0, res.dintegration_db.template block<3, 6>(7, 0));
        res.dz_db.setBlock(BO, 0, Mat<6, 6>::Identity());
        res.dz_dx1.setBlock(0, BO, res.dz_db);
        res.dz_dx2.setBlock(BO, BO, -Mat<6, 6>::Identity());
        for (int k = 0; k < 6; ++k) res.residual[BO + k] = pose1.b[k] - pose2.b[k];
      }
      res.mahalanobis_distance = (res.residual.transpose() * res.cov_inv * res.residual)(0, 0);
    }
  }

  // ---- BuildProblem, src/BundleAdjuster.cpp:1166-1803 ------------------------------
  void BuildProblem() {
    const uint32_t num_poses = num_active_poses_;
    const uint32_t num_lm = num_active_landmarks_;
    const uint32_t num_proj_res = proj_residuals_.size();
    const uint32_t num_bin_res = binary_residuals_.size();
    const uint32_t num_un_res = unary_residuals_.size();
    const uint32_t num_im_res = inertial_residuals_.size();
    if (num_proj_res > 0) {
      j_pr_.resize(num_proj_res, num_poses); jt_pr.resize(num_poses, num_proj_res); j_l_.resize(num_proj_res, num_lm);
      r_pr_.assign(num_proj_res * 2, 0.0);
    }
    if (num_bin_res > 0) { j_pp_.resize(num_bin_res, num_poses); jt_pp_.resize(num_poses, num_bin_res); r_pp_.assign(num_bin_res * 6, 0.0); }
    if (num_un_res > 0) { j_u_.resize(num_un_res, num_poses); jt_u_.resize(num_poses, num_un_res); r_u_.assign(num_un_res * 6, 0.0); }
    if (num_im_res > 0) { j_i_.resize(num_im_res, num_poses); jt_i_.resize(num_poses, num_im_res); r_i_.assign(num_im_res * D, 0.0); }
    is_param_mask_used_ = false;

    // cpp:1239-1259 (note the break at the first inactive pose, Q8)
    bool are_all_active = true;
    for (Pose& pose : poses_) {
      for (size_t ii = 0; ii < rig_.size(); ++ii) pose.GetTsw(ii, rig_);
      if (pose.is_active == false) { are_all_active = false; break; }
      if (pose.proj_residuals.size() == 0 && pose.binary_residuals.size() == 0 && pose.unary_residuals.size() == 0 &&
          pose.inertial_residuals.size() == 0) {
        pose.is_param_mask_used = true;
        pose.param_mask.assign(kPoseDim, false);
      }
    }
    // cpp:1263-1279
    if (kVelInState) {
      for (Pose& pose : poses_) {
        if (pose.inertial_residuals.empty() && pose.is_active) {
          pose.is_param_mask_used = true;
          pose.param_mask.assign(kPoseDim, true);
          pose.param_mask[6] = pose.param_mask[7] = pose.param_mask[8] = false;
          if (kBiasInState) for (int k = 9; k < 15 && k < D; ++k) pose.param_mask[k] = false;
        }
      }
    }
    // cpp:1285-1330
    if (are_all_active && num_un_res == 0 && opt.enable_auto_regularization) {
      Pose& root_pose = poses_[root_pose_id_];
      root_pose.is_param_mask_used = true;
      root_pose.param_mask.assign(kPoseDim, true);
      root_pose.param_mask[0] = root_pose.param_mask[1] = root_pose.param_mask[2] = false;
      if (kBiasInState && opt.regularize_biases_in_batch) for (int k = 9; k < 15 && k < D; ++k) root_pose.param_mask[k] = false;
      if (!kVelInState) {
        root_pose.param_mask[3] = root_pose.param_mask[4] = root_pose.param_mask[5] = false;
      } else {
        uint32_t reg_dim = GetGravityRegularizationDimension(root_pose_id_);
        root_pose.param_mask[reg_dim] = false;
      }
    }

    errors_.reserve(num_proj_res); errors_.clear();
    std::vector<double> cond_errors; cond_errors.reserve(num_proj_res);
    proj_error_ = 0;
    double t0 = now_s();
    std::vector<double> proj_errors; proj_errors.reserve(num_proj_res);
    ProjectionResiduals(proj_errors, cond_errors);
    errors_ = proj_errors;
    phase_s[BAGPU_PH_RESIDUAL] += now_s() - t0; t0 = now_s();
    last_median_ = 0;
    // cpp:1355-1388
    if (errors_.size() > 0) {
      auto it = errors_.begin() + (size_t)std::floor(errors_.size() * 0.5);
      std::nth_element(errors_.begin(), it, errors_.end());
      const double sigma = std::sqrt(*it);
      last_median_ = *it;
      double cond_sigma = 0;
      if (cond_errors.size() > 0) {
        auto it2 = cond_errors.begin() + (size_t)std::floor(cond_errors.size() * 0.5);
        std::nth_element(cond_errors.begin(), it2, cond_errors.end());
        cond_sigma = std::sqrt(*it2);
      }
      const double c_huber = 1.2107 * sigma;
      const double cond_c_huber = 1.2107 * cond_sigma;
      for (ProjRes& res : proj_residuals_) {
        const double e = std::sqrt(res.mahalanobis_distance);
        const bool use_robust = opt.use_robust_norm_for_proj_residuals;
        const bool is_outlier = e > (res.is_conditioning ? cond_c_huber : c_huber);
        res.weight *= (is_outlier && use_robust ? c_huber / e : 1.0);
        res.mahalanobis_distance = res.residual.squaredNorm() * res.weight;
        r_pr_[res.residual_offset] = res.residual[0] * std::sqrt(res.weight);
        r_pr_[res.residual_offset + 1] = res.residual[1] * std::sqrt(res.weight);
        proj_error_ += res.mahalanobis_distance;
      }
    }
    errors_.clear();
    phase_s[BAGPU_PH_MEDIAN] += now_s() - t0; t0 = now_s();

    // binary, cpp:1392-1429
    binary_error_ = 0;
    for (BinaryResidual& res : binary_residuals_) {
      const SE3& t_w1 = poses_[res.x1_id].t_wp;
      const SE3& t_w2 = poses_[res.x2_id].t_wp;
      const SE3 t_1w = t_w1.inverse();
      const SE3 t_12 = t_1w * t_w2;
      res.residual = res.cov_inv_sqrt * log_decoupled(t_12, res.t_12);
      const Mat<6, 7> dlog_dt1 = dLog_decoupled_dt1(t_12, res.t_12);
      res.dz_dx1 = dlog_dt1 * dt1_t2_dt1(t_1w, t_w2) * dinv_exp_decoupled_dx(t_w1);
      res.dz_dx2 = dlog_dt1 * dt1_t2_dt2(t_1w) * dexp_decoupled_dx(t_w2);
      if (!res.use_rotation) {
        for (int k = 3; k < 6; ++k) { res.residual[k] = 0; for (int c = 0; c < 6; ++c) { res.dz_dx1(k, c) = 0; res.dz_dx2(k, c) = 0; } }
      }
      res.weight = res.orig_weight;
      for (int k = 0; k < 6; ++k) r_pp_[res.residual_offset + k] = res.residual[k];
      res.mahalanobis_distance = (res.residual.transpose() * res.cov_inv * res.residual)(0, 0);
      binary_error_ += res.mahalanobis_distance * res.weight;
    }
    // unary, cpp:1431-1484
    unary_error_ = 0;
    errors_.clear();
    for (UnaryResidual& res : unary_residuals_) {
      SE3& t_wp = poses_[res.pose_id].t_wp;
      res.dz_dx = dlog_decoupled_dx(t_wp, res.t_wp);
      res.residual = log_decoupled(t_wp, res.t_wp);
      if (!res.use_rotation) { for (int k = 3; k < 6; ++k) { res.residual[k] = 0; for (int c = 0; c < 6; ++c) res.dz_dx(k, c) = 0; } }
      res.weight = res.orig_weight;
      res.mahalanobis_distance = (res.residual.transpose() * res.cov_inv * res.residual)(0, 0);
      errors_.push_back(res.mahalanobis_distance);
    }
    if (errors_.size() > 0) {
      auto it = errors_.begin() + (size_t)std::floor(errors_.size() * 0.5);
      std::nth_element(errors_.begin(), it, errors_.end());
      const double sigma = std::sqrt(*it);
      const double c_huber = 1.2107 * sigma;
      for (UnaryResidual& res : unary_residuals_) {
        const double e = std::sqrt(res.mahalanobis_distance);
        const double weight = ((e > c_huber) ? c_huber / e : 1.0);
        res.cov_inv = res.cov_inv * weight;        // compounds over iterations (Q6)
        res.cov_inv_sqrt = sqrtm_spd(res.cov_inv);
        Vec6 res_std_form = res.cov_inv_sqrt * res.residual;
        for (int k = 0; k < 6; ++k) r_u_[res.residual_offset + k] = res_std_form[k];
        res.mahalanobis_distance = res_std_form.squaredNorm();
        unary_error_ += res.mahalanobis_distance;
      }
    }
    errors_.clear();
    // inertial, cpp:1486-1541
    inertial_error_ = 0;
    std::vector<double> imu_errors;
    InertialResiduals(imu_errors);
    errors_ = proj_errors;  // Q3: the list is taken from the *projection* functor (cpp:1497)
    if (errors_.size() > 0) {
      auto it = errors_.begin() + (size_t)std::floor(errors_.size() * 0.5);
      std::nth_element(errors_.begin(), it, errors_.end());
      const double sigma = std::sqrt(*it);
      const double c_huber = 1.2107 * sigma;
      for (ImuRes& res : inertial_residuals_) {
        const bool use_robust = opt.use_robust_norm_for_inertial_residuals;
        const bool is_cond = !poses_[res.pose1_id].is_active && poses_[res.pose2_id].is_active;
        const double e = std::sqrt(res.mahalanobis_distance);
        const double weight = ((e > c_huber) && !is_cond && use_robust ? c_huber / e : 1.0);
        res.cov_inv = res.cov_inv * weight;
        res.cov_inv_sqrt = sqrtm_spd(res.cov_inv);
        Mat<D, 1> res_std_form = res.cov_inv_sqrt * res.residual;
        for (int k = 0; k < D; ++k) r_i_[res.residual_offset + k] = res_std_form[k];
        res.mahalanobis_distance = res_std_form.squaredNorm();
        inertial_error_ += res.mahalanobis_distance;
      }
    }
    errors_.clear();
    phase_s[BAGPU_PH_OTHER_RES] += now_s() - t0; t0 = now_s();

    // J insertion, cpp:1552-1802
    for (Pose& pose : poses_) {
      if (pose.is_active) {
        std::sort(pose.proj_residuals.begin(), pose.proj_residuals.end());
        for (const int id : pose.proj_residuals) {
          ProjRes& res = proj_residuals_[id];
          Mat<2, 6>& dz_dx = res.x_meas_id == pose.id ? res.dz_dx_meas : res.dz_dx_ref;
          if (pose.is_param_mask_used) {
            is_param_mask_used_ = true;
            for (uint32_t ii = 0; ii < kPrPoseDim; ++ii) if (!pose.param_mask[ii]) dz_dx.setColZero(ii);
          }
          j_pr_.insert(res.residual_id, pose.opt_id) = dz_dx * std::sqrt(res.weight);
          jt_pr.insert(pose.opt_id, res.residual_id) = dz_dx.transpose() * std::sqrt(res.weight);
        }
        std::sort(pose.binary_residuals.begin(), pose.binary_residuals.end());
        for (const int id : pose.binary_residuals) {
          BinaryResidual& res = binary_residuals_[id];
          Mat<6, 6>& dz_dz = res.x1_id == pose.id ? res.dz_dx1 : res.dz_dx2;
          if (pose.is_param_mask_used) {
            is_param_mask_used_ = true;
            for (int ii = 0; ii < 6; ++ii) if (!pose.param_mask[ii]) dz_dz.setColZero(ii);
          }
          Mat<6, D>& a = j_pp_.insert(res.residual_id, pose.opt_id);
          a.setBlock(0, 0, res.cov_inv_sqrt * dz_dz);
          Mat<D, 6>& at = jt_pp_.insert(pose.opt_id, res.residual_id);
          at.setBlock(0, 0, dz_dz.transpose() * res.cov_inv_sqrt * res.weight);
        }
        std::sort(pose.unary_residuals.begin(), pose.unary_residuals.end());
        for (const int id : pose.unary_residuals) {
          UnaryResidual& res = unary_residuals_[id];
          if (pose.is_param_mask_used) {
            is_param_mask_used_ = true;
            for (int ii = 0; ii < 6; ++ii) if (!pose.param_mask[ii]) res.dz_dx.setColZero(ii);
          }
          Mat<6, D>& a = j_u_.insert(res.residual_id, pose.opt_id);
          a.setBlock(0, 0, res.cov_inv_sqrt * res.dz_dx);
          Mat<D, 6>& at = jt_u_.insert(pose.opt_id, res.residual_id);
          at.setBlock(0, 0, res.dz_dx.transpose() * res.cov_inv_sqrt);
        }
        std::sort(pose.inertial_residuals.begin(), pose.inertial_residuals.end());
        for (const int id : pose.inertial_residuals) {
          ImuRes& res = inertial_residuals_[id];
          Mat<D, D> dz_dz = res.pose1_id == pose.id ? res.dz_dx1 : res.dz_dx2;
          if (pose.is_param_mask_used) {
            is_param_mask_used_ = true;
            for (uint32_t ii = 0; ii < (uint32_t)kPoseDim; ++ii) if (!pose.param_mask[ii]) dz_dz.setColZero(ii);
          }
          j_i_.insert(res.residual_id, pose.opt_id) = res.cov_inv_sqrt * dz_dz;
          const Mat<D, D> trans = dz_dz.transpose();
          jt_i_.insert(pose.opt_id, res.residual_id) = trans * res.cov_inv_sqrt;
        }
      }
    }
    for (LandmarkT& lm : landmarks_) {
      if (lm.is_active) {
        std::sort(lm.proj_residuals.begin(), lm.proj_residuals.end());
        for (const int id : lm.proj_residuals) {
          const ProjRes& res = proj_residuals_[id];
          j_l_.insert(res.residual_id, lm.opt_id) = res.dz_dlm * std::sqrt(res.weight);
        }
      }
    }
    phase_s[BAGPU_PH_ACCUMULATE] += now_s() - t0;
  }

  // ---- problem set-up inside Solve, src/BundleAdjuster.cpp:305-598 ---------------------
  void SetupProblem() {
    double t0 = now_s();
    const uint32_t num_poses = num_active_poses_;
    const uint32_t num_pose_params = num_poses * kPoseDim;
    const uint32_t num_lm = num_active_landmarks_;
    rhs_p_.assign(num_pose_params, 0.0);
    rhs_k_.clear();
    vi_.resize(num_lm, num_lm);
    rhs_p_sc_.assign(num_pose_params, 0.0);
    jt_l_j_pr_.resize(num_lm, num_poses);
    BlockMat<6, LMS> jt_pr_j_l_vi(num_poses, num_lm);
    u_.resize(num_poses, num_poses);
    schur_term_.resize(num_poses, num_poses);

    if (proj_residuals_.size() > 0 && num_poses > 0) {
      BlockMat<6, 6> jt_pr_j_pr(num_poses, num_poses);
      SparseBlockProduct(jt_pr, j_pr_, jt_pr_j_pr, opt.use_triangular_matrices != 0);
      BlockMat<D, D> temp_u = u_;
      SparseBlockAdd(temp_u, jt_pr_j_pr, u_);
      VecX jt_pr_r_pr(num_pose_params);
      SparseBlockVectorProductDenseResult(jt_pr, r_pr_, jt_pr_r_pr, -1, kPoseDim);
      for (uint32_t i = 0; i < num_pose_params; ++i) rhs_p_[i] += jt_pr_r_pr[i];
    }
    if (binary_residuals_.size() > 0) {
      BlockMat<D, D> jt_pp_j_pp(num_poses, num_poses);
      SparseBlockProduct(jt_pp_, j_pp_, jt_pp_j_pp, opt.use_triangular_matrices != 0);
      BlockMat<D, D> temp_u = u_;
      SparseBlockAdd(temp_u, jt_pp_j_pp, u_);
      VecX v(num_pose_params);
      SparseBlockVectorProductDenseResult(jt_pp_, r_pp_, v);
      for (uint32_t i = 0; i < num_pose_params; ++i) rhs_p_[i] += v[i];
    }
    if (unary_residuals_.size() > 0) {
      BlockMat<D, D> jt_u_j_u(num_poses, num_poses);
      SparseBlockProduct(jt_u_, j_u_, jt_u_j_u, opt.use_triangular_matrices != 0);
      BlockMat<D, D> temp_u = u_;
      SparseBlockAdd(temp_u, jt_u_j_u, u_);
      VecX v(num_pose_params);
      SparseBlockVectorProductDenseResult(jt_u_, r_u_, v);
      for (uint32_t i = 0; i < num_pose_params; ++i) rhs_p_[i] += v[i];
    }
    if (inertial_residuals_.size() > 0) {
      BlockMat<D, D> jt_i_j_i(num_poses, num_poses);
      SparseBlockProduct(jt_i_, j_i_, jt_i_j_i, opt.use_triangular_matrices != 0);
      BlockMat<D, D> temp_u = u_;
      SparseBlockAdd(temp_u, jt_i_j_i, u_);
      VecX v(num_pose_params);
      SparseBlockVectorProductDenseResult(jt_i_, r_i_, v);
      for (uint32_t i = 0; i < num_pose_params; ++i) rhs_p_[i] += v[i];
    }

    auto init_profile = [&](const BlockMat<6, 6>* term) {
      std::vector<int> first(num_pose_params);
      for (uint32_t j = 0; j < num_pose_params; ++j) first[j] = j;
      for (uint32_t bj = 0; bj < num_poses; ++bj) {
        int fb = bj;
        if (!u_.col[bj].empty()) fb = std::min(fb, u_.col[bj].front().index);
        if (term && !term->col[bj].empty()) fb = std::min(fb, term->col[bj].front().index);
        for (int c = 0; c < D; ++c) first[bj * D + c] = std::min<int>(bj * D + c, fb * D);
      }
      s_.init(num_pose_params, first);
    };

    if (kLmDim > 0 && num_lm > 0) {
      rhs_l_.assign(num_lm * LMS, 0.0);
      Mat<LMS, 1> jtr_l;
      for (uint32_t ii = 0; ii < landmarks_.size(); ++ii) {
        if (!landmarks_[ii].is_active) continue;
        landmarks_[ii].jtj.setZero();
        jtr_l.setZero();
        for (const int id : landmarks_[ii].proj_residuals) {
          const ProjRes& res = proj_residuals_[id];
          landmarks_[ii].jtj += (res.dz_dlm.transpose() * res.dz_dlm) * res.weight;
          Vec2 rr; rr[0] = r_pr_[res.residual_id * 2]; rr[1] = r_pr_[res.residual_id * 2 + 1];
          jtr_l += (res.dz_dlm.transpose() * std::sqrt(res.weight)) * rr;
        }
        for (int k = 0; k < LMS; ++k) rhs_l_[landmarks_[ii].opt_id * LMS + k] = jtr_l[k];
        if (kLmDim == 1) {
          if (std::fabs(landmarks_[ii].jtj(0, 0)) < 1e-6) landmarks_[ii].jtj(0, 0) += 1e-6;
        } else {
          if (landmarks_[ii].jtj.norm() < 1e-6) for (int k = 0; k < LMS; ++k) landmarks_[ii].jtj(k, k) += 1e-6;
        }
        vi_.insert(landmarks_[ii].opt_id, landmarks_[ii].opt_id) = inverse(landmarks_[ii].jtj);
      }
      if (num_poses > 0) {
        jt_pr_j_l_.resize(num_poses, num_lm);
        SparseBlockProduct(jt_pr, j_l_, jt_pr_j_l_);
        BlockMat<LMS, 6>::forceTranspose(jt_pr_j_l_, jt_l_j_pr_);
        SparseBlockDiagonalRhsProduct(jt_pr_j_l_, vi_, jt_pr_j_l_vi);
        SparseBlockProduct(jt_pr_j_l_vi, jt_l_j_pr_, schur_term_, opt.use_triangular_matrices != 0);
        init_profile(&schur_term_);
        // SparseBlockSubtractDenseResult(u_, term, s_) — SparseBlockMatrixOps.h:318-380
        for (uint32_t j = 0; j < num_poses; ++j) {
          for (const auto& e : schur_term_.col[j]) s_.addBlock(e.index * D, j * D, e.value, -1.0);
          for (const auto& e : u_.col[j]) s_.addBlock(e.index * D, j * D, e.value, 1.0);
        }
        VecX jt_pr_j_l_vi_bll(num_pose_params);
        SparseBlockVectorProductDenseResult(jt_pr_j_l_vi, rhs_l_, jt_pr_j_l_vi_bll, -1, kPoseDim);
        for (uint32_t i = 0; i < num_pose_params; ++i) rhs_p_sc_[i] = rhs_p_[i] - jt_pr_j_l_vi_bll[i];
      } else {
        init_profile(nullptr);
      }
    } else {
      init_profile(nullptr);
      for (uint32_t j = 0; j < num_poses; ++j) for (const auto& e : u_.col[j]) s_.addBlock(e.index * D, j * D, e.value, 1.0);
      rhs_p_sc_ = rhs_p_;
    }
    // cpp:587-598
    if (is_param_mask_used_) {
      for (Pose& pose : poses_) {
        if (pose.is_active && pose.is_param_mask_used) {
          for (uint32_t ii = 0; ii < pose.param_mask.size(); ++ii) {
            if (!pose.param_mask[ii]) { const int idx = pose.opt_id * kPoseDim + ii; s_.at(idx, idx) = 1e6; }
          }
        }
      }
    }
    phase_s[BAGPU_PH_SCHUR] += now_s() - t0;
  }

  // ---- GetLandmarkDelta, src/BundleAdjuster.cpp:709-744 ------------------------------
  void GetLandmarkDelta(const Delta& delta, const uint32_t num_poses, const uint32_t num_lm, VecX& delta_l) {
    if (num_lm > 0) {
      delta_l.assign(num_lm * LMS, 0.0);
      VecX rhs_l_sc = rhs_l_;
      if (num_poses > 0) {
        VecX wt_delta_p_k(num_lm * LMS);
        SparseBlockVectorProductDenseResult(jt_l_j_pr_, delta.delta_p, wt_delta_p_k, kPoseDim, -1);
        for (size_t i = 0; i < rhs_l_sc.size(); ++i) rhs_l_sc[i] -= wt_delta_p_k[i];
      }
      for (size_t ii = 0; ii < num_lm; ++ii) {
        const Mat<LMS, LMS> v = vi_.coeff(ii, ii);
        for (int r = 0; r < LMS; ++r) { double s = 0; for (int c = 0; c < LMS; ++c) s += v(r, c) * rhs_l_sc[ii * LMS + c]; delta_l[ii * LMS + r] = s; }
      }
    }
  }

  // ---- CalculateGn, src/BundleAdjuster.cpp:748-833 ------------------------------------
  void CalculateGn(const VecX& rhs_p, Delta& delta) {
    summary.result = BAGPU_Success;
    s_factor_ = s_;
    VecX x;
    if (!skyline_ldlt_solve(s_factor_, rhs_p, x)) { summary.result = BAGPU_FactorizationError; x.assign(rhs_p.size(), 0.0); }
    delta.delta_p = x;
    delta.delta_k.clear();
  }

  // ---- SolveInternal, src/BundleAdjuster.cpp:838-1161 ---------------------------------
  bool SolveInternal(VecX rhs_p_sc, const double gn_damping, const bool error_increase_allowed, const bool use_dogleg,
                     bagpu_iter_trace& tr, Delta& applied) {
    bool gn_computed = false;
    Delta delta_sd, delta_dl, delta_gn;
    double proj_error, binary_error, unary_error, inertial_error;
    if (use_dogleg) {
      double numerator = squaredNorm(rhs_p_) + squaredNorm(rhs_l_) + squaredNorm(rhs_k_);
      VecX j_p_rhs_p(2 * proj_residuals_.size(), 0.0), j_pp_rhs_p(6 * binary_residuals_.size(), 0.0),
          j_u_rhs_p(6 * unary_residuals_.size(), 0.0), j_i_rhs_p(D * inertial_residuals_.size(), 0.0),
          j_l_rhs_l(2 * proj_residuals_.size(), 0.0);
      if (num_active_poses_ > 0) {
        if (proj_residuals_.size() > 0) SparseBlockVectorProductDenseResult(j_pr_, rhs_p_, j_p_rhs_p, kPoseDim);
        if (inertial_residuals_.size() > 0) SparseBlockVectorProductDenseResult(j_i_, rhs_p_, j_i_rhs_p);
        if (binary_residuals_.size() > 0) SparseBlockVectorProductDenseResult(j_pp_, rhs_p_, j_pp_rhs_p);
        if (unary_residuals_.size() > 0) SparseBlockVectorProductDenseResult(j_u_, rhs_p_, j_u_rhs_p);
      }
      if (num_active_landmarks_ > 0 && proj_residuals_.size() > 0) SparseBlockVectorProductDenseResult(j_l_, rhs_l_, j_l_rhs_l);
      double denominator = 0;
      for (size_t i = 0; i < j_p_rhs_p.size(); ++i) { const double v = j_p_rhs_p[i] + j_l_rhs_l[i]; denominator += v * v; }
      denominator += squaredNorm(j_pp_rhs_p) + squaredNorm(j_u_rhs_p) + squaredNorm(j_i_rhs_p);
      double factor = numerator / denominator;
      delta_sd.delta_p = rhs_p_; for (double& v : delta_sd.delta_p) v *= factor;
      delta_sd.delta_k.clear();
      delta_sd.delta_l = rhs_l_; for (double& v : delta_sd.delta_l) v *= factor;
      double delta_sd_norm = std::sqrt(squaredNorm(delta_sd.delta_p) + squaredNorm(delta_sd.delta_l));
      tr.sd_numerator = numerator; tr.sd_denominator = denominator; tr.sd_norm = delta_sd_norm;

      uint32_t iteration_count = 0;
      while (1) {
        iteration_count++;
        if (iteration_count > opt.dogleg_max_inner_iterations) break;
        tr.inner_iterations = iteration_count;
        if (delta_sd_norm > trust_region_size_ && trust_region_size_ != kTrustRegionAuto) {
          double f = trust_region_size_ / delta_sd_norm;
          delta_dl.delta_p = delta_sd.delta_p; for (double& v : delta_dl.delta_p) v *= f;
          delta_dl.delta_l = delta_sd.delta_l; for (double& v : delta_dl.delta_l) v *= f;
        } else {
          if (!gn_computed) {
            double t0 = now_s();
            if (num_active_poses_ > 0) {
              CalculateGn(rhs_p_sc, delta_gn);
              if (!(summary.result != BAGPU_SolverError && summary.result != BAGPU_FactorizationError)) return false;
            }
            GetLandmarkDelta(delta_gn, num_active_poses_, num_active_landmarks_, delta_gn.delta_l);
            gn_computed = true;
            phase_s[BAGPU_PH_SOLVE] += now_s() - t0;
          }
          double delta_gn_norm = std::sqrt(squaredNorm(delta_gn.delta_p) + squaredNorm(delta_gn.delta_k) + squaredNorm(delta_gn.delta_l));
          tr.gn_norm = delta_gn_norm;
          const bool delta_gn_good = !std::isnan(delta_gn_norm) && !std::isinf(delta_gn_norm);
          if (delta_gn_good && trust_region_size_ == kTrustRegionAuto) trust_region_size_ = delta_gn_norm;
          if (delta_gn_good && delta_gn_norm <= trust_region_size_) {
            delta_dl = delta_gn;
          } else {
            const size_t np = delta_gn.delta_p.size(), nl = delta_gn.delta_l.size();
            VecX diff_p(np), diff_l(nl);
            for (size_t i = 0; i < np; ++i) diff_p[i] = delta_gn.delta_p[i] - delta_sd.delta_p[i];
            for (size_t i = 0; i < nl; ++i) diff_l[i] = delta_gn.delta_l[i] - delta_sd.delta_l[i];
            double a = squaredNorm(diff_p) + squaredNorm(diff_l);
            double dotp = 0; for (size_t i = 0; i < np; ++i) dotp += diff_p[i] * delta_sd.delta_p[i];
            double dotl = 0; for (size_t i = 0; i < nl; ++i) dotl += diff_l[i] * delta_sd.delta_l[i];
            double b = 2 * (dotp + dotl);
            double c = (squaredNorm(delta_sd.delta_p) + squaredNorm(delta_sd.delta_l)) - trust_region_size_ * trust_region_size_;
            double beta = 0;
            if (b * b > 4 * a * c && a > 1e-10) beta = (-(b * b) + std::sqrt(b * b - 4 * a * c)) / (2 * a);  // Q4
            delta_dl.delta_p.resize(np); delta_dl.delta_l.resize(nl);
            for (size_t i = 0; i < np; ++i) delta_dl.delta_p[i] = delta_sd.delta_p[i] + beta * diff_p[i];
            for (size_t i = 0; i < nl; ++i) delta_dl.delta_l[i] = delta_sd.delta_l[i] + beta * diff_l[i];
          }
        }
        double t0 = now_s();
        auto landmarks_copy = landmarks_;
        auto poses_copy = poses_;
        auto imu_copy = imu_;
        EvaluateResiduals(&proj_error, &binary_error, &unary_error, &inertial_error);
        summary.pre_solve_norm = proj_error + inertial_error + binary_error + unary_error;
        if (opt.apply_results) ApplyUpdate(delta_dl, false);
        EvaluateResiduals(&proj_error, &binary_error, &unary_error, &inertial_error);
        summary.post_solve_norm = proj_error + inertial_error + binary_error + unary_error;
        applied = delta_dl;
        phase_s[BAGPU_PH_STEP] += now_s() - t0;
        if (summary.post_solve_norm > summary.pre_solve_norm) {
          if (opt.apply_results) { landmarks_ = landmarks_copy; poses_ = poses_copy; imu_ = imu_copy; }
          trust_region_size_ /= 2;
        } else {
          proj_error_ = proj_error; unary_error_ = unary_error; binary_error_ = binary_error; inertial_error_ = inertial_error;
          trust_region_size_ *= 2;
          break;
        }
      }
    } else {
      Delta delta;
      if (num_active_poses_ > 0) {
        CalculateGn(rhs_p_sc, delta);
        if (!(summary.result != BAGPU_SolverError && summary.result != BAGPU_FactorizationError)) return false;
      }
      auto landmarks_copy = landmarks_;
      auto poses_copy = poses_;
      auto imu_copy = imu_;
      GetLandmarkDelta(delta, num_active_poses_, num_active_landmarks_, delta.delta_l);
      for (double& v : delta.delta_l) v *= gn_damping;
      for (double& v : delta.delta_p) v *= gn_damping;
      tr.gn_norm = std::sqrt(squaredNorm(delta.delta_p) + squaredNorm(delta.delta_l));
      EvaluateResiduals(&proj_error, &binary_error, &unary_error, &inertial_error);
      const double prev_error = proj_error + inertial_error + binary_error + unary_error;
      if (opt.apply_results) ApplyUpdate(delta, false);
      EvaluateResiduals(&proj_error, &binary_error, &unary_error, &inertial_error);
      const double postError = proj_error + inertial_error + binary_error + unary_error;
      applied = delta;
      tr.inner_iterations = 1;
      // the reference leaves summary_.pre/post_solve_norm untouched on this path (cpp:1117-1132)
      tr.pre_solve_norm = prev_error; tr.post_solve_norm = postError;
      if (postError > prev_error && !error_increase_allowed) {
        if (opt.apply_results) { landmarks_ = landmarks_copy; poses_ = poses_copy; imu_ = imu_copy; }
        summary.result = BAGPU_ErrorIncreased;
        return false;
      } else {
        proj_error_ = proj_error; unary_error_ = unary_error; binary_error_ = binary_error; inertial_error_ = inertial_error;
      }
    }
    return true;
  }

  void RecordBuild(bagpu_iter_trace& tr) {
    memset(&tr, 0, sizeof(tr));
    tr.proj_error_build = proj_error_; tr.binary_error_build = binary_error_;
    tr.unary_error_build = unary_error_; tr.inertial_error_build = inertial_error_;
    tr.median_maha = last_median_;
  }

  void PrepareXs() {  // cpp:287-296
    if (kLmDim == 1) {
      for (LandmarkT& lm : landmarks_) {
        lm.x_s = MultHomogeneous(poses_[lm.ref_pose_id].GetTsw(lm.ref_cam_id, rig_), lm.x_w);
        const double length = std::sqrt(lm.x_s[0] * lm.x_s[0] + lm.x_s[1] * lm.x_s[1] + lm.x_s[2] * lm.x_s[2]);
        lm.x_s = lm.x_s / length;
      }
    }
  }

  void BuildOnly() override {
    PrepareXs();
    BuildProblem();
    SetupProblem();
  }

  // ---- Solve, src/BundleAdjuster.cpp:278-705 -------------------------------------------
  void Solve(uint32_t uMaxIter, double gn_damping, bool error_increase_allowed) override {
    trace.clear(); rec_delta_p.clear(); rec_delta_l.clear();
    summary.iterations = 0;
    if (proj_residuals_.empty() && binary_residuals_.empty() && unary_residuals_.empty() && inertial_residuals_.empty()) return;
    PrepareXs();
    for (uint32_t kk = 0; kk < uMaxIter; ++kk) {
      BuildProblem();
      SetupProblem();
      if (opt.write_reduced_camera_matrix) dump(".");   // cpp:600-620
      bagpu_iter_trace tr; RecordBuild(tr);
      Delta applied;
      const bool ok = SolveInternal(rhs_p_sc_, gn_damping, error_increase_allowed, opt.use_dogleg != 0, tr, applied);
      summary.iterations = kk + 1;
      if (opt.use_dogleg) { tr.pre_solve_norm = summary.pre_solve_norm; tr.post_solve_norm = summary.post_solve_norm; }
      tr.delta_norm = summary.delta_norm; tr.trust_region_after = trust_region_size_;
      tr.delta_p_norm = std::sqrt(squaredNorm(applied.delta_p)); tr.delta_l_norm = std::sqrt(squaredNorm(applied.delta_l));
      if (opt.record_deltas) { rec_delta_p.push_back(applied.delta_p); rec_delta_l.push_back(applied.delta_l); }
      if (!ok) { tr.result = summary.result; trace.push_back(tr); break; }
      if ((std::fabs(summary.post_solve_norm - summary.pre_solve_norm) / summary.pre_solve_norm) < opt.error_change_threshold) {
        summary.result = BAGPU_ErrorChangeBelowThreshold; tr.result = summary.result; trace.push_back(tr); break;
      }
      if (summary.delta_norm < opt.param_change_threshold) {
        summary.result = BAGPU_ParamChangeBelowThreshold; tr.result = summary.result; trace.push_back(tr); break;
      }
      tr.result = summary.result; trace.push_back(tr);
    }
    if (kBiasInState && poses_.size() > 0) {
      for (int k = 0; k < 3; ++k) { imu_.b_g[k] = poses_.back().b[k]; imu_.b_a[k] = poses_.back().b[3 + k]; }
    }
    if (kLmDim == 1) {
      for (LandmarkT& lm : landmarks_) lm.x_w = MultHomogeneous(poses_[lm.ref_pose_id].GetTsw(lm.ref_cam_id, rig_).inverse(), lm.x_s);
    }
    summary.cond_inertial_error = 0; summary.cond_proj_error = 0;
    summary.num_cond_inertial_residuals = conditioning_inertial_residuals_.size();
    summary.num_inertial_residuals = inertial_residuals_.size();
    summary.inertial_error = inertial_error_;
    for (uint32_t id : conditioning_inertial_residuals_) summary.cond_inertial_error += inertial_residuals_[id].mahalanobis_distance;
    summary.num_cond_proj_residuals = conditioning_proj_residuals_.size();
    summary.num_proj_residuals = proj_residuals_.size();
    summary.proj_error = proj_error_;
    for (uint32_t id : conditioning_proj_residuals_) summary.cond_proj_error += proj_residuals_[id].mahalanobis_distance / proj_residuals_[id].weight;
    summary.unary_error = unary_error_; summary.binary_error = binary_error_;
    summary.trust_region_size = trust_region_size_;
  }

  // ---- read-back -----------------------------------------------------------------
  int get_poses(double* t, double* v, double* b, uint8_t* mu, uint8_t* m) override {
    for (size_t i = 0; i < poses_.size(); ++i) {
      if (t) poses_[i].t_wp.to7(t + 7 * i);
      if (v) for (int k = 0; k < 3; ++k) v[3 * i + k] = poses_[i].v_w[k];
      if (b) for (int k = 0; k < 6; ++k) b[6 * i + k] = poses_[i].b[k];
      if (mu) mu[i] = poses_[i].is_param_mask_used;
      if (m) for (int k = 0; k < D; ++k) m[D * i + k] = (poses_[i].is_param_mask_used && poses_[i].param_mask.size() == (size_t)D) ? poses_[i].param_mask[k] : 1;
    }
    return 0;
  }
  int get_landmarks(double* xw, double* xs, uint8_t* rel, uint32_t* nout) override {
    for (size_t i = 0; i < landmarks_.size(); ++i) {
      if (xw) for (int k = 0; k < 4; ++k) xw[4 * i + k] = landmarks_[i].x_w[k];
      if (xs) for (int k = 0; k < 4; ++k) xs[4 * i + k] = landmarks_[i].x_s[k];
      if (rel) rel[i] = landmarks_[i].is_reliable;
      if (nout) nout[i] = landmarks_[i].num_outlier_residuals;
    }
    return 0;
  }
  int get_proj(double* r, double* w, double* m) override {
    for (size_t i = 0; i < proj_residuals_.size(); ++i) {
      if (r) { r[2 * i] = proj_residuals_[i].residual[0]; r[2 * i + 1] = proj_residuals_[i].residual[1]; }
      if (w) w[i] = proj_residuals_[i].weight;
      if (m) m[i] = proj_residuals_[i].mahalanobis_distance;
    }
    return 0;
  }
  // res.residual / res.mahalanobis_distance / res.cov_inv of the unary, binary and inertial residuals as Solve() leaves them
  // (reference cpp:190-256 EvaluateResiduals, cpp:1469,1526 cov_inv *= weight)
  int get_other(double* ur, double* um, double* uc, double* br, double* bm, double* ir, double* im, double* ic) override {
    for (size_t i = 0; i < unary_residuals_.size(); ++i) {
      const UnaryResidual& r = unary_residuals_[i];
      if (ur) for (int k = 0; k < 6; ++k) ur[6 * i + k] = r.residual[k];
      if (um) um[i] = r.mahalanobis_distance;
      if (uc) for (int a = 0; a < 6; ++a) for (int b = 0; b < 6; ++b) uc[36 * i + 6 * a + b] = r.cov_inv(a, b);
    }
    for (size_t i = 0; i < binary_residuals_.size(); ++i) {
      const BinaryResidual& r = binary_residuals_[i];
      if (br) for (int k = 0; k < 6; ++k) br[6 * i + k] = r.residual[k];
      if (bm) bm[i] = r.mahalanobis_distance;
    }
    for (size_t i = 0; i < inertial_residuals_.size(); ++i) {
      const auto& r = inertial_residuals_[i];
      if (ir) for (int k = 0; k < D; ++k) ir[(size_t)D * i + k] = r.residual[k];
      if (im) im[i] = r.mahalanobis_distance;
      if (ic) for (int a = 0; a < D; ++a) for (int b = 0; b < D; ++b) ic[(size_t)D * D * i + D * a + b] = r.cov_inv(a, b);
    }
    return 0;
  }
  int get_pattern(int which, uint32_t* ncols, uint32_t* nnzb, uint32_t* colptr, uint32_t* rowidx) override {
    std::vector<uint32_t> cp, ri;
    if (which == BAGPU_PATTERN_U) u_.pattern(cp, ri);
    else if (which == BAGPU_PATTERN_W) jt_pr_j_l_.pattern(cp, ri);
    else if (which == BAGPU_PATTERN_S) {
      // union of u_ and the Schur term: the blocks SparseBlockSubtractDenseResult touches
      BlockMat<6, 6> uni(u_.rows(), u_.cols());
      for (int j = 0; j < u_.cols(); ++j) {
        std::vector<int> rows;
        for (const auto& e : u_.col[j]) rows.push_back(e.index);
        if (schur_term_.cols() == u_.cols()) for (const auto& e : schur_term_.col[j]) rows.push_back(e.index);
        std::sort(rows.begin(), rows.end()); rows.erase(std::unique(rows.begin(), rows.end()), rows.end());
        for (int i : rows) uni.insertBack(i, j, Mat<6, 6>());
      }
      uni.pattern(cp, ri);
    } else return BAGPU_ERR_INVALID;
    if (ncols) *ncols = cp.size() - 1;
    if (nnzb) *nnzb = ri.size();
    if (colptr) memcpy(colptr, cp.data(), cp.size() * 4);
    if (rowidx) memcpy(rowidx, ri.data(), ri.size() * 4);
    return 0;
  }
  int get_system(double* s, double* rsc, double* rp, double* rl, double* vi) override {
    const int n = s_.n;
    if (s) { for (int j = 0; j < n; ++j) for (int i = 0; i < n; ++i) s[(size_t)j * n + i] = (i <= j) ? s_.get(i, j) : 0.0; }
    if (rsc) memcpy(rsc, rhs_p_sc_.data(), rhs_p_sc_.size() * 8);
    if (rp) memcpy(rp, rhs_p_.data(), rhs_p_.size() * 8);
    if (rl && LM > 0) memcpy(rl, rhs_l_.data(), rhs_l_.size() * 8);
    if (vi && LM > 0) for (uint32_t l = 0; l < num_active_landmarks_; ++l) { const Mat<LMS, LMS> v = vi_.coeff(l, l); memcpy(vi + (size_t)l * LMS * LMS, v.a, sizeof(double) * LMS * LMS); }
    return 0;
  }

  // write_reduced_camera_matrix block of Solve (reference src/BundleAdjuster.cpp:600-620): s.txt, rhs.txt,
  // j_pr.txt, r_pr.txt, j_l.txt in Eigen's kLongCsvFmt (Utils.h:66: ", " between coefficients, "\n" between
  // rows, no trailing newline).  Digits: %.17g (round-trip) where Eigen's FullPrecision prints 15-16
  // significant digits.  The Jacobians are densified like Eigen::LoadDenseFromSparse does, so this is for
  // parity-sized problems only; j_kpr.txt is not written (CalibSize = 0).
  template <int R, int C>
  static int write_dense_blocks(const std::string& path, const BlockMat<R, C>& m) {
    const size_t rows = (size_t)m.rows() * R, cols = (size_t)m.cols() * C;
    if (rows * cols > ((size_t)1 << 27)) return BAGPU_ERR_INVALID;
    std::vector<double> d(rows * cols, 0.0);
    for (int j = 0; j < m.cols(); ++j)
      for (const auto& e : m.col[j])
        for (int r = 0; r < R; ++r) for (int c = 0; c < C; ++c) d[((size_t)e.index * R + r) * cols + (size_t)j * C + c] = e.value(r, c);
    FILE* f = fopen(path.c_str(), "w");
    if (!f) return BAGPU_ERR_INVALID;
    for (size_t i = 0; i < rows; ++i) {
      if (i) fputc('\n', f);
      for (size_t j = 0; j < cols; ++j) fprintf(f, j ? ", %.17g" : "%.17g", d[i * cols + j]);
    }
    fclose(f);
    return 0;
  }
  static int write_vector(const std::string& path, const VecX& v) {
    FILE* f = fopen(path.c_str(), "w");
    if (!f) return BAGPU_ERR_INVALID;
    for (size_t i = 0; i < v.size(); ++i) fprintf(f, i ? "\n%.17g" : "%.17g", v[i]);
    fclose(f);
    return 0;
  }
  int dump(const char* dir) override {
    const std::string d = dir ? dir : ".";
    const int n = s_.n;
    FILE* f = fopen((d + "/s.txt").c_str(), "w");
    if (!f) return BAGPU_ERR_INVALID;
    for (int i = 0; i < n; ++i) {
      if (i) fputc('\n', f);
      for (int j = 0; j < n; ++j) fprintf(f, j ? ", %.17g" : "%.17g", (i <= j) ? s_.get(i, j) : 0.0);
    }
    fclose(f);
    int rc = write_vector(d + "/rhs.txt", rhs_p_sc_);
    if (rc) return rc;
    if (proj_residuals_.size() == 0) return 0;
    if ((rc = write_dense_blocks(d + "/j_pr.txt", j_pr_))) return rc;
    if ((rc = write_vector(d + "/r_pr.txt", r_pr_))) return rc;
    if (LM > 0 && (rc = write_dense_blocks(d + "/j_l.txt", j_l_))) return rc;
    return 0;
  }
};

static thread_local std::string g_err;

}  // namespace orc

using namespace orc;

struct baoracle_handle_s { OracleBase* impl; bagpu_dims dims; };

extern "C" {

const char* baoracle_last_error(void) { return g_err.c_str(); }

void baoracle_default_options(bagpu_options* o) {
  memset(o, 0, sizeof(*o));
  o->trust_region_size = -1;  // BundleAdjuster.h:75
  o->gyro_sigma = IMU_GYRO_SIGMA; o->accel_sigma = IMU_ACCEL_SIGMA;
  o->gyro_bias_sigma = IMU_GYRO_BIAS_SIGMA; o->accel_bias_sigma = IMU_ACCEL_BIAS_SIGMA;
  o->projection_outlier_threshold = 1.0; o->error_change_threshold = 0.01; o->param_change_threshold = 1e-3;
  o->dogleg_max_inner_iterations = 100; o->apply_results = 1; o->use_dogleg = 1; o->use_triangular_matrices = 1;
  o->use_sparse_solver = 1; o->regularize_biases_in_batch = 1; o->enable_auto_regularization = 1;
  o->use_robust_norm_for_proj_residuals = 1;
  o->pcg_tolerance = 1e-12; o->pcg_max_iterations = 1000;
}

int baoracle_create(const bagpu_dims* d, baoracle_handle* out) {
  OracleBase* impl = nullptr;
  if (d->scalar_bytes != 8 || d->calib_dim != 0 || d->do_tvs != 0) { g_err = "oracle: only double, CalibSize=0, DoTvs=false"; return BAGPU_ERR_UNSUPPORTED; }
#define ORC_CASE(L, D) if (d->lm_dim == L && d->pose_dim == D) impl = new BundleAdjuster<L, D>();
  ORC_CASE(1, 6) ORC_CASE(3, 6) ORC_CASE(0, 6) ORC_CASE(0, 9) ORC_CASE(1, 15) ORC_CASE(0, 15)
#undef ORC_CASE
  if (!impl) { g_err = "oracle: unsupported <LmSize,PoseSize>"; return BAGPU_ERR_UNSUPPORTED; }
  baoracle_default_options(&impl->opt);
  impl->init_after_options();
  *out = new baoracle_handle_s{impl, *d};
  return 0;
}
void baoracle_destroy(baoracle_handle h) { if (h) { delete h->impl; delete h; } }
int baoracle_set_options(baoracle_handle h, const bagpu_options* o) { h->impl->opt = *o; h->impl->init_after_options(); return 0; }
int baoracle_set_cameras(baoracle_handle h, uint32_t n, const double* p, const double* t) { return h->impl->set_cameras(n, p, t); }
int baoracle_set_poses(baoracle_handle h, uint32_t n, const double* t, const double* v, const double* b, const uint8_t* a, const uint8_t* mu, const uint8_t* m) { return h->impl->set_poses(n, t, v, b, a, mu, m); }
int baoracle_set_landmarks(baoracle_handle h, uint32_t n, const double* x, const uint32_t* rp, const uint32_t* rc, const uint8_t* a) { return h->impl->set_landmarks(n, x, rp, rc, a); }
int baoracle_set_proj_residuals(baoracle_handle h, uint32_t n, const double* z, const uint32_t* mp, const uint32_t* lm, const uint32_t* cam, const double* w) { return h->impl->set_proj(n, z, mp, lm, cam, w); }
int baoracle_set_unary_residuals(baoracle_handle h, uint32_t n, const uint32_t* pid, const double* t, const double* ci, const uint8_t* ur) { return h->impl->set_unary(n, pid, t, ci, ur); }
int baoracle_set_binary_residuals(baoracle_handle h, uint32_t n, const uint32_t* p1, const uint32_t* p2, const double* t, const double* ci, const double* w, const uint8_t* ur) { return h->impl->set_binary(n, p1, p2, t, ci, w, ur); }
int baoracle_set_imu_residuals(baoracle_handle h, uint32_t n, const uint32_t* p1, const uint32_t* p2, const uint32_t* ms, const double* m7, const double* w) { return h->impl->set_imu(n, p1, p2, ms, m7, w); }
int baoracle_solve(baoracle_handle h, uint32_t it, double damping, int allow, bagpu_summary* out) {
  for (int i = 0; i < BAGPU_PH_COUNT; ++i) h->impl->phase_s[i] = 0;
  const double t0 = now_s();
  h->impl->Solve(it, damping, allow != 0);
  h->impl->phase_s[BAGPU_PH_TOTAL] = now_s() - t0;
  if (out) *out = h->impl->summary;
  return 0;
}
int baoracle_build(baoracle_handle h, float* ms) {
  double t0 = now_s();
  h->impl->BuildOnly();
  if (ms) *ms = (float)((now_s() - t0) * 1e3);
  return 0;
}
int baoracle_get_poses(baoracle_handle h, double* t, double* v, double* b, uint8_t* mu, uint8_t* m) { return h->impl->get_poses(t, v, b, mu, m); }
int baoracle_get_landmarks(baoracle_handle h, double* xw, double* xs, uint8_t* rel, uint32_t* n) { return h->impl->get_landmarks(xw, xs, rel, n); }
int baoracle_get_proj_residuals(baoracle_handle h, double* r, double* w, double* m) { return h->impl->get_proj(r, w, m); }
int baoracle_get_other_residuals(baoracle_handle h, double* ur, double* um, double* uc, double* br, double* bm, double* ir, double* im, double* ic) {
  return h->impl->get_other(ur, um, uc, br, bm, ir, im, ic);
}
int baoracle_get_trace(baoracle_handle h, uint32_t maxr, bagpu_iter_trace* out, uint32_t* n) {
  uint32_t c = std::min<uint32_t>(maxr, h->impl->trace.size());
  if (out) memcpy(out, h->impl->trace.data(), c * sizeof(bagpu_iter_trace));
  if (n) *n = h->impl->trace.size();
  return 0;
}
int baoracle_get_deltas(baoracle_handle h, uint32_t iter, double* dp, double* dl) {
  if (iter >= h->impl->rec_delta_p.size()) return BAGPU_ERR_INVALID;
  if (dp) memcpy(dp, h->impl->rec_delta_p[iter].data(), h->impl->rec_delta_p[iter].size() * 8);
  if (dl) memcpy(dl, h->impl->rec_delta_l[iter].data(), h->impl->rec_delta_l[iter].size() * 8);
  return 0;
}
int baoracle_get_pattern(baoracle_handle h, int which, uint32_t* nc, uint32_t* nnzb, uint32_t* cp, uint32_t* ri) { return h->impl->get_pattern(which, nc, nnzb, cp, ri); }
int baoracle_get_system(baoracle_handle h, double* s, double* rsc, double* rp, double* rl, double* vi) { return h->impl->get_system(s, rsc, rp, rl, vi); }
int baoracle_dump(baoracle_handle h, const char* dir) { return h->impl->dump(dir); }
int baoracle_sync_state(baoracle_handle) { return 0; }
int baoracle_get_phase_ms(baoracle_handle h, float* ms, uint64_t* launches) {
  for (int i = 0; i < BAGPU_PH_COUNT; ++i) ms[i] = (float)(h->impl->phase_s[i] * 1e3);
  if (launches) *launches = 0;
  return 0;
}

// Standalone helpers so the tests can run the reference's finite-difference checks
// (reference include/ba/BundleAdjusterTest.h, Utils.h:189-219) against this restatement.
void baoracle_math_log_decoupled(const double* a7, const double* b7, double* out6) { Vec6 r = log_decoupled(SE3::from7(a7), SE3::from7(b7)); memcpy(out6, r.a, 48); }
void baoracle_math_exp_decoupled(const double* a7, const double* x6, double* out7) { Vec6 x; memcpy(x.a, x6, 48); exp_decoupled(SE3::from7(a7), x).to7(out7); }
void baoracle_math_compose(const double* a7, const double* b7, double* out7) { (SE3::from7(a7) * SE3::from7(b7)).to7(out7); }
void baoracle_math_inverse(const double* a7, double* out7) { SE3::from7(a7).inverse().to7(out7); }
void baoracle_math_so3_log(const double* q4, double* out3) { SO3 s = SO3::raw(Quat(q4[3], q4[0], q4[1], q4[2])); Vec3 r = s.log(); memcpy(out3, r.a, 24); }
void baoracle_math_so3_exp(const double* w3, double* q4) { Vec3 w; memcpy(w.a, w3, 24); SO3 s = SO3::exp(w); q4[0] = s.q.x; q4[1] = s.q.y; q4[2] = s.q.z; q4[3] = s.q.w; }
void baoracle_math_dlog_dq(const double* q4, double* out12) { Mat<3, 4> m = dlog_dq(Quat(q4[3], q4[0], q4[1], q4[2])); memcpy(out12, m.a, 96); }
void baoracle_math_dlog_decoupled_dx(const double* a7, const double* b7, double* out36) { Mat<6, 6> m = dlog_decoupled_dx(SE3::from7(a7), SE3::from7(b7)); memcpy(out36, m.a, 288); }
void baoracle_math_sqrtm6(const double* a36, double* out36) { Mat<6, 6> m; memcpy(m.a, a36, 288); Mat<6, 6> r = sqrtm_spd(m); memcpy(out36, r.a, 288); }
void baoracle_math_inverse6(const double* a36, double* out36) { Mat<6, 6> m; memcpy(m.a, a36, 288); Mat<6, 6> r = inverse(m); memcpy(out36, r.a, 288); }
// Block-sparse operations on caller-supplied dense matrices, the checks of the reference's math demo
// (applications/math_test/main.cpp:157-314: 70x50 blocks of 6x3 against 50x70 blocks of 3x6, dense and ~5 % sparse).
// a: (nr*6) x (nc*3) row-major, b: (nc*3) x (nr*6) row-major, x: nc*3, y: nr*6.  A block is stored iff it has a non-zero
// coefficient (LoadSparseFromDense, SparseBlockMatrixOps.h:400-414).  Outputs, all dense row-major:
// ab (nr*6)^2 = A*B; ab_upper = A*B with upper_triangular (blocks above the break of Ops.h:236-238 only);
// at (nc*3) x (nr*6) = forceTranspose(A); ax = A*x; aty = A^T*y through the transposed copy; sum = A*B + 2*(A*B);
// a_scaled (nr*6) x (nc*3) = A * blockdiag(d) with d = the nc diagonal 3x3 blocks given in dblocks (nc*9).
}  // extern "C"
template <int R, int C>
static void load_sparse_from_dense(const double* d, int nrb, int ncb, BlockMat<R, C>& m) {
  m.resize(nrb, ncb);
  const size_t ld = (size_t)ncb * C;
  for (int j = 0; j < ncb; ++j) for (int i = 0; i < nrb; ++i) {
    Mat<R, C> blk; bool nz = false;
    for (int r = 0; r < R; ++r) for (int c = 0; c < C; ++c) { blk(r, c) = d[((size_t)i * R + r) * ld + (size_t)j * C + c]; nz |= blk(r, c) != 0.0; }
    if (nz) m.insertBack(i, j, blk);
  }
}
template <int R, int C>
static void load_dense_from_sparse(const BlockMat<R, C>& m, double* d) {
  const size_t ld = (size_t)m.cols() * C;
  std::fill(d, d + (size_t)m.rows() * R * ld, 0.0);
  for (int j = 0; j < m.cols(); ++j) for (const auto& e : m.col[j])
    for (int r = 0; r < R; ++r) for (int c = 0; c < C; ++c) d[((size_t)e.index * R + r) * ld + (size_t)j * C + c] = e.value(r, c);
}
extern "C" {
void baoracle_math_block_sparse_ops(int nr, int nc, const double* a, const double* b, const double* x, const double* y, const double* dblocks,
                                    double* ab, double* ab_upper, double* at, double* ax, double* aty, double* sum, double* a_scaled,
                                    uint32_t* nnzb_out /* A, B, A*B, upper A*B */) {
  BlockMat<6, 3> A; BlockMat<3, 6> B, At; BlockMat<6, 6> AB, ABu, S;
  load_sparse_from_dense(a, nr, nc, A);
  load_sparse_from_dense(b, nc, nr, B);
  SparseBlockProduct(A, B, AB);
  SparseBlockProduct(A, B, ABu, true);
  BlockMat<3, 6>::forceTranspose(A, At);
  load_dense_from_sparse(AB, ab); load_dense_from_sparse(ABu, ab_upper); load_dense_from_sparse(At, at);
  VecX vx(x, x + (size_t)nc * 3), vy(y, y + (size_t)nr * 6), rax((size_t)nr * 6), raty((size_t)nc * 3);
  SparseBlockVectorProductDenseResult(A, vx, rax);
  SparseBlockVectorProductDenseResult(At, vy, raty);
  memcpy(ax, rax.data(), rax.size() * 8); memcpy(aty, raty.data(), raty.size() * 8);
  SparseBlockAdd(AB, AB, S, 2);
  load_dense_from_sparse(S, sum);
  BlockMat<3, 3> D(nc, nc);
  for (int j = 0; j < nc; ++j) { Mat<3, 3> blk; memcpy(blk.a, dblocks + (size_t)j * 9, 72); D.insertBack(j, j, blk); }
  BlockMat<6, 3> AD;
  SparseBlockDiagonalRhsProduct(A, D, AD);
  load_dense_from_sparse(AD, a_scaled);
  if (nnzb_out) { nnzb_out[0] = (uint32_t)A.nonZeros(); nnzb_out[1] = (uint32_t)B.nonZeros(); nnzb_out[2] = (uint32_t)AB.nonZeros(); nnzb_out[3] = (uint32_t)ABu.nonZeros(); }
}
// binary residual + Jacobians exactly as BuildProblem computes them (cpp:1396-1411)
void baoracle_math_binary(const double* tw1_7, const double* tw2_7, const double* t12_7, double* r6, double* j1_36, double* j2_36) {
  const SE3 t_w1 = SE3::from7(tw1_7), t_w2 = SE3::from7(tw2_7), t12m = SE3::from7(t12_7);
  const SE3 t_1w = t_w1.inverse(); const SE3 t_12 = t_1w * t_w2;
  Vec6 r = log_decoupled(t_12, t12m);
  const Mat<6, 7> dlog_dt1 = dLog_decoupled_dt1(t_12, t12m);
  Mat<6, 6> j1 = dlog_dt1 * dt1_t2_dt1(t_1w, t_w2) * dinv_exp_decoupled_dx(t_w1);
  Mat<6, 6> j2 = dlog_dt1 * dt1_t2_dt2(t_1w) * dexp_decoupled_dx(t_w2);
  memcpy(r6, r.a, 48); memcpy(j1_36, j1.a, 288); memcpy(j2_36, j2.a, 288);
}
// IMU integration: state y = [t(3) q(4) v(3)], with d y / d bias (10x6) and the covariance
void baoracle_math_imu_integrate(const double* t_wp7, const double* v3, const double* bg3, const double* ba3, const double* g3,
                                 uint32_t n_meas, const double* meas7, const double* r6diag, double* y10, double* dy_db60, double* c100) {
  std::vector<ImuMeasurement> ms(n_meas);
  for (uint32_t k = 0; k < n_meas; ++k) { for (int c = 0; c < 3; ++c) { ms[k].w[c] = meas7[7 * k + c]; ms[k].a[c] = meas7[7 * k + 3 + c]; } ms[k].time = meas7[7 * k + 6]; }
  Vec3 v, bg, ba, g; memcpy(v.a, v3, 24); memcpy(bg.a, bg3, 24); memcpy(ba.a, ba3, 24); memcpy(g.a, g3, 24);
  Diag6 r; for (int i = 0; i < 6; ++i) r(i, i) = r6diag ? r6diag[i] : 0.0;
  std::vector<ImuPose> poses; Mat<10, 6> db; Mat<10, 10> c;
  ImuPose out = (dy_db60 || c100) ? ImuResidual<15>::IntegrateResidual(ImuPose(SE3::from7(t_wp7), v, Vec3(), 0), ms, bg, ba, g, poses, &db, nullptr, &c, &r)
                                  : ImuResidual<15>::IntegrateResidual(ImuPose(SE3::from7(t_wp7), v, Vec3(), 0), ms, bg, ba, g, poses);
  for (int i = 0; i < 3; ++i) { y10[i] = out.t_wp.t[i]; y10[7 + i] = out.v_w[i]; }
  y10[3] = out.t_wp.r.q.x; y10[4] = out.t_wp.r.q.y; y10[5] = out.t_wp.r.q.z; y10[6] = out.t_wp.r.q.w;
  if (dy_db60) memcpy(dy_db60, db.a, 480);
  if (c100) memcpy(c100, c.a, 800);
}

}  // extern "C"
